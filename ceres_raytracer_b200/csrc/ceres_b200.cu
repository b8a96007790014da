// ceres_b200.cu -- B200 (sm_100a) implementation of the ceres-raytracer render loop behind the C ABI of
// include/ceres_b200.h.  Written from scratch; citations are to the reference behaviour each piece reproduces
// (paths relative to the upstream repository).
//
// Pipeline per frame: ONE persistent kernel (frame_fused.cuh) whose warps first trace 8x4-pixel tiles of the scene's
// conservative screen rectangle -- camera ray generation (render.hpp:105-113), ordered eager-leaf closest-hit BVH
// traversal (single_ray_traverser.hpp:65-126), Moeller-Trumbore (triangle.hpp:95-115); misses store their zeros, hits
// are ballot-compacted into a 16-byte record queue -- and then, 32 queue records at a time, rebuild the shadow rays
// (render.hpp:120-136), run the any-hit traversal (only bool(hit) is used, render.hpp:137-150) and shade + store the
// pixels (render.hpp:56-84) together with the frame counters (render.hpp:115,119,138,148,155).  Pixels outside the
// rectangle are zeroed by clear_outside_kernel where the output buffer is not known to hold zeros there already.
//
// Arithmetic: every floating-point operation of the path is issued through the round-to-nearest intrinsics
// (__fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn/__fsqrt_rn never contract), the slab test uses __fmaf_rn exactly where
// the reference calls fast_multiply_add (utilities.hpp:37-40), so results are those of the reference built with
// -ffp-contract=off regardless of nvcc's -fmad setting.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cfloat>
#include <string>
#include <vector>
#include <atomic>
#include <mutex>
#include <memory>
#include <immintrin.h>
#include <chrono>
#include <new>
#include <algorithm>
#include <thread>

#include "../../include/ceres_b200.h"
#include "host_fill.hpp"
#include "checked.cuh"

// ------------------------------------------------------------------------------------------------ errors
static thread_local std::string g_err;
static std::atomic<uint64_t> g_launches{0};

static int fail(int code, const std::string& msg) { g_err = msg; return code; }
static thread_local int tl_sequence_depth = 0;       // this thread is inside ceres_render_sequence (its own frames are not "concurrent")
// A scene handle is not thread-safe; the one place where the LIBRARY runs a second thread over it is ceres_render_sequence, so
// that is where a concurrent call from the caller's side would race silently (geometry moved by a refit under the second
// thread's frames, per-frame state shared by two renders).  Such calls fail loudly instead.
#define CERES_REJECT_DURING_SEQUENCE(s, who)                                                                 \
    do {                                                                                                     \
        if ((s) && (s)->in_sequence.load(std::memory_order_acquire) && tl_sequence_depth == 0)               \
            return fail(CERES_ERR_INVALID, std::string(who) + ": ceres_render_sequence is running on this scene in another thread"); \
    } while (0)
#define CUDA_TRY(expr)                                                                                      \
    do {                                                                                                    \
        cudaError_t e_ = (expr);                                                                            \
        if (e_ != cudaSuccess)                                                                              \
            return fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? CERES_ERR_NO_DEVICE  \
                                                                                     : CERES_ERR_CUDA,      \
                        std::string(#expr) + ": " + cudaGetErrorString(e_));                                \
    } while (0)

// ------------------------------------------------------------------------------------------------ device data
// Node pair = the two sibling nodes of the reference (bvh.hpp:13-17,25-30) in one 64-byte, 64-byte-aligned record:
//   q[0] = L.minx L.maxx L.miny L.maxy   q[1] = L.minz L.maxz L.count L.first
//   q[2] = R.minx R.maxx R.miny R.maxy   q[3] = R.minz R.maxz R.count R.first
// `first` of an inner node is the index of its child PAIR; of a leaf the first slot in the leaf-ordered triangle
// array (= index into Bvh::primitive_indices, so the per-test indirection of primitive_intersectors.hpp:17-20
// disappears).  Triangles are the reference's 48-byte {p0,e1,e2,n} records (triangle.hpp:27) permuted into leaf
// order and read as 3 x float4.
struct DevScene {
    const float4* pairs;
    const float4* tris;     // 3 float4 per slot
    const float4* norms;    // 3 float4 per slot: n0.xyz n1.x | n1.yz n2.xy | n2.z 0 0 0
    const uint32_t* slot_id;  // slot -> primitive index of the reference (bvh.primitive_indices[slot])
    uint32_t root_leaf_count;  // != 0: the root itself is a leaf (single_ray_traverser.hpp:72-73)
    uint32_t stack_depth;      // entries of traversal stack needed (<= 64, single_ray_traverser.hpp:14)
    unsigned long long* hist;  // statistics mode: 2 x 64 bins of node-pair steps per ray (bin = steps / 4), primary then shadow; may be null
    uint32_t n_pairs, n_slots;     // bounds of pairs[] / tris[] (checked build)
    unsigned long long* check;     // checked build: CHECK_* bits are ORed in here (the scene's error word); null otherwise
    unsigned long long* timeline;  // CERES_TIMELINE_WORDS words per warp of the frame kernel (ceres_scene_set_timeline); null = off
};

struct FrameCounters {
    uint32_t qcount;        // primary hits appended to the queue
    uint32_t shadow_next;   // work-stealing cursor of the shadow kernel
    uint32_t occluded;      // shadow rays that hit something
    uint32_t done_blocks;   // shade kernel: blocks finished (last one publishes the counters)
    uint32_t primary_next;  // work-stealing cursor of the persistent primary wave
    uint32_t ext[4];        // min column, max column, min local row, max local row over the band's hit pixels
    uint32_t primary_done;  // frame kernel: primary tiles retired (the warp that retires the last one closes the queue)
    uint32_t error;         // frame kernel: FRAME_ERR_* bits (a bounded wait expired)
    uint32_t pad_[1];
    unsigned long long stats[6];  // S_miss,T_miss,S_hit,T_hit,S_shadow,T_shadow (statistics mode only)
};

// Control block of a peer frame (lives in rank 0's device memory, mapped by every rank; include/ceres_b200.h).
struct PeerCtl {
    uint32_t cleared_epoch;                          // rank 0: my kernels of this epoch have begun (the other ranks may store)
    uint32_t clear_blocks;                           // rank 0: CTAs of the clear pass that have finished
    uint32_t error;                                  // a bounded wait expired (1 = cleared wait, 2 = done wait)
    uint32_t pad_;
    uint32_t done_epoch[CERES_PEER_MAX_RANKS];       // rank r: all my pixels of this epoch are visible system-wide
    unsigned long long counters[CERES_PEER_MAX_RANKS][2];   // rank r: (rays, hits) of my shard
};

struct Frame {
    float eye[3], dir[3], iu[3], iv[3], sun[3];
    uint32_t W, H;             // full frame
    uint32_t rows_local;       // rows of this shard
    uint32_t row_begin;        // first local row of this launch (a frame may be rendered in several row bands)
    uint32_t row_end;          // one past the last local row of this launch
    uint32_t accumulate;       // add this launch's counters to out_counters instead of overwriting them
    uint32_t block_rows, shard_index, shard_count;
    uint32_t compact;          // outputs packed by local row
    uint32_t no_flip;          // quantised payloads: output row = frame row (anim.cpp:117-119) instead of height-1-row (static.cpp:135-147)
    uint32_t shadows;
    // persistent primary wave: conservative screen rectangle of the scene, in local rows / columns of this launch,
    // cut into 8x4-pixel tiles (work item = tile * 32 + pixel in tile)
    uint32_t rect_x0, rect_x1, rect_row0, rect_row1, tiles_x, tiles_y, work_items;
    uint32_t refill_min;                            // refill a warp when at least this many ray slots are idle (1..32)
    uint32_t chunk;                                 // work items a warp takes from the global cursor at a time (multiple of 32)
    uint32_t cull_i0, cull_i1, cull_l0, cull_l1;   // host only: the rectangle for the whole frame (columns, LOCAL rows)
    float* pixels;
    uint32_t* ids;
    unsigned long long* out_counters;   // [rays, hits]
    unsigned long long* out_stats;      // 8 x u64 or null
    uint4* queue;
    FrameCounters* ctr;
    // peer-frame mode (ceres_render_peer): pixels/ids point into rank 0's memory
    PeerCtl* peer;
    uint32_t peer_epoch, peer_rank;
    uint32_t skip_clear;       // experiment kernels: the rows were zeroed by somebody else (rank 0 in peer mode)
    uint32_t peer_wait_world;  // frame kernel, rank 0 of a peer frame: wait for this many ranks' done flags at the end (0 = no)
    unsigned long long* peer_out;   // ... and put the frame's summed (rays, hits) here (nullable)
    uint32_t* gate;            // peer frames, ranks > 0: LOCAL relay of rank 0's cleared_epoch {flag, poller election} (frame_fused.cuh)
    uint32_t* ready;           // frame kernel: per-32-record-chunk fill counters of this launch's queue region (frame_fused.cuh)
    unsigned long long* out_error;  // frame kernel: FRAME_ERR_* bits are ORed in here (nullable)
    // host-frame path: where the band's last shade CTA publishes the extents of its hit pixels (mapped page-locked host
    // memory, 8 words per band: ext[4], then the frame's epoch as the ready flag); null = not wanted
    uint32_t* host_ext;
    uint32_t ext_epoch;
    // frame kernel, shadow-subtree sharing (frame_fused.cuh): steps between two rounds of bookkeeping; idle lanes needed for a hand-over round
    uint32_t share_period, share_min_idle;
    uint32_t chunk_shift;      // frame kernel: a shadow work item is 1 << chunk_shift queue records (3..5)
    uint32_t clear_rows;       // frame kernel: zero the pixels of rows [row_begin, row_end) outside the rectangle itself
    uint32_t queue_cap, ready_cap, pix_rows;   // checked build: records / ready[] words of this launch's regions, rows of the output buffer
};

// ------------------------------------------------------------------------------------------------ device math
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float rcp(float a) { return __fdiv_rn(1.0f, a); }
// vector.hpp:134-141: sum = a0*b0; sum += a1*b1; sum += a2*b2
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return add(add(mul(ax, bx), mul(ay, by)), mul(az, bz));
}
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 vadd(V3 a, V3 b) { return mk(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return mk(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
__device__ __forceinline__ V3 vscale(V3 a, float s) { return mk(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
__device__ __forceinline__ float vdot(V3 a, V3 b) { return dot3(a.x, a.y, a.z, b.x, b.y, b.z); }
// vector.hpp:159-167
__device__ __forceinline__ V3 vcross(V3 a, V3 b) {
    return mk(sub(mul(a.y, b.z), mul(a.z, b.y)), sub(mul(a.z, b.x), mul(a.x, b.z)), sub(mul(a.x, b.y), mul(a.y, b.x)));
}
// vector.hpp:143-154: inv = 1/sqrt(dot(v,v)); v*inv
__device__ __forceinline__ V3 vnormalize(V3 a) {
    float inv = rcp(__fsqrt_rn(vdot(a, a)));
    return vscale(a, inv);
}

struct Tri { V3 p0, e1, e2, n; };
__device__ __forceinline__ Tri load_tri(const float4* __restrict__ tris, uint32_t slot) {
    const float4* p = tris + 3ull * slot;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    Tri t;
    t.p0 = mk(a.x, a.y, a.z); t.e1 = mk(a.w, b.x, b.y); t.e2 = mk(b.z, b.w, c.x); t.n = mk(c.y, c.z, c.w);
    return t;
}

struct Hit { uint32_t slot; float t, u, v; };

// triangle.hpp:95-115.  Returns true and updates (t,u,v) when the ray hits within [tmin, tmax].
__device__ __forceinline__ bool tri_test(const Tri& tr, V3 o, V3 d, float tmin, float tmax, float& t, float& u, float& v) {
    V3 c = vsub(tr.p0, o);
    V3 r = vcross(d, c);
    float inv_det = rcp(vdot(tr.n, d));
    float uu = mul(vdot(r, tr.e2), inv_det);
    float vv = mul(vdot(r, tr.e1), inv_det);
    float ww = sub(sub(1.0f, uu), vv);
    if (uu >= 0.0f && vv >= 0.0f && ww >= 0.0f) {
        float tt = mul(vdot(tr.n, c), inv_det);
        if (tt >= tmin && tt <= tmax) { t = tt; u = uu; v = vv; return true; }
    }
    return false;
}

// Ordered, eager-leaf, stack-based traversal (single_ray_traverser.hpp:65-126) over node pairs.
//   * both children are slab-tested with the tmax seen at the start of the step (:86-87);
//   * a hit leaf is intersected immediately, left before right, slots ascending (:89-107, :51-62), `t <= tmax`
//     lets a later equal-t triangle replace an earlier one (triangle.hpp:110);
//   * both inner children alive: the nearer entry goes first, ties keep the left one, the other's child pair is
//     pushed (:109-115).
// Slab test = node_intersectors.hpp:34-47,96-100: entry_k = fma(bounds[2k+oct_k], inv_k, so_k),
// exit_k = fma(bounds[2k+1-oct_k], inv_k, so_k); the robust_max/min chains (utilities.hpp:57-67) are evaluated with
// fmaxf/fminf, which return the same value whenever the right-hand accumulator is not NaN (tmin/tmax never are);
// only the sign of a zero result can differ, and entry/exit are only ever compared.
// ANY = true: return at the first accepted triangle (used for shadow rays, where only the boolean is consumed; the
// boolean does not depend on traversal order because tmax never shrinks before the first hit).
template <bool ANY, bool STATS>
__device__ __forceinline__ bool traverse(const DevScene& sc, V3 o, V3 d, float tmin, float tmax,
                                         uint32_t* __restrict__ stack, uint32_t stride, Hit& best,
                                         uint32_t& n_steps, uint32_t& n_tests) {
    bool found = false;
    const float4* __restrict__ tris = sc.tris;
    if (sc.root_leaf_count) {  // single_ray_traverser.hpp:72-73
        if (STATS) n_tests += sc.root_leaf_count;
        for (uint32_t s = 0; s < sc.root_leaf_count; ++s) {
            Tri tr = load_tri(tris, s);
            float t, u, v;
            if (tri_test(tr, o, d, tmin, tmax, t, u, v)) {
                best.slot = s; best.t = t; best.u = u; best.v = v; found = true; tmax = t;
                if (ANY) return true;
            }
        }
        return found;
    }
    // FastNodeIntersector (node_intersectors.hpp:20-26,89-94; vector.hpp:69-74)
    const bool ox = signbit(d.x), oy = signbit(d.y), oz = signbit(d.z);
    const float eps = FLT_EPSILON;
    const float ix = rcp(fabsf(d.x) < eps ? copysignf(eps, d.x) : d.x);
    const float iy = rcp(fabsf(d.y) < eps ? copysignf(eps, d.y) : d.y);
    const float iz = rcp(fabsf(d.z) < eps ? copysignf(eps, d.z) : d.z);
    const float sx = mul(-o.x, ix), sy = mul(-o.y, iy), sz = mul(-o.z, iz);

    const float4* __restrict__ pairs = sc.pairs;
    uint32_t cur = 0;  // child pair of the root
    uint32_t sp = 0;
    for (;;) {
        if (STATS) n_steps++;
        CERES_CLAMP(cur, sc.n_pairs, CHECK_PAIR_INDEX, sc);
        const float4* pp = pairs + 4ull * cur;
        const float4 q0 = __ldg(pp), q1 = __ldg(pp + 1), q2 = __ldg(pp + 2), q3 = __ldg(pp + 3);
        float le, lx, re, rx;
        {
            float e0 = __fmaf_rn(ox ? q0.y : q0.x, ix, sx), x0 = __fmaf_rn(ox ? q0.x : q0.y, ix, sx);
            float e1 = __fmaf_rn(oy ? q0.w : q0.z, iy, sy), x1 = __fmaf_rn(oy ? q0.z : q0.w, iy, sy);
            float e2 = __fmaf_rn(oz ? q1.y : q1.x, iz, sz), x2 = __fmaf_rn(oz ? q1.x : q1.y, iz, sz);
            le = fmaxf(e0, fmaxf(e1, fmaxf(e2, tmin)));
            lx = fminf(x0, fminf(x1, fminf(x2, tmax)));
        }
        {
            float e0 = __fmaf_rn(ox ? q2.y : q2.x, ix, sx), x0 = __fmaf_rn(ox ? q2.x : q2.y, ix, sx);
            float e1 = __fmaf_rn(oy ? q2.w : q2.z, iy, sy), x1 = __fmaf_rn(oy ? q2.z : q2.w, iy, sy);
            float e2 = __fmaf_rn(oz ? q3.y : q3.x, iz, sz), x2 = __fmaf_rn(oz ? q3.x : q3.y, iz, sz);
            re = fmaxf(e0, fmaxf(e1, fmaxf(e2, tmin)));
            rx = fminf(x0, fminf(x1, fminf(x2, tmax)));
        }
        const uint32_t cL = __float_as_uint(q1.z), fL = __float_as_uint(q1.w);
        const uint32_t cR = __float_as_uint(q3.z), fR = __float_as_uint(q3.w);
        const bool hitL = le <= lx, hitR = re <= rx;

        // leaves: left range then right range
        uint32_t s = 0, e = 0, s2 = 0, e2 = 0;
        if (hitL && cL) { s = fL; e = fL + cL; }
        if (hitR && cR) { s2 = fR; e2 = fR + cR; }
        if (s == e) { s = s2; e = e2; s2 = e2; }
        while (s < e) {
            if (STATS) n_tests++;
            CERES_CK(s < sc.n_slots, CHECK_SLOT_INDEX, sc);
            Tri tr = load_tri(tris, s < sc.n_slots ? s : 0u);
            float t, u, v;
            if (tri_test(tr, o, d, tmin, tmax, t, u, v)) {
                best.slot = s; best.t = t; best.u = u; best.v = v; found = true; tmax = t;
                if (ANY) return true;
            }
            ++s;
            if (s == e) { s = s2; e = e2; s2 = e2; }
        }

        const bool goL = hitL && !cL, goR = hitR && !cR;
        if (goL && goR) {
            const bool swap = le > re;  // ties keep left first
            const uint32_t nearc = swap ? fR : fL, farc = swap ? fL : fR;
            CERES_CK(sp < sc.stack_depth, CHECK_STACK_DEPTH, sc);
            stack[sp * stride] = farc;
            ++sp;
            cur = nearc;
        } else if (goL || goR) {
            cur = goL ? fL : fR;
        } else {
            if (sp == 0) break;
            --sp;
            cur = stack[sp * stride];
        }
    }
    return found;
}

// ------------------------------------------------------------------------------------------------ ray setup
// local row -> frame row for a row-block shard
__device__ __forceinline__ uint32_t frame_row(const Frame& f, uint32_t lrow) {
    if (f.shard_count <= 1) return lrow;
    uint32_t blk = lrow / f.block_rows;
    return (blk * f.shard_count + f.shard_index) * f.block_rows + (lrow - blk * f.block_rows);
}

// render.hpp:109-111
__device__ __forceinline__ V3 primary_dir(const Frame& f, uint32_t i, uint32_t j) {
    float u = sub(__fdiv_rn(mul(2.0f, add((float)i, 0.5f)), (float)f.W), 1.0f);
    float v = sub(__fdiv_rn(mul(2.0f, add((float)j, 0.5f)), (float)f.H), 1.0f);
    V3 iu = mk(f.iu[0], f.iu[1], f.iu[2]), iv = mk(f.iv[0], f.iv[1], f.iv[2]), dir = mk(f.dir[0], f.dir[1], f.dir[2]);
    return vnormalize(vadd(vadd(vscale(iu, u), vscale(iv, v)), dir));
}

// render.hpp:120-135: hit point with the reference's barycentric convention (Q1), offset along -normal, sun line
__device__ __forceinline__ void shadow_setup(const Tri& tr, float u, float v, V3 sun, V3& P, V3& sun_line) {
    V3 normal = vnormalize(tr.n);
    V3 p1 = vsub(tr.p0, tr.e1), p2 = vadd(tr.p0, tr.e2);       // triangle.hpp:36-37
    float w = sub(sub(1.0f, u), v);
    P = vadd(vadd(vscale(tr.p0, u), vscale(p1, v)), vscale(p2, w));
    const float scale = -0.00001f;
    P = vadd(P, vscale(normal, scale));
    sun_line = vnormalize(vsub(sun, P));
}

// ------------------------------------------------------------------------------------------------ kernels
#define PRIMARY_THREADS 256
#define OCCLUDED_BIT 0x80000000u

#ifdef CERES_EXPERIMENTS   // tuning variant 1: the first kernels of round 1 (one thread per pixel, separate waves)
// One thread per pixel; a warp covers an 8x4 pixel tile, a CTA 32x8 pixels.
template <bool STATS>
__global__ void __launch_bounds__(PRIMARY_THREADS) primary_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t i = blockIdx.x * 32u + (warp & 3u) * 8u + (lane & 7u);
    const uint32_t lrow = f.row_begin + blockIdx.y * 8u + (warp >> 2) * 4u + (lane >> 3);
    const bool valid = i < f.W && lrow < f.row_end;
    const uint32_t j = frame_row(f, lrow);

    bool found = false;
    Hit best; best.slot = 0; best.t = best.u = best.v = 0.0f;
    uint32_t n_steps = 0, n_tests = 0;
    if (valid) {
        V3 view = primary_dir(f, i, j);
        V3 eye = mk(f.eye[0], f.eye[1], f.eye[2]);
        found = traverse<false, STATS>(sc, eye, view, 0.0f, FLT_MAX, smem_stack + tid, PRIMARY_THREADS, best, n_steps, n_tests);
    }
    __syncwarp();
    const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
    if (valid && !found) {   // render.hpp:116-117
        float* p = f.pixels + 3 * opix;
        p[0] = 0.0f; p[1] = 0.0f; p[2] = 0.0f;
        if (f.ids) f.ids[opix] = CERES_MISS;
    }
    // ballot compaction of the hits into the record queue (one atomic per warp)
    const unsigned m = __ballot_sync(0xFFFFFFFFu, found);
    if (m) {
        const int leader = __ffs(m) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&f.ctr->qcount, (uint32_t)__popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (found) {
            const uint32_t at = base + __popc(m & ((1u << lane) - 1u));
            // pixel is stored as (local row, column) so that the later passes can rebuild both the frame row and
            // the output position
            f.queue[at] = make_uint4(lrow * f.W + i, best.slot, __float_as_uint(best.u), __float_as_uint(best.v));
        }
    }
    if (STATS) {
        unsigned long long s_m = found ? 0 : n_steps, t_m = found ? 0 : n_tests;
        unsigned long long s_h = found ? n_steps : 0, t_h = found ? n_tests : 0;
        for (int o = 16; o; o >>= 1) {
            s_m += __shfl_xor_sync(0xFFFFFFFFu, s_m, o); t_m += __shfl_xor_sync(0xFFFFFFFFu, t_m, o);
            s_h += __shfl_xor_sync(0xFFFFFFFFu, s_h, o); t_h += __shfl_xor_sync(0xFFFFFFFFu, t_h, o);
        }
        if (lane == 0) {
            atomicAdd(&f.ctr->stats[0], s_m); atomicAdd(&f.ctr->stats[1], t_m);
            atomicAdd(&f.ctr->stats[2], s_h); atomicAdd(&f.ctr->stats[3], t_h);
        }
    }
}

// Persistent warps; each steals 32 consecutive records (= one ballot-compacted group of neighbouring pixels).
#define SHADOW_THREADS 256
template <bool STATS>
__global__ void __launch_bounds__(SHADOW_THREADS) shadow_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t total = f.ctr->qcount;
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);
    uint32_t occl = 0;
    unsigned long long st_s = 0, st_t = 0;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&f.ctr->shadow_next, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= total) break;
        const uint32_t r = base + lane;
        if (r < total) {
            uint4 rec = f.queue[r];
            Tri tr = load_tri(sc.tris, rec.y);
            V3 P, sun_line;
            shadow_setup(tr, __uint_as_float(rec.z), __uint_as_float(rec.w), sun, P, sun_line);
            Hit h; h.slot = 0; h.t = h.u = h.v = 0.0f;
            uint32_t n_steps = 0, n_tests = 0;
            bool hit = traverse<true, STATS>(sc, P, sun_line, 0.0f, FLT_MAX, smem_stack + tid, SHADOW_THREADS, h, n_steps, n_tests);
            if (hit) { f.queue[r].y = rec.y | OCCLUDED_BIT; occl++; }
            if (STATS) { st_s += n_steps; st_t += n_tests; }
        }
        __syncwarp();
    }
    for (int o = 16; o; o >>= 1) occl += __shfl_xor_sync(0xFFFFFFFFu, occl, o);
    if (lane == 0 && occl) atomicAdd(&f.ctr->occluded, occl);
    if (STATS) {
        for (int o = 16; o; o >>= 1) { st_s += __shfl_xor_sync(0xFFFFFFFFu, st_s, o); st_t += __shfl_xor_sync(0xFFFFFFFFu, st_t, o); }
        if (lane == 0) { atomicAdd(&f.ctr->stats[4], st_s); atomicAdd(&f.ctr->stats[5], st_t); }
    }
}

#endif  // CERES_EXPERIMENTS

// render.hpp:46-54,56-84.  pow(double(x), 24) by repeated squaring in double (5 roundings, relative error
// < 3e-16, then narrowed to float like the reference's `return std::pow(...)` in a float function).
__device__ __forceinline__ float spec_pow24(float x) {
    double b = (double)x;
    double b2 = __dmul_rn(b, b), b3 = __dmul_rn(b2, b), b6 = __dmul_rn(b3, b3), b12 = __dmul_rn(b6, b6);
    return (float)__dmul_rn(b12, b12);
}
__device__ __forceinline__ float clamp01(float x) { return x < 0.0f ? 0.0f : (1.0f < x ? 1.0f : x); }

// ---- system-scope flag helpers of the peer-frame protocol ------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define PEER_WAIT_NS 4000000000ull
// Bounded wait until *flag >= epoch (epochs are compared as a signed difference so that they may wrap).
__device__ __forceinline__ bool peer_wait_flag(const uint32_t* flag, uint32_t epoch) {
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
        __nanosleep(200);
        if (global_ns() - t0 > PEER_WAIT_NS) return false;
    }
    return true;
}

#ifdef CERES_EXPERIMENTS   // separate shade pass + rank-0 clear / wait kernels of the wave pipelines (variants 1-6)
#define SHADE_THREADS 256
__global__ void __launch_bounds__(SHADE_THREADS) shade_kernel(const DevScene sc, const Frame f) {
    const uint32_t total = f.ctr->qcount;
    if (f.peer && f.peer_rank != 0u) {
        // my stores land in rank 0's frame: they must not overtake rank 0's clear of this epoch
        if (threadIdx.x == 0 && !peer_wait_flag(&f.peer->cleared_epoch, f.peer_epoch)) f.peer->error = 1u;
        __syncthreads();
    }
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);
    uint32_t e_i0 = 0xFFFFFFFFu, e_i1 = 0u, e_r0 = 0xFFFFFFFFu, e_r1 = 0u;
    for (uint32_t r = blockIdx.x * SHADE_THREADS + threadIdx.x; r < total; r += gridDim.x * SHADE_THREADS) {
        const uint4 rec = f.queue[r];
        const uint32_t lrow = rec.x / f.W, i = rec.x - lrow * f.W;
        e_i0 = min(e_i0, i); e_i1 = max(e_i1, i); e_r0 = min(e_r0, lrow); e_r1 = max(e_r1, lrow);
        const uint32_t j = frame_row(f, lrow);
        const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
        const uint32_t slot = rec.y & ~OCCLUDED_BIT;
        float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
        if (!(rec.y & OCCLUDED_BIT)) {
            const float u = __uint_as_float(rec.z), v = __uint_as_float(rec.w);
            Tri tr = load_tri(sc.tris, slot);
            V3 P, sun_line;
            shadow_setup(tr, u, v, sun, P, sun_line);
            V3 view = primary_dir(f, i, j);
            V3 mview = vscale(view, -1.0f);                      // render.hpp:66
            V3 h = vnormalize(vadd(sun_line, mview));            // render.hpp:53
            const float4* np = sc.norms + 3ull * slot;
            float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2);
            V3 nk[3] = {mk(a.x, a.y, a.z), mk(a.w, b.x, b.y), mk(b.z, b.w, c.x)};
            float wk[3] = {u, v, sub(sub(1.0f, u), v)};
            const float amb = 0.2f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                // lambertian (render.hpp:46-49): abs of the left-to-right sum
                float diffuse = mul(0.5f, fabsf(add(add(mul(sun_line.x, nk[k].x), mul(sun_line.y, nk[k].y)), mul(sun_line.z, nk[k].z))));
                float specular = mul(0.8f, spec_pow24(vdot(nk[k], h)));
                float ad = add(amb, diffuse);
                c0 = add(c0, mul(wk[k], clamp01(add(mul(ad, 0.5f), specular))));
                c1 = add(c1, mul(wk[k], clamp01(add(mul(ad, 0.0f), specular))));
                c2 = add(c2, mul(wk[k], clamp01(add(mul(ad, 0.8f), specular))));
            }
        }
        float* p = f.pixels + 3 * opix;
        p[0] = c0; p[1] = c1; p[2] = c2;
        if (f.ids) f.ids[opix] = __ldg(sc.slot_id + slot);
    }
    if (f.host_ext) {        // uniform branch: extents of the hit pixels this warp wrote
        e_i0 = __reduce_min_sync(0xFFFFFFFFu, e_i0); e_i1 = __reduce_max_sync(0xFFFFFFFFu, e_i1);
        e_r0 = __reduce_min_sync(0xFFFFFFFFu, e_r0); e_r1 = __reduce_max_sync(0xFFFFFFFFu, e_r1);
        if ((threadIdx.x & 31u) == 0u && e_i0 <= e_i1) {
            atomicMin(&f.ctr->ext[0], e_i0); atomicMax(&f.ctr->ext[1], e_i1); atomicMin(&f.ctr->ext[2], e_r0); atomicMax(&f.ctr->ext[3], e_r1);
        }
    }
    // the last CTA publishes the frame counters and re-arms the per-frame state for the next frame
    __shared__ bool last;
    if (f.peer) __threadfence_system();     // my pixel stores (NVLink) are performed before the CTA is counted done
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last = atomicAdd(&f.ctr->done_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        const unsigned long long hits = total, occ = f.ctr->occluded;
        const unsigned long long prim = (unsigned long long)f.W * (f.row_end - f.row_begin);
        const unsigned long long rays = f.shadows ? prim + hits : prim;                          // render.hpp:115,138
        const unsigned long long nh = hits + occ;                                                 // render.hpp:119,148
        // several launches (row bands, possibly on concurrent streams) add into the same frame counters; the host
        // zeroes them before the first launch of a frame
        if (f.out_counters) { atomicAdd(&f.out_counters[0], rays); atomicAdd(&f.out_counters[1], nh); }
        if (f.out_stats) {
            for (int k = 0; k < 6; ++k) atomicAdd(&f.out_stats[k], f.ctr->stats[k]);
            atomicAdd(&f.out_stats[6], rays); atomicAdd(&f.out_stats[7], nh);
        }
        f.ctr->qcount = 0; f.ctr->shadow_next = 0; f.ctr->occluded = 0; f.ctr->done_blocks = 0; f.ctr->primary_next = 0;
        for (int k = 0; k < 6; ++k) f.ctr->stats[k] = 0;
        if (f.host_ext) {
            // every CTA fenced its pixel stores and its extent atomics before it was counted: tell the host which
            // rectangle of this band holds anything but zeros, then raise the flag it is polling
            volatile uint32_t* e = f.ctr->ext;
            for (int k = 0; k < 4; ++k) f.host_ext[k] = e[k];
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(f.host_ext + 4), "r"(f.ext_epoch) : "memory");
        }
        f.ctr->ext[0] = 0xFFFFFFFFu; f.ctr->ext[1] = 0u; f.ctr->ext[2] = 0xFFFFFFFFu; f.ctr->ext[3] = 0u;
        __threadfence();
        if (f.peer) {
            // every CTA fenced its stores at system scope before it was counted; publish counters, then the flag
            f.peer->counters[f.peer_rank][0] = rays;
            f.peer->counters[f.peer_rank][1] = nh;
            __threadfence_system();
            st_release_sys(&f.peer->done_epoch[f.peer_rank], f.peer_epoch);
        }
    }
}

#endif  // CERES_EXPERIMENTS

// Known-answer kernel: closest-hit traversal of caller-supplied rays.
__global__ void __launch_bounds__(256) trace_rays_kernel(const DevScene sc, const float* __restrict__ rays, uint32_t n,
                                                          uint32_t* __restrict__ out_id, float* __restrict__ out_tuv) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t k = blockIdx.x * 256u + threadIdx.x;
    if (k >= n) return;
    const float* r = rays + 8ull * k;
    Hit best; best.slot = 0; best.t = best.u = best.v = 0.0f;
    uint32_t a = 0, b = 0;
    bool found = traverse<false, false>(sc, mk(r[0], r[1], r[2]), mk(r[3], r[4], r[5]), r[6], r[7], smem_stack + threadIdx.x, 256, best, a, b);
    out_id[k] = found ? __ldg(sc.slot_id + best.slot) : CERES_MISS;
    out_tuv[3ull * k] = found ? best.t : 0.0f;
    out_tuv[3ull * k + 1] = found ? best.u : 0.0f;
    out_tuv[3ull * k + 2] = found ? best.v : 0.0f;
}

// static.cpp:135-147: rows top-down from j = height..1, uint8(max(min(p*255,255),0)) (truncation) -- one thread
// per output pixel.
__global__ void __launch_bounds__(256) quantize_kernel(const float* __restrict__ px, uint32_t W, uint32_t H, uint8_t* __restrict__ out) {
    const size_t n = (size_t)W * H;
    for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < n; k += (size_t)gridDim.x * 256u) {
        const uint32_t row_out = (uint32_t)(k / W), i = (uint32_t)(k - (size_t)row_out * W);
        const uint32_t j = H - 1u - row_out;
        const float* p = px + 3 * ((size_t)j * W + i);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float x = mul(p[c], 255.0f);
            // std::min(x, 255) = (255 < x) ? 255 : x ; std::max(y, 0) = (y < 0) ? 0 : y   (first arg returned on NaN)
            float y = (255.0f < x) ? 255.0f : x;
            float z = (y < 0.0f) ? 0.0f : y;
            out[3 * k + c] = (uint8_t)(int)z;
        }
    }
}

// Same quantisation restricted to frame rows [j0, j1) x columns [i0, i1) (the screen rectangle of the scene): the
// host-frame P6 path copies only that rectangle and fills the rest with zeros on the host.
template <int BPC>       // bytes per channel: 1 = static.cpp:135-147 (x 255), 2 = the x 65535 quantum scaling of anim.cpp:117-119, big-endian
__global__ void __launch_bounds__(256) quantize_rect_kernel(const float* __restrict__ px, const Frame f, uint32_t i0, uint32_t i1,
                                                            uint32_t r0, uint32_t r1, uint8_t* __restrict__ out) {
    // local rows [r0, r1) x columns [i0, i1) of the (possibly packed) float staging -> their place in the flipped full-size payload
    const uint32_t rw = i1 - i0;
    const size_t n = (size_t)rw * (r1 - r0);
    const float top = BPC == 1 ? 255.0f : 65535.0f;
    for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < n; k += (size_t)gridDim.x * 256u) {
        const uint32_t rj = (uint32_t)(k / rw), i = i0 + (uint32_t)(k - (size_t)rj * rw);
        const uint32_t lr = r0 + rj, j = frame_row(f, lr), row_out = f.no_flip ? j : f.H - 1u - j;
        const float* p = px + 3 * ((size_t)(f.compact ? lr : j) * f.W + i);
        uint8_t* o = out + 3 * BPC * ((size_t)row_out * f.W + i);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float x = mul(p[c], top);
            float y = (top < x) ? top : x;
            float z = (y < 0.0f) ? 0.0f : y;
            const unsigned q = (unsigned)(int)z;
            if (BPC == 1) o[c] = (uint8_t)q;
            else { o[2 * c] = (uint8_t)(q >> 8); o[2 * c + 1] = (uint8_t)(q & 0xFFu); }
        }
    }
}

#include "ray_step.cuh"
#include "frame_fused.cuh"
#ifdef CERES_EXPERIMENTS
// Wave pipelines kept for A/B measurements (profiles/kbench.py against libceres_b200_exp.so); none of them is in the
// product library.  See DESIGN.md "kernel variants" and profiles/r1/sweeps.md for what each one lost by.
#include "experiments/trace_waves.cuh"
#include "experiments/trace_batched.cuh"
#include "experiments/trace_repack.cuh"
#include "experiments/trace_shadow_spec.cuh"
#endif
#include "scene_prep.cuh"
#include "lbvh_build.cuh"

// ------------------------------------------------------------------------------------------------ host side
#define CERES_MAX_BANDS 32
#define CERES_MAX_SUB 8
struct ceres_scene {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;                     // D2H of finished row bands
    cudaEvent_t band_done[CERES_MAX_BANDS] = {};
    cudaStream_t sub_stream[CERES_MAX_SUB] = {};           // concurrent sub-frames (row bands) of one frame
    cudaEvent_t sub_done[CERES_MAX_SUB] = {};
    cudaEvent_t fork_ev = nullptr;
    DevScene dev{};
    void* d_pairs = nullptr; void* d_tris = nullptr; void* d_norms = nullptr; void* d_slot_id = nullptr;
    size_t scene_bytes = 0;
    uint32_t tri_count = 0;
    // scene preparation on the device (scene_prep.cuh): structure kept for refit / download
    size_t node_count = 0;
    uint32_t n_pairs = 0;
    std::vector<uint32_t> h_pair_first_child;               // pair -> reference node index of its left child
    std::vector<uint32_t> h_slot_id;                        // slot -> primitive index
    uint32_t* d_parent = nullptr; uint32_t* d_need = nullptr; uint32_t* d_arrived = nullptr;   // made by the first refit
    float* d_root_bounds = nullptr;
    unsigned long long* d_hist = nullptr;                   // step histogram (ceres_scene_get_step_histogram)
    unsigned long long* d_timeline = nullptr; size_t timeline_warps = 0;   // per-warp timestamps of the last statistics frame (ceres_scene_get_timeline)
    FrameCounters* d_ctr = nullptr;
    uint4* d_queue = nullptr; size_t queue_cap = 0;
    uint32_t* d_ready = nullptr; size_t ready_cap = 0;     // per-chunk fill counters of the queue (frame_fused.cuh), all-zero between launches
    cudaEvent_t frame_done = nullptr; cudaStream_t last_stream = nullptr;   // stream order between the frames of this scene
    float* d_pixels = nullptr; size_t pixels_cap = 0;      // staging for the host-buffer API
    uint32_t* d_ids = nullptr; size_t ids_cap = 0;
    uint8_t* d_rgb8 = nullptr; size_t rgb8_cap = 0;        // staging of the 8-bit host-frame path (ceres_render_ppm8)
    unsigned long long* d_out = nullptr;                    // 2 counters + 8 stats + 1 error word (FRAME_ERR_* bits)
    unsigned long long* h_out = nullptr;                    // pinned mirror
    bool stats = false;
    bool profile = false;                                   // CUDA events around each kernel of every frame
    bool profile_kernels = true;                            // false while sub-frames run concurrently (frame total only)
    std::vector<cudaEvent_t> prof_events;                   // 4 per frame: start, after primary, after shadow, after shade
    size_t prof_frames = 0;
    ceres_tuning tune{};
    int shadow_blocks_per_sm = 0, shadow_blocks_per_sm_stats = 0;
    int trace_blocks_per_sm[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // frame kernel: [stats + 2 * left-first + 4 * no-share], 8/9 timeline build; experiments: [wave*2 + stats]
    float bounds[6] = {0, 0, 0, 0, 0, 0};                   // root bounding box (nodes[0].bounds)
    uint64_t last_d2h_bytes = 0;                            // bytes DMA-copied to the host by the last host-frame render
    uint32_t* h_ext = nullptr;                              // mapped page-locked: 8 words per band (Frame::host_ext)
    // CERES_FRAME_RETAINED: what the last host frame left in the caller's buffers (rectangles per band; zeros elsewhere)
    struct Retained {
        bool valid = false;
        const void* px = nullptr; const void* ids = nullptr;
        size_t W = 0, H = 0;
        uint32_t sh[3] = {0, 0, 0}; bool compact = false;
        uint32_t i0 = 0, i1 = 0, l0 = 0, l1 = 0, nb = 0, trpb = 0;
        uint32_t ext[CERES_MAX_BANDS][4] = {};              // ci0, ci1, cr0, cr1 (inclusive) of band b; ci0 > ci1: nothing written
    } kept;
    ceres_scene* seq_view = nullptr;                        // second set of per-frame resources over the same geometry (sequences)
    bool borrowed = false;                                  // a view: the geometry belongs to the scene it was made from
    std::atomic<int> in_sequence{0};                        // ceres_render_sequence is running on this scene (its second thread reads the geometry)
    uint32_t ext_epoch = 0;
    uint64_t rays_outside = 0;                              // primary rays of the last frame that were never launched (host path)
    uint32_t cur_nsub = 1;                                  // row bands of the frame being enqueued (render_rows)
    uint32_t last_rect[4] = {0, 0, 0, 0};                   // conservative screen rectangle of the last frame: columns [0],[1]; local rows [2],[3]
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct RefNode { float bounds[6]; uint32_t primitive_count; uint32_t first_child_or_primitive; };
static_assert(sizeof(RefNode) == 32, "bvh.hpp:25-30");

// render.hpp:91-97 in binary32, source order (this translation unit's host code is compiled with
// -ffp-contract=off; tanf is the host libm's, as in the reference)
void frame_setup_host(const ceres_camera* cam, size_t width, size_t height, float* dir, float* iu, float* iv) {
    auto dot = [](const float* a, const float* b) { float s = a[0] * b[0]; s += a[1] * b[1]; s += a[2] * b[2]; return s; };
    auto normalize = [&](float* a) { float inv = 1.0f / std::sqrt(dot(a, a)); a[0] = a[0] * inv; a[1] = a[1] * inv; a[2] = a[2] * inv; };
    auto cross = [](const float* a, const float* b, float* c) {
        c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
    };
    for (int k = 0; k < 3; ++k) dir[k] = cam->dir[k];
    normalize(dir);
    cross(dir, cam->up, iu); normalize(iu);
    cross(iu, dir, iv); normalize(iv);
    float image_w = std::tan(cam->fov * float(3.14159265 * (1.0 / 180.0) * 0.5));
    float ratio = float(height) / float(width);
    for (int k = 0; k < 3; ++k) { iu[k] = iu[k] * image_w; iv[k] = iv[k] * image_w * ratio; }
}

uint32_t shard_rows(size_t H, const ceres_shard* sh) {
    if (!sh || sh->count <= 1) return (uint32_t)H;
    uint32_t n = 0;
    const uint32_t br = sh->block_rows;
    for (size_t b = sh->index; b * br < H; b += sh->count) {
        size_t lo = b * br, hi = lo + br; if (hi > H) hi = H;
        n += (uint32_t)(hi - lo);
    }
    return n;
}

template <typename T>
int ensure(T*& p, size_t& cap, size_t need) {
    if (need <= cap) return CERES_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    CUDA_TRY(cudaMalloc((void**)&p, need * sizeof(T)));
    cap = need;
    return CERES_OK;
}

// Host framebuffers: a buffer the caller page-locked (ceres_host_alloc / ceres_host_register, or any
// cudaHostAlloc'ed memory) is written by DMA straight from the device; anything else goes through the driver's
// pageable-copy path.  The library never keeps a registration of caller memory across calls: the caller may free
// and re-map the same virtual address between frames (numpy does), and a stale registration would then DMA into
// pages that no longer back the buffer.
bool host_is_pinned(const void* p) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) return attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return false;
}

}  // namespace

// Dynamic shared memory above the 48 KB default needs an opt-in per kernel (a 64-deep stack x 256 threads = 64 KB).
template <class K>
static int smem_opt_in(K kernel, size_t bytes) {
    if (bytes > 48u * 1024u) CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return CERES_OK;
}

extern "C" {

int ceres_abi_version(void) { return CERES_B200_ABI_VERSION; }
const char* ceres_last_error(void) { return g_err.c_str(); }
uint64_t ceres_kernel_launch_count(void) { return g_launches.load(); }

int ceres_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

void* ceres_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); g_err = "cudaMallocHost failed"; return nullptr; }
    return p;
}
void ceres_host_free(void* p) { if (p) cudaFreeHost(p); }
int ceres_host_register(void* p, size_t bytes) {
    if (!p || !bytes) return fail(CERES_ERR_INVALID, "ceres_host_register: null/zero argument");
    CUDA_TRY(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
    return CERES_OK;
}
int ceres_host_unregister(void* p) {
    if (!p) return fail(CERES_ERR_INVALID, "ceres_host_unregister: null argument");
    CUDA_TRY(cudaHostUnregister(p));
    return CERES_OK;
}
int ceres_host_is_pinned(const void* p) { return p && host_is_pinned(p) ? 1 : 0; }

int ceres_frame_setup(const ceres_camera* camera, size_t width, size_t height, float out[9]) {
    if (!camera || !out || !width || !height) return fail(CERES_ERR_INVALID, "ceres_frame_setup: null/zero argument");
    frame_setup_host(camera, width, height, out, out + 3, out + 6);
    return CERES_OK;
}

// A scene object with its streams, events and per-frame counters, and no geometry yet.
static int scene_new(int device, int sm_count, ceres_scene** out) {
    ceres_scene* s = new (std::nothrow) ceres_scene;
    if (!s) return fail(CERES_ERR_INVALID, "out of host memory");
    s->device = device;
    s->sm_count = sm_count;
    auto bail = [&](int code) { ceres_scene_destroy(s); return code; };
#define TRY_S(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); return bail(CERES_ERR_CUDA); } } while (0)
    TRY_S(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    TRY_S(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    for (int k = 0; k < CERES_MAX_BANDS; ++k) TRY_S(cudaEventCreateWithFlags(&s->band_done[k], cudaEventDisableTiming));
    // earlier bands get higher priority: they finish (and start their copy) first, later bands fill their tails
    int prio_lo = 0, prio_hi = 0;
    TRY_S(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));       // lo = numerically largest = least urgent
    for (int k = 0; k < CERES_MAX_SUB; ++k) {
        const int prio = std::min(prio_lo, prio_hi + k);
        TRY_S(cudaStreamCreateWithPriority(&s->sub_stream[k], cudaStreamNonBlocking, prio));
        TRY_S(cudaEventCreateWithFlags(&s->sub_done[k], cudaEventDisableTiming));
    }
    TRY_S(cudaEventCreateWithFlags(&s->fork_ev, cudaEventDisableTiming));
    TRY_S(cudaEventCreateWithFlags(&s->frame_done, cudaEventDisableTiming));
    TRY_S(cudaMalloc((void**)&s->d_ctr, CERES_MAX_SUB * sizeof(FrameCounters)));
    TRY_S(cudaMalloc((void**)&s->d_out, 11 * sizeof(unsigned long long)));
    TRY_S(cudaMallocHost((void**)&s->h_out, 11 * sizeof(unsigned long long)));
    {
        FrameCounters init[CERES_MAX_SUB] = {};
        for (auto& c : init) { c.ext[0] = 0xFFFFFFFFu; c.ext[2] = 0xFFFFFFFFu; }
        TRY_S(cudaMemcpy(s->d_ctr, init, sizeof init, cudaMemcpyHostToDevice));
    }
    TRY_S(cudaHostAlloc((void**)&s->h_ext, CERES_MAX_BANDS * 8 * sizeof(uint32_t), cudaHostAllocMapped));
    std::memset(s->h_ext, 0, CERES_MAX_BANDS * 8 * sizeof(uint32_t));
    TRY_S(cudaMemset(s->d_out, 0, 11 * sizeof(unsigned long long)));
#undef TRY_S
#ifdef CERES_CHECKED
    s->dev.check = s->d_out + 10;
#endif
    // the initialisations above ran on the legacy stream; the scene's own streams are non-blocking and not ordered after it
    if (cudaDeviceSynchronize() != cudaSuccess) { fail(CERES_ERR_CUDA, "scene_new: cudaDeviceSynchronize failed"); ceres_scene_destroy(s); return CERES_ERR_CUDA; }
    *out = s;
    return CERES_OK;
}

int ceres_scene_create(const void* nodes_v, size_t node_count, const size_t* prim_idx, const void* tris_v,
                       const void* norms_v, size_t tri_count, int device, ceres_scene** out) {
    if (!out) return fail(CERES_ERR_INVALID, "ceres_scene_create: out is null");
    *out = nullptr;
    if (!nodes_v || !prim_idx || !tris_v || !node_count || !tri_count)
        return fail(CERES_ERR_INVALID, "ceres_scene_create: null or empty scene arrays");
    if (tri_count >= 0x7FFFFFFFull || node_count >= 0xFFFFFFFFull)
        return fail(CERES_ERR_INVALID, "ceres_scene_create: scene too large for 31-bit slots");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return fail(CERES_ERR_NO_DEVICE, "no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(CERES_ERR_INVALID, "ceres_scene_create: bad device index");

    const RefNode* nodes = static_cast<const RefNode*>(nodes_v);
    const float* tris = static_cast<const float*>(tris_v);
    const float* norms = static_cast<const float*>(norms_v);

    // ---- flatten: walk the tree from the root, give every inner node's children one pair record ----------------
    std::vector<float> pairs;            // 16 floats per pair
    std::vector<uint32_t> pair_first_child;
    uint32_t root_leaf_count = 0, depth_max = 1;
    std::vector<uint32_t> slot_id(tri_count);
    for (size_t k = 0; k < tri_count; ++k) {
        if (prim_idx[k] >= tri_count) return fail(CERES_ERR_INVALID, "ceres_scene_create: primitive index out of range");
        slot_id[k] = (uint32_t)prim_idx[k];
    }
    const RefNode& root = nodes[0];
    if (root.primitive_count != 0) {
        if (root.first_child_or_primitive != 0 || root.primitive_count != tri_count)
            return fail(CERES_ERR_INVALID, "ceres_scene_create: leaf root must cover all primitives from slot 0");
        root_leaf_count = root.primitive_count;
        pairs.assign(16, 0.0f);
    } else {
        // depth-first, left child first: pair index = order of first visit.  (pair_of_node, depth) work list.
        struct Item { uint32_t first_child; uint32_t pair; uint32_t depth; };
        std::vector<Item> work;
        pairs.reserve((node_count / 2 + 1) * 16);
        pair_first_child.reserve(node_count / 2 + 1);
        auto new_pair = [&](uint32_t first_child) { uint32_t id = (uint32_t)(pairs.size() / 16); pairs.resize(pairs.size() + 16); pair_first_child.push_back(first_child); return id; };
        size_t visited = 1;
        work.push_back({root.first_child_or_primitive, new_pair(root.first_child_or_primitive), 1});
        while (!work.empty()) {
            Item it = work.back(); work.pop_back();
            if ((size_t)it.first_child + 1 >= node_count)
                return fail(CERES_ERR_INVALID, "ceres_scene_create: child index out of range");
            visited += 2;
            if (visited > node_count) return fail(CERES_ERR_INVALID, "ceres_scene_create: BVH is not a tree (cycle or shared nodes)");
            if (it.depth > depth_max) depth_max = it.depth;
            uint32_t child_pair[2] = {0, 0};
            for (int c = 0; c < 2; ++c) {
                const RefNode& n = nodes[it.first_child + c];
                float* q = &pairs[(size_t)it.pair * 16 + 8 * c];
                for (int b = 0; b < 6; ++b) q[b] = n.bounds[b];
                uint32_t cnt = n.primitive_count, first = n.first_child_or_primitive;
                if (cnt != 0) {
                    if ((size_t)first + cnt > tri_count) return fail(CERES_ERR_INVALID, "ceres_scene_create: leaf range out of bounds");
                } else {
                    first = new_pair(first);
                    child_pair[c] = first;
                }
                std::memcpy(&pairs[(size_t)it.pair * 16 + 8 * c + 6], &cnt, 4);
                std::memcpy(&pairs[(size_t)it.pair * 16 + 8 * c + 7], &first, 4);
            }
            for (int c = 1; c >= 0; --c) {
                const RefNode& n = nodes[it.first_child + c];
                if (n.primitive_count == 0) work.push_back({n.first_child_or_primitive, child_pair[c], it.depth + 1});
            }
        }
    }
    if (depth_max > 64) return fail(CERES_ERR_UNSUPPORTED, "ceres_scene_create: BVH deeper than the reference's 64-entry stack");

    // leaf-ordered triangles and normals
    std::vector<float> tri_slots(tri_count * 12), norm_slots(tri_count * 12, 0.0f);
    for (size_t s = 0; s < tri_count; ++s) {
        std::memcpy(&tri_slots[s * 12], tris + (size_t)slot_id[s] * 12, 48);
        if (norms) std::memcpy(&norm_slots[s * 12], norms + (size_t)slot_id[s] * 9, 36);
    }

    DeviceGuard guard(device);
    if (!guard.ok) return fail(CERES_ERR_CUDA, "cudaSetDevice failed");
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(CERES_ERR_NO_DEVICE, std::string("device is not Blackwell-class (sm_100a required): ") + prop.name);

    ceres_scene* s = nullptr;
    if (int rc = scene_new(device, prop.multiProcessorCount, &s)) return rc;
    s->tri_count = (uint32_t)tri_count;
    for (int b = 0; b < 6; ++b) s->bounds[b] = root.bounds[b];
    auto bail = [&](int code) { ceres_scene_destroy(s); return code; };
#define TRY_S(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); return bail(CERES_ERR_CUDA); } } while (0)
    const size_t pb = pairs.size() * 4, tb = tri_slots.size() * 4, nb = norm_slots.size() * 4, ib = slot_id.size() * 4;
    TRY_S(cudaMalloc(&s->d_pairs, pb));
    TRY_S(cudaMalloc(&s->d_tris, tb));
    TRY_S(cudaMalloc(&s->d_norms, nb));
    TRY_S(cudaMalloc(&s->d_slot_id, ib));
    TRY_S(cudaMemcpy(s->d_pairs, pairs.data(), pb, cudaMemcpyHostToDevice));
    TRY_S(cudaMemcpy(s->d_tris, tri_slots.data(), tb, cudaMemcpyHostToDevice));
    TRY_S(cudaMemcpy(s->d_norms, norm_slots.data(), nb, cudaMemcpyHostToDevice));
    TRY_S(cudaMemcpy(s->d_slot_id, slot_id.data(), ib, cudaMemcpyHostToDevice));
#undef TRY_S
    s->scene_bytes = pb + tb + nb + ib;
    s->dev.pairs = (const float4*)s->d_pairs;
    s->dev.tris = (const float4*)s->d_tris;
    s->dev.norms = (const float4*)s->d_norms;
    s->dev.slot_id = (const uint32_t*)s->d_slot_id;
    s->dev.root_leaf_count = root_leaf_count;
    s->dev.stack_depth = depth_max;
    s->dev.n_pairs = (uint32_t)(pairs.size() / 16); s->dev.n_slots = (uint32_t)tri_count;
#ifdef CERES_CHECKED
    if (std::getenv("CERES_CHECK_SELFTEST")) s->dev.n_slots = (uint32_t)(tri_count / 2);    // the checks must fire: proves they are live
#endif
    s->node_count = node_count;
    s->n_pairs = root_leaf_count ? 0u : (uint32_t)(pairs.size() / 16);
    s->h_pair_first_child = std::move(pair_first_child);
    s->h_slot_id = std::move(slot_id);
    CUDA_TRY(cudaDeviceSynchronize());          // uploads on the legacy stream are complete before a non-blocking stream reads them
    *out = s;
    return CERES_OK;
}

void ceres_scene_destroy(ceres_scene* s) {
    if (!s) return;
    DeviceGuard guard(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->copy_stream) { cudaStreamSynchronize(s->copy_stream); cudaStreamDestroy(s->copy_stream); }
    for (int k = 0; k < CERES_MAX_BANDS; ++k) if (s->band_done[k]) cudaEventDestroy(s->band_done[k]);
    for (int k = 0; k < CERES_MAX_SUB; ++k) {
        if (s->sub_stream[k]) { cudaStreamSynchronize(s->sub_stream[k]); cudaStreamDestroy(s->sub_stream[k]); }
        if (s->sub_done[k]) cudaEventDestroy(s->sub_done[k]);
    }
    if (s->fork_ev) cudaEventDestroy(s->fork_ev);
    if (s->frame_done) cudaEventDestroy(s->frame_done);
    if (s->seq_view) { ceres_scene_destroy(s->seq_view); s->seq_view = nullptr; }
    if (s->borrowed) { s->d_pairs = s->d_tris = s->d_norms = s->d_slot_id = nullptr; s->d_parent = s->d_need = s->d_arrived = nullptr; s->d_root_bounds = nullptr; s->d_hist = nullptr; }
    cudaFree(s->d_hist); cudaFree(s->d_timeline);
    cudaFree(s->d_parent); cudaFree(s->d_need); cudaFree(s->d_arrived); cudaFree(s->d_root_bounds);
    cudaFree(s->d_pairs); cudaFree(s->d_tris); cudaFree(s->d_norms); cudaFree(s->d_slot_id);
    cudaFree(s->d_ctr); cudaFree(s->d_queue); cudaFree(s->d_ready); cudaFree(s->d_pixels); cudaFree(s->d_ids); cudaFree(s->d_rgb8); cudaFree(s->d_out);
    if (s->h_out) cudaFreeHost(s->h_out);
    if (s->h_ext) cudaFreeHost(s->h_ext);
    for (cudaEvent_t e : s->prof_events) cudaEventDestroy(e);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
}

size_t ceres_scene_device_bytes(const ceres_scene* s) { return s ? s->scene_bytes : 0; }
uint32_t ceres_scene_stack_depth(const ceres_scene* s) { return s ? s->dev.stack_depth : 0u; }
size_t ceres_scene_node_count(const ceres_scene* s) { return s ? s->node_count : 0; }
uint64_t ceres_scene_last_d2h_bytes(const ceres_scene* s) { return s ? s->last_d2h_bytes : 0; }

// ---- scene preparation on the device (scene_prep.cuh) ---------------------------------------------
// All four calls are synchronous on the scene's stream and must not overlap frames of this scene in flight.
static int prep_triangles(ceres_scene* s, const PrepParams& q) {
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_scene_rotate/transform");
    DeviceGuard guard(s->device);
    const unsigned grid = (unsigned)std::min<size_t>(((size_t)s->tri_count + 255) / 256, (size_t)s->sm_count * 16);
    prep_triangles_kernel<<<grid, 256, 0, s->stream>>>((float4*)s->d_tris, s->tri_count, q);
    g_launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return CERES_OK;
}

int ceres_scene_rotate(ceres_scene* s, int axis, float degrees) {
    if (!s || axis < 0 || axis > 2) return fail(CERES_ERR_INVALID, "ceres_scene_rotate: bad argument");
    PrepParams q{};
    q.mode = axis;
    const float pi = 3.14159265359f;                       // render.hpp:26-28, float arithmetic, host cosf/sinf
    q.c = std::cos(degrees * pi / 180.0f);
    q.s = std::sin(degrees * pi / 180.0f);
    return prep_triangles(s, q);
}

int ceres_scene_transform(ceres_scene* s, const float a[9], const float v[3]) {
    if (!s || !a || !v) return fail(CERES_ERR_INVALID, "ceres_scene_transform: null argument");
    PrepParams q{};
    q.mode = PREP_AFFINE;
    for (int k = 0; k < 9; ++k) q.a[k] = a[k];
    for (int k = 0; k < 3; ++k) q.v[k] = v[k];
    return prep_triangles(s, q);
}

int ceres_scene_refit(ceres_scene* s) {
    if (!s) return fail(CERES_ERR_INVALID, "ceres_scene_refit: null scene");
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_scene_refit");
    DeviceGuard guard(s->device);
    if (!s->d_root_bounds) CUDA_TRY(cudaMalloc((void**)&s->d_root_bounds, 6 * sizeof(float)));
    if (s->dev.root_leaf_count) {
        // the root is the only node: its box is the leaf box of all triangles (one thread; such scenes are tiny)
        return fail(CERES_ERR_UNSUPPORTED, "ceres_scene_refit: single-leaf scenes have no hierarchy to refit");
    }
    const unsigned grid = (s->n_pairs + 255u) / 256u;
    if (!s->d_parent) {
        CUDA_TRY(cudaMalloc((void**)&s->d_parent, (size_t)s->n_pairs * 4));
        CUDA_TRY(cudaMalloc((void**)&s->d_need, (size_t)s->n_pairs * 4));
        CUDA_TRY(cudaMalloc((void**)&s->d_arrived, (size_t)s->n_pairs * 4));
        CUDA_TRY(cudaMemsetAsync(s->d_arrived, 0, (size_t)s->n_pairs * 4, s->stream));
        refit_links_kernel<<<grid, 256, 0, s->stream>>>((const float4*)s->d_pairs, s->n_pairs, s->d_parent, s->d_need);
        g_launches++;
    }
    refit_kernel<<<grid, 256, 0, s->stream>>>((float*)s->d_pairs, (const float4*)s->d_tris, s->d_parent, s->d_need, s->d_arrived,
                                              s->n_pairs, s->d_root_bounds);
    g_launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(s->bounds, s->d_root_bounds, 6 * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));       // s->bounds feeds the screen-rectangle culling of the next frame
    return CERES_OK;
}

int ceres_scene_download(ceres_scene* s, void* nodes_out, size_t node_count, void* tris_out, size_t tri_count) {
    if (!s) return fail(CERES_ERR_INVALID, "ceres_scene_download: null scene");
    if ((nodes_out && node_count != s->node_count) || (tris_out && tri_count != s->tri_count))
        return fail(CERES_ERR_INVALID, "ceres_scene_download: node/triangle count differs from the scene's");
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (tris_out) {
        std::vector<float> slots((size_t)s->tri_count * 12);
        CUDA_TRY(cudaMemcpy(slots.data(), s->d_tris, slots.size() * 4, cudaMemcpyDeviceToHost));
        float* out = static_cast<float*>(tris_out);
        for (size_t k = 0; k < s->tri_count; ++k) std::memcpy(out + (size_t)s->h_slot_id[k] * 12, &slots[k * 12], 48);
    }
    if (nodes_out) {
        RefNode* out = static_cast<RefNode*>(nodes_out);
        if (s->dev.root_leaf_count) {
            for (int b = 0; b < 6; ++b) out[0].bounds[b] = s->bounds[b];
            out[0].primitive_count = s->dev.root_leaf_count; out[0].first_child_or_primitive = 0;
            return CERES_OK;
        }
        std::vector<float> pairs((size_t)s->n_pairs * 16);
        CUDA_TRY(cudaMemcpy(pairs.data(), s->d_pairs, pairs.size() * 4, cudaMemcpyDeviceToHost));
        for (int b = 0; b < 6; ++b) out[0].bounds[b] = s->bounds[b];
        out[0].primitive_count = 0; out[0].first_child_or_primitive = s->h_pair_first_child[0];
        for (size_t p = 0; p < s->n_pairs; ++p) {
            for (int c = 0; c < 2; ++c) {
                RefNode& n = out[(size_t)s->h_pair_first_child[p] + c];
                const float* q = &pairs[p * 16 + 8 * c];
                for (int b = 0; b < 6; ++b) n.bounds[b] = q[b];
                uint32_t cnt, first;
                std::memcpy(&cnt, q + 6, 4); std::memcpy(&first, q + 7, 4);
                n.primitive_count = cnt;
                n.first_child_or_primitive = cnt ? first : s->h_pair_first_child[first];
            }
        }
    }
    return CERES_OK;
}

// ---- BVH construction on the GPU (lbvh_build.cuh) -----------------------------------------------------
namespace {
// Working set of one GPU build (lbvh_build.cuh); everything stays on the device until the caller picks what it needs.
int g_build_algorithm = 2, g_build_radius = 8;      // ceres_build_options: 2 = top-down binned SAH (default), 3 = PLOC over the Morton order, 1 = radix tree only
int g_build_max_leaf = 8; float g_build_ci = 1.5f;   // leaf collapsing: triangles per leaf at most (1 = off), cost of a triangle test in node steps
struct LbvhWork {
    float* box = nullptr; unsigned int* scene = nullptr;
    uint32_t* cid[2] = {nullptr, nullptr}; float* cbox[2] = {nullptr, nullptr}; uint32_t* nn = nullptr;       // PLOC
    unsigned long long *flag = nullptr, *scan = nullptr; PlocState* st = nullptr; void* scan_tmp = nullptr; size_t scan_tmp_bytes = 0;
    int rounds = 0;
    SahTask* s_task = nullptr; uint32_t *s_node_of = nullptr, *s_open[2] = {nullptr, nullptr}, *s_open_few[2] = {nullptr, nullptr};   // top-down SAH
    uint32_t *s_bins = nullptr, *s_bins_few = nullptr; SahState* s_st = nullptr;
    uint32_t *c_cnt = nullptr, *c_col = nullptr, *c_rank = nullptr, *c_arr = nullptr, *idx3 = nullptr; float* c_cost = nullptr;   // leaf collapsing
    uint32_t *c_keep = nullptr, *c_newid = nullptr, *c_fl = nullptr; LbvhNode* nodes2 = nullptr;
    size_t node_count = 0;                                   // nodes of the final tree (2 leaves - 1)
    uint32_t *code = nullptr, *idx = nullptr, *code2 = nullptr, *idx2 = nullptr, *child = nullptr, *pi = nullptr, *pl = nullptr, *arr = nullptr;
    LbvhNode* nodes = nullptr; void* tmp = nullptr; size_t tmp_bytes = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ~LbvhWork() {
        cudaFree(box); cudaFree(scene); cudaFree(code); cudaFree(idx); cudaFree(code2); cudaFree(idx2);
        cudaFree(child); cudaFree(pi); cudaFree(pl); cudaFree(arr); cudaFree(nodes); cudaFree(tmp);
        cudaFree(c_cnt); cudaFree(c_col); cudaFree(c_rank); cudaFree(c_arr); cudaFree(idx3); cudaFree(c_cost); cudaFree(c_keep); cudaFree(c_newid); cudaFree(c_fl);
        cudaFree(nodes2);
        cudaFree(s_task); cudaFree(s_node_of); cudaFree(s_open[0]); cudaFree(s_open[1]); cudaFree(s_open_few[0]); cudaFree(s_open_few[1]);
        cudaFree(s_bins); cudaFree(s_bins_few); cudaFree(s_st);
        cudaFree(cid[0]); cudaFree(cid[1]); cudaFree(cbox[0]); cudaFree(cbox[1]); cudaFree(nn); cudaFree(flag); cudaFree(scan); cudaFree(st); cudaFree(scan_tmp);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    }
};

// Builds the tree of the n triangles at d_tris (legacy stream); w.nodes / w.idx2 hold the reference layout afterwards.
int lbvh_run(const float4* d_tris, int n, LbvhWork& w, float* ms_out) {
    const size_t tc = (size_t)n;
#define TRY_B(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); } while (0)
    TRY_B(cudaMalloc((void**)&w.box, tc * 24));
    TRY_B(cudaMalloc((void**)&w.scene, 6 * sizeof(unsigned int)));
    TRY_B(cudaMalloc((void**)&w.code, tc * 4)); TRY_B(cudaMalloc((void**)&w.idx, tc * 4));
    TRY_B(cudaMalloc((void**)&w.code2, tc * 4)); TRY_B(cudaMalloc((void**)&w.idx2, tc * 4));
    TRY_B(cudaMalloc((void**)&w.child, 2 * tc * 4)); TRY_B(cudaMalloc((void**)&w.pi, tc * 4));
    TRY_B(cudaMalloc((void**)&w.pl, tc * 4)); TRY_B(cudaMalloc((void**)&w.arr, tc * 4));
    TRY_B(cudaMalloc((void**)&w.nodes, (2 * tc - 1) * sizeof(LbvhNode)));
    TRY_B(cub::DeviceRadixSort::SortPairs(nullptr, w.tmp_bytes, w.code, w.code2, w.idx, w.idx2, n, 0, 30));
    TRY_B(cudaMalloc(&w.tmp, w.tmp_bytes ? w.tmp_bytes : 16));
    const bool sah = g_build_algorithm == 2 && n > 2;
    const bool ploc = g_build_algorithm == 3 && n > 2;
    const size_t sah_max_open = tc / (SAH_FEW + 1) + 1, sah_max_few = tc / 3 + 1;      // an open task has more than SAH_FEW members, or at least 3
    if (sah) {
        TRY_B(cudaMalloc((void**)&w.cid[0], tc * 4)); TRY_B(cudaMalloc((void**)&w.cbox[0], tc * 24));
        TRY_B(cudaMalloc((void**)&w.s_task, (2 * tc + 2) * sizeof(SahTask)));
        TRY_B(cudaMalloc((void**)&w.s_node_of, tc * 4));
        for (int k = 0; k < 2; ++k) { TRY_B(cudaMalloc((void**)&w.s_open[k], sah_max_open * 4)); TRY_B(cudaMalloc((void**)&w.s_open_few[k], sah_max_few * 4)); }
        TRY_B(cudaMalloc((void**)&w.s_bins, sah_max_open * SAH_TASK_WORDS * 4));
        TRY_B(cudaMalloc((void**)&w.s_bins_few, sah_max_few * SAH_TASK_WORDS_FEW * 4));
        TRY_B(cudaMalloc((void**)&w.s_st, sizeof(SahState)));
    }
    if (ploc) {
        for (int k = 0; k < 2; ++k) { TRY_B(cudaMalloc((void**)&w.cid[k], tc * 4)); TRY_B(cudaMalloc((void**)&w.cbox[k], tc * 24)); }
        TRY_B(cudaMalloc((void**)&w.nn, tc * 4));
        TRY_B(cudaMalloc((void**)&w.flag, tc * 8)); TRY_B(cudaMalloc((void**)&w.scan, tc * 8));
        TRY_B(cudaMalloc((void**)&w.st, sizeof(PlocState)));
        TRY_B(cub::DeviceScan::ExclusiveSum(nullptr, w.scan_tmp_bytes, w.flag, w.scan, n));
        TRY_B(cudaMalloc(&w.scan_tmp, w.scan_tmp_bytes ? w.scan_tmp_bytes : 16));
    }
    const bool collapse = n > 2;        // max_leaf 1: nothing collapses, the pass still puts leaves and inner nodes in depth-first order
    if (collapse) {
        TRY_B(cudaMalloc((void**)&w.c_cnt, tc * 4)); TRY_B(cudaMalloc((void**)&w.c_col, tc * 4)); TRY_B(cudaMalloc((void**)&w.c_rank, tc * 4));
        TRY_B(cudaMalloc((void**)&w.c_arr, tc * 4)); TRY_B(cudaMalloc((void**)&w.idx3, tc * 4)); TRY_B(cudaMalloc((void**)&w.c_cost, tc * 4));
        TRY_B(cudaMalloc((void**)&w.c_keep, tc * 4)); TRY_B(cudaMalloc((void**)&w.c_newid, tc * 4)); TRY_B(cudaMalloc((void**)&w.c_fl, tc * 4));
        TRY_B(cudaMalloc((void**)&w.nodes2, (2 * tc - 1) * sizeof(LbvhNode)));
    }
    w.node_count = 2 * tc - 1;
    TRY_B(cudaEventCreate(&w.e0)); TRY_B(cudaEventCreate(&w.e1));
    const unsigned int scene_init[6] = {0xFFFFFFFFu, 0u, 0xFFFFFFFFu, 0u, 0xFFFFFFFFu, 0u};
    TRY_B(cudaMemcpy(w.scene, scene_init, sizeof scene_init, cudaMemcpyHostToDevice));
    TRY_B(cudaEventRecord(w.e0, 0));
    const unsigned grid = (unsigned)((tc + 255) / 256);
    lbvh_boxes_kernel<<<grid, 256>>>(d_tris, (uint32_t)n, w.box, w.scene);
    lbvh_codes_kernel<<<grid, 256>>>(w.box, (uint32_t)n, w.scene, w.code, w.idx);
    TRY_B(cub::DeviceRadixSort::SortPairs(w.tmp, w.tmp_bytes, w.code, w.code2, w.idx, w.idx2, n, 0, 30));
    g_launches += 2;
    if (n > 1) {
        TRY_B(cudaMemsetAsync(w.arr, 0, tc * 4, 0));
        if (ploc) {
            // locally-ordered clustering over the Morton order (lbvh_build.cuh): rounds until one cluster is left
            const PlocState init{(uint32_t)n, (uint32_t)(n - 1), 0u, 0u};
            TRY_B(cudaMemcpyAsync(w.st, &init, sizeof init, cudaMemcpyHostToDevice, 0));
            ploc_init_kernel<<<grid, 256>>>(w.idx2, w.box, (uint32_t)n, w.cid[0], w.cbox[0]);
            g_launches++;
            uint32_t m = (uint32_t)n;
            int cur = 0;
            while (m > 1) {
                const unsigned g = (m + 255u) / 256u;
                ploc_nn_kernel<<<g, 256>>>(w.cbox[cur], w.st, g_build_radius, w.nn);
                ploc_flag_kernel<<<g, 256>>>(w.nn, w.st, w.flag);
                TRY_B(cub::DeviceScan::ExclusiveSum(w.scan_tmp, w.scan_tmp_bytes, w.flag, w.scan, (int)m));
                ploc_totals_kernel<<<1, 1>>>(w.flag, w.scan, w.st);
                ploc_merge_kernel<<<g, 256>>>(w.cid[cur], w.cbox[cur], w.nn, w.flag, w.scan, w.st, w.cid[cur ^ 1], w.cbox[cur ^ 1], w.child, w.pi, w.pl);
                ploc_advance_kernel<<<1, 1>>>(w.st);
                g_launches += 5;
                PlocState now;
                TRY_B(cudaMemcpy(&now, w.st, sizeof now, cudaMemcpyDeviceToHost));
                if (now.m >= m || now.m == 0) return fail(CERES_ERR_CUDA, "ceres_build_lbvh: clustering made no progress (internal error)");
                m = now.m; cur ^= 1; ++w.rounds;
            }
        } else if (sah) {
            // top-down binned SAH over the Morton-sorted triangles, one level per iteration (lbvh_build.cuh)
            ploc_init_kernel<<<grid, 256>>>(w.idx2, w.box, (uint32_t)n, w.cid[0], w.cbox[0]);
            const SahState init{1u, 1u, 0u, 0u, 0u, {0u, 0u, 0u}};
            TRY_B(cudaMemcpyAsync(w.s_st, &init, sizeof init, cudaMemcpyHostToDevice, 0));
            const uint32_t zero_open = 0u;
            const bool root_few = (uint32_t)n <= SAH_FEW;
            TRY_B(cudaMemcpyAsync(root_few ? w.s_open_few[0] : w.s_open[0], &zero_open, 4, cudaMemcpyHostToDevice, 0));      // level 0: the root task
            sah_root_kernel<<<grid, 256>>>(w.scene, (uint32_t)n, w.s_task, w.s_node_of);
            g_launches += 2;
            uint32_t n_open = root_few ? 0u : 1u, n_few = root_few ? 1u : 0u, first_new = 1u;
            int cur = 0;
            int sms = 148;
            { int dev_ = 0; if (cudaGetDevice(&dev_) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_); }
            const unsigned persistent = (unsigned)sms * 4u;        // 43 KB of shared memory per CTA at most: 4-5 CTAs per SM
            static const bool trace = std::getenv("CERES_BUILD_TRACE") != nullptr;     // per-level wall time (every level ends with a readback)
            static const bool cta_allowed = std::getenv("CERES_BUILD_WARP_PASSES") == nullptr;   // A/B: the warp-aggregated passes at every level
            const unsigned cta_grid = (unsigned)((tc + SAH_CTA_MEMBERS - 1) / SAH_CTA_MEMBERS);
            auto t_prev = std::chrono::steady_clock::now();
            while (n_open + n_few) {
                if (++w.rounds > 160) return fail(CERES_ERR_CUDA, "ceres_build_bvh: the top-down build did not finish in 160 levels (internal error)");
                if (n_open) TRY_B(cudaMemsetAsync(w.s_bins, 0, (size_t)n_open * SAH_TASK_WORDS * 4, 0));
                if (n_few) TRY_B(cudaMemsetAsync(w.s_bins_few, 0, (size_t)n_few * SAH_TASK_WORDS_FEW * 4, 0));
                const bool small = !n_few && n_open <= SAH_SMALL_LEVEL && grid > persistent;       // few tasks, many members each: privatised in shared memory
                // many tasks of a few hundred members or less: a CTA's members meet a handful of records each -> CTA-level accumulation
                // (measured on the 10.5 M-triangle scene: better from the level with 320 members per task on, worse above)
                const bool cta_pass = cta_allowed && (uint64_t)(n_open + n_few) * 320u >= (uint64_t)n;
                if (small) sah_bin_small_kernel<<<persistent, 256, (size_t)n_open * SAH_TASK_WORDS * 4>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of, w.s_bins, n_open);
                else if (cta_pass) sah_bin_cta_kernel<<<cta_grid, 256>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of, w.s_bins, w.s_bins_few);
                else sah_bin_kernel<<<grid, 256>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of, w.s_bins, w.s_bins_few);
                if (n_open) sah_split_kernel<SAH_BINS><<<(n_open + 127u) / 128u, 128>>>(w.s_open[cur], n_open, w.s_task, w.s_bins, w.s_st);
                if (n_few) sah_split_kernel<SAH_BINS_FEW><<<(n_few + 127u) / 128u, 128>>>(w.s_open_few[cur], n_few, w.s_task, w.s_bins_few, w.s_st);
                if (small) sah_assign_small_kernel<<<persistent, 256, (size_t)2 * n_open * SAH_ACC_WORDS * 4>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of, w.s_open[cur], n_open);
                else if (cta_pass) sah_assign_cta_kernel<<<cta_grid, 256>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of);
                else sah_assign_kernel<<<grid, 256>>>(w.cbox[0], (uint32_t)n, w.s_task, w.s_node_of);
                const uint32_t created = 2u * (n_open + n_few);
                sah_link_kernel<<<(created + 255u) / 256u, 256>>>(first_new, created, w.s_task, w.cid[0], w.child, w.pi, w.pl, w.s_open[cur ^ 1], w.s_open_few[cur ^ 1], w.s_st);
                g_launches += 5;
                SahState now;
                TRY_B(cudaMemcpy(&now, w.s_st, sizeof now, cudaMemcpyDeviceToHost));
                if (now.error || now.next_task != first_new + created || now.n_open > sah_max_open || now.n_open_few > sah_max_few)
                    return fail(CERES_ERR_CUDA, "ceres_build_bvh: the top-down build lost a member (internal error)");
                if (trace) {
                    const auto t_now = std::chrono::steady_clock::now();
                    std::fprintf(stderr, "sah level %d: %u + %u open tasks -> %u + %u, %.3f ms\n", w.rounds, n_open, n_few, now.n_open, now.n_open_few,
                                 std::chrono::duration<double, std::milli>(t_now - t_prev).count());
                    t_prev = t_now;
                }
                first_new += created; n_open = now.n_open; n_few = now.n_open_few; cur ^= 1;
                const uint32_t z[2] = {0u, 0u};
                TRY_B(cudaMemcpy(&w.s_st->n_open, z, 8, cudaMemcpyHostToDevice));
                if (!(n_open + n_few) && now.next_inner != (uint32_t)n - 1u) return fail(CERES_ERR_CUDA, "ceres_build_bvh: the top-down build made the wrong number of nodes (internal error)");
            }
        } else
        lbvh_tree_kernel<<<grid, 256>>>(w.code2, n, w.child, w.pi, w.pl);
        lbvh_emit_kernel<<<grid, 256>>>(w.child, w.idx2, w.box, n, w.nodes);
        if (!collapse) lbvh_fit_kernel<<<grid, 256>>>(w.pi, w.pl, w.arr, n, w.nodes);
        g_launches += 3;
        if (collapse) {
            // bottom-up boxes + SAH leaf collapsing (one climb), depth-first leaf order, compaction of the surviving inner nodes
            const CollapseParams q{1.0f, g_build_ci, (uint32_t)g_build_max_leaf};
            TRY_B(cudaMemsetAsync(w.nodes2, 0, (2 * tc - 1) * sizeof(LbvhNode), 0));
            bvhc_fit_cost_kernel<<<grid, 256>>>(w.child, w.pi, w.pl, w.nodes, n, q, w.arr, w.c_cnt, w.c_cost, w.c_col, w.c_fl);
            bvhc_rank_kernel<<<(unsigned)((2 * tc - 1 + 255) / 256), 256>>>(w.child, w.pi, w.pl, w.c_cnt, w.c_col, w.c_fl, w.idx2, n, w.idx3, w.c_rank, w.c_keep, w.c_newid);
            bvhc_emit_kernel<<<grid, 256>>>(w.child, w.c_cnt, w.c_col, w.c_rank, w.c_keep, w.c_newid, w.nodes, n, w.nodes2);
            g_launches += 3;
            uint32_t leaves = 0;                                  // leaves of the final tree = fl[root]
            TRY_B(cudaMemcpy(&leaves, w.c_fl, 4, cudaMemcpyDeviceToHost));
            w.node_count = 2 * (size_t)leaves - 1;
            std::swap(w.nodes, w.nodes2); std::swap(w.idx2, w.idx3);
        }
    }
    TRY_B(cudaEventRecord(w.e1, 0));
    TRY_B(cudaGetLastError());
    TRY_B(cudaEventSynchronize(w.e1));
    float ms = 0.0f;
    TRY_B(cudaEventElapsedTime(&ms, w.e0, w.e1));
    if (ms_out) *ms_out = ms;
#undef TRY_B
    return CERES_OK;
}

int lbvh_check_args(const void* tris, size_t tri_count, int device, const char* who) {
    if (!tris || !tri_count) return fail(CERES_ERR_INVALID, std::string(who) + ": null or empty argument");
    if (tri_count >= 0x40000000ull) return fail(CERES_ERR_INVALID, std::string(who) + ": too many triangles");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return fail(CERES_ERR_NO_DEVICE, "no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(CERES_ERR_INVALID, std::string(who) + ": bad device index");
    return CERES_OK;
}
}  // namespace

int ceres_build_options(int algorithm, int radius, int max_leaf, float triangle_cost) {
    if (algorithm < 0 || algorithm > 3 || radius < 0 || radius > 256 || max_leaf < 0 || max_leaf > 16 || !(triangle_cost >= 0.0f))
        return fail(CERES_ERR_INVALID, "ceres_build_options: algorithm 0 = 2 (top-down binned SAH), 1 (radix tree) or 3 (clustering), radius 1..256, max_leaf 1..16, triangle_cost > 0 (0 = defaults)");
    g_build_algorithm = algorithm ? algorithm : 2;
    g_build_radius = radius ? radius : 8;
    g_build_max_leaf = max_leaf ? max_leaf : 8;
    g_build_ci = triangle_cost > 0.0f ? triangle_cost : 1.5f;
    return CERES_OK;
}

int ceres_build_lbvh(const void* tris, size_t tri_count, int device, void* nodes_out, size_t* prim_idx_out, double* build_ms) {
    return ceres_build_bvh(tris, tri_count, device, nodes_out, nullptr, prim_idx_out, build_ms);
}

int ceres_build_bvh(const void* tris, size_t tri_count, int device, void* nodes_out, size_t* node_count_out, size_t* prim_idx_out, double* build_ms) {
    if (!nodes_out || !prim_idx_out) return fail(CERES_ERR_INVALID, "ceres_build_lbvh: null or empty argument");
    if (int rc = lbvh_check_args(tris, tri_count, device, "ceres_build_lbvh")) return rc;
    DeviceGuard guard(device);
    const int n = (int)tri_count;
    float4* d_tris = nullptr;
    LbvhWork w;
    float ms = 0.0f;
    int rc = CERES_OK;
    if (cudaMalloc((void**)&d_tris, tri_count * 48) != cudaSuccess || cudaMemcpy(d_tris, tris, tri_count * 48, cudaMemcpyHostToDevice) != cudaSuccess)
        rc = fail(CERES_ERR_CUDA, std::string("ceres_build_lbvh: triangle upload: ") + cudaGetErrorString(cudaGetLastError()));
    if (!rc) rc = lbvh_run(d_tris, n, w, &ms);
    if (!rc && n == 1) {
        LbvhNode root;
        if (cudaMemcpy(root.bounds, w.box, 24, cudaMemcpyDeviceToHost) != cudaSuccess) rc = fail(CERES_ERR_CUDA, "ceres_build_lbvh: download");
        root.primitive_count = 1; root.first_child_or_primitive = 0;
        std::memcpy(nodes_out, &root, sizeof root);
        prim_idx_out[0] = 0;
        ms = 0.0f;
        if (node_count_out) *node_count_out = 1;
    } else if (!rc) {
        if (node_count_out) *node_count_out = w.node_count;
        std::vector<uint32_t> idx(tri_count);
        if (cudaMemcpy(nodes_out, w.nodes, (2 * tri_count - 1) * sizeof(LbvhNode), cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(idx.data(), w.idx2, tri_count * 4, cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = fail(CERES_ERR_CUDA, std::string("ceres_build_lbvh: download: ") + cudaGetErrorString(cudaGetLastError()));
        for (size_t k = 0; k < tri_count; ++k) prim_idx_out[k] = idx[k];
    }
    if (build_ms) *build_ms = ms;
    cudaFree(d_tris);
    return rc;
}

// ---- a scene straight from the GPU build: the tree never visits the host ------------------------------------------------
// pair i = the two children of radix-tree inner node i (reference nodes 1 + 2 i, 2 + 2 i); an inner child's
// first_child_or_primitive 1 + 2 c becomes pair c.
__global__ void __launch_bounds__(256) lbvh_pairs_kernel(const LbvhNode* __restrict__ nodes, int n, float4* __restrict__ pairs) {
    const int i = (int)(blockIdx.x * 256u + threadIdx.x);
    if (i >= n - 1) return;
    for (int c = 0; c < 2; ++c) {
        const LbvhNode nd = nodes[1 + 2 * i + c];
        const uint32_t first = nd.primitive_count ? nd.first_child_or_primitive : (nd.first_child_or_primitive - 1u) >> 1;
        pairs[4 * (size_t)i + 2 * c] = make_float4(nd.bounds[0], nd.bounds[1], nd.bounds[2], nd.bounds[3]);
        pairs[4 * (size_t)i + 2 * c + 1] = make_float4(nd.bounds[4], nd.bounds[5], __uint_as_float(nd.primitive_count), __uint_as_float(first));
    }
}

// slot k holds triangle idx[k] (48 B) and its vertex normals (9 floats on the host side, padded to 12 here)
__global__ void __launch_bounds__(256) lbvh_slots_kernel(const float4* __restrict__ tris, const float* __restrict__ norms9,
                                                         const uint32_t* __restrict__ idx, uint32_t n, float4* __restrict__ tri_slots,
                                                         float4* __restrict__ norm_slots) {
    const uint32_t k = blockIdx.x * 256u + threadIdx.x;
    if (k >= n) return;
    const uint32_t t = idx[k];
    for (int q = 0; q < 3; ++q) tri_slots[3ull * k + q] = tris[3ull * t + q];
    float v[12];
    for (int q = 0; q < 12; ++q) v[q] = (norms9 && q < 9) ? norms9[9ull * t + q] : 0.0f;
    for (int q = 0; q < 3; ++q) norm_slots[3ull * k + q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
}

// number of inner nodes above each leaf = the deepest traversal stack the tree can need
__global__ void __launch_bounds__(256) lbvh_depth_kernel(const uint32_t* __restrict__ par_inner, const uint32_t* __restrict__ par_leaf,
                                                         int n, unsigned int* depth_max) {
    const int k = (int)(blockIdx.x * 256u + threadIdx.x);
    unsigned int d = 0;
    if (k < n) {
        d = 1;
        for (uint32_t p = par_leaf[k] >> 1; p != 0u; p = par_inner[p] >> 1) ++d;
    }
    for (int o = 16; o; o >>= 1) d = max(d, __shfl_xor_sync(0xFFFFFFFFu, d, o));
    if ((threadIdx.x & 31u) == 0u && d) atomicMax(depth_max, d);
}

int ceres_scene_create_lbvh(const void* tris, const void* norms, size_t tri_count, int device, ceres_scene** out, double* build_ms) {
    if (!out) return fail(CERES_ERR_INVALID, "ceres_scene_create_lbvh: out is null");
    *out = nullptr;
    if (int rc = lbvh_check_args(tris, tri_count, device, "ceres_scene_create_lbvh")) return rc;
    if (tri_count == 1) {                       // a leaf root has no pair: take the general path
        RefNode root; size_t idx0 = 0;
        if (int rc = ceres_build_lbvh(tris, 1, device, &root, &idx0, build_ms)) return rc;
        return ceres_scene_create(&root, 1, &idx0, tris, norms, 1, device, out);
    }
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CERES_ERR_CUDA, "cudaSetDevice failed");
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(CERES_ERR_NO_DEVICE, std::string("device is not Blackwell-class (sm_100a required): ") + prop.name);
    const int n = (int)tri_count;
    ceres_scene* s = nullptr;
    if (int rc = scene_new(device, prop.multiProcessorCount, &s)) return rc;
    float4* d_src = nullptr; float* d_norms9 = nullptr; unsigned int* d_depth = nullptr;
    auto bail = [&](int code) { cudaFree(d_src); cudaFree(d_norms9); cudaFree(d_depth); ceres_scene_destroy(s); return code; };
#define TRY_S(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); return bail(CERES_ERR_CUDA); } } while (0)
    TRY_S(cudaMalloc((void**)&d_src, tri_count * 48));
    TRY_S(cudaMemcpy(d_src, tris, tri_count * 48, cudaMemcpyHostToDevice));
    if (norms) {
        TRY_S(cudaMalloc((void**)&d_norms9, tri_count * 36));
        TRY_S(cudaMemcpy(d_norms9, norms, tri_count * 36, cudaMemcpyHostToDevice));
    }
    float ms = 0.0f;
    size_t n_pairs = tri_count - 1;
    {
        LbvhWork w;
        if (int rc = lbvh_run(d_src, n, w, &ms)) return bail(rc);
        n_pairs = (w.node_count - 1) / 2;                    // leaf collapsing leaves fewer nodes than 2 n - 1
        const size_t pb = n_pairs * 64, tb = tri_count * 48, ib = tri_count * 4;
        TRY_S(cudaMalloc(&s->d_pairs, pb));
        TRY_S(cudaMalloc(&s->d_tris, tb));
        TRY_S(cudaMalloc(&s->d_norms, tb));
        TRY_S(cudaMalloc((void**)&d_depth, sizeof(unsigned int)));
        TRY_S(cudaMemset(d_depth, 0, sizeof(unsigned int)));
        const unsigned grid = (unsigned)((tri_count + 255) / 256);
        lbvh_pairs_kernel<<<(unsigned)((n_pairs + 255) / 256), 256>>>(w.nodes, (int)n_pairs + 1, (float4*)s->d_pairs);
        lbvh_slots_kernel<<<grid, 256>>>(d_src, d_norms9, w.idx2, (uint32_t)n, (float4*)s->d_tris, (float4*)s->d_norms);
        lbvh_depth_kernel<<<grid, 256>>>(w.pi, w.pl, n, d_depth);
        g_launches += 3;
        TRY_S(cudaGetLastError());
        unsigned int depth = 0;
        TRY_S(cudaMemcpy(&depth, d_depth, sizeof depth, cudaMemcpyDeviceToHost));
        TRY_S(cudaMemcpy(s->bounds, w.nodes[0].bounds, 24, cudaMemcpyDeviceToHost));
        s->h_slot_id.resize(tri_count);
        TRY_S(cudaMemcpy(s->h_slot_id.data(), w.idx2, ib, cudaMemcpyDeviceToHost));
        s->d_slot_id = w.idx2; w.idx2 = nullptr;            // the sorted primitive indices ARE slot -> primitive
        if (depth > 64) { fail(CERES_ERR_UNSUPPORTED, "ceres_scene_create_lbvh: tree deeper than the reference's 64-entry stack"); return bail(CERES_ERR_UNSUPPORTED); }
        s->dev.stack_depth = depth;
        s->scene_bytes = pb + 2 * tb + ib;
    }
#undef TRY_S
    cudaFree(d_src); cudaFree(d_norms9); cudaFree(d_depth);
    s->tri_count = (uint32_t)tri_count;
    s->dev.pairs = (const float4*)s->d_pairs;
    s->dev.tris = (const float4*)s->d_tris;
    s->dev.norms = (const float4*)s->d_norms;
    s->dev.slot_id = (const uint32_t*)s->d_slot_id;
    s->dev.root_leaf_count = 0;
    s->dev.n_pairs = (uint32_t)n_pairs; s->dev.n_slots = (uint32_t)tri_count;
    s->node_count = 2 * n_pairs + 1;
    s->n_pairs = (uint32_t)n_pairs;
    s->h_pair_first_child.resize(n_pairs);
    for (size_t i = 0; i < n_pairs; ++i) s->h_pair_first_child[i] = (uint32_t)(1 + 2 * i);
    if (build_ms) *build_ms = ms;
    CUDA_TRY(cudaDeviceSynchronize());          // build kernels on the legacy stream are complete before a non-blocking stream reads the scene
    *out = s;
    return CERES_OK;
}

// ---- read-bandwidth probe ("a measured L2 peak ... with an L2-resident 128-bit __ldg read kernel") ------
__global__ void __launch_bounds__(256) read_probe_kernel(const float4* __restrict__ buf, size_t n4, int repeats, float* sink) {
    float acc = 0.0f;
    const size_t stride = (size_t)gridDim.x * 256u;
    for (int r = 0; r < repeats; ++r)
        // every repeat a CTA reads the segments another CTA (on another SM) read before: hits in L2, not in its own L1
        for (size_t k = (size_t)((blockIdx.x + 37u * (unsigned)r) % gridDim.x) * 256u + threadIdx.x; k < n4; k += stride) {
            const float4 v = __ldg(buf + k);
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 123.456f) *sink = acc;        // never true: keeps the loads alive
}

int ceres_probe_read_bandwidth(int device, size_t working_set_bytes, int repeats, double* gb_per_s) {
    if (!gb_per_s || working_set_bytes < 4096 || repeats < 1) return fail(CERES_ERR_INVALID, "ceres_probe_read_bandwidth: bad argument");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return fail(CERES_ERR_NO_DEVICE, "no CUDA device");
    if (device < 0 || device >= ndev) return fail(CERES_ERR_INVALID, "bad device index");
    DeviceGuard guard(device);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    const size_t n4 = working_set_bytes / 16;
    float4* buf = nullptr; float* sink = nullptr; cudaEvent_t e0 = nullptr, e1 = nullptr;
    auto cleanup = [&]() { cudaFree(buf); cudaFree(sink); if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); };
#define TRY_P(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); cleanup(); return CERES_ERR_CUDA; } } while (0)
    TRY_P(cudaMalloc((void**)&buf, n4 * 16));
    TRY_P(cudaMalloc((void**)&sink, 4));
    TRY_P(cudaMemset(buf, 0, n4 * 16));
    TRY_P(cudaEventCreate(&e0)); TRY_P(cudaEventCreate(&e1));
    const unsigned grid = (unsigned)prop.multiProcessorCount * 8u;
    read_probe_kernel<<<grid, 256>>>(buf, n4, 2, sink);                  // warm the caches
    double best = 0.0;
    for (int t = 0; t < 5; ++t) {
        TRY_P(cudaEventRecord(e0, 0));
        read_probe_kernel<<<grid, 256>>>(buf, n4, repeats, sink);
        TRY_P(cudaEventRecord(e1, 0));
        TRY_P(cudaEventSynchronize(e1));
        float ms = 0.0f;
        TRY_P(cudaEventElapsedTime(&ms, e0, e1));
        if (ms > 0.0f) best = std::max(best, (double)n4 * 16.0 * repeats / (ms * 1e-3) / 1e9);
    }
    g_launches += 6;
#undef TRY_P
    *gb_per_s = best;
    cleanup();
    return CERES_OK;
}

int ceres_scene_last_rectangle(const ceres_scene* s, uint32_t rect[4]) {
    if (!s || !rect) return fail(CERES_ERR_INVALID, "ceres_scene_last_rectangle: null argument");
    for (int k = 0; k < 4; ++k) rect[k] = s->last_rect[k];
    return CERES_OK;
}

// Streaming-store bandwidth of the host-fill pool into page-locked memory: the denominator of the host-frame floor
// (every byte of a host frame is written into host DRAM by somebody: DMA for the rectangle, this pool for the zeros).
int ceres_probe_host_fill_bandwidth(size_t bytes, int repeats, double* gb_per_s) {
    if (!gb_per_s || bytes < (1u << 20) || repeats < 1) return fail(CERES_ERR_INVALID, "ceres_probe_host_fill_bandwidth: bad argument");
    void* buf = nullptr;
    CUDA_TRY(cudaMallocHost(&buf, bytes));
    ceres_host::Pool& pool = ceres_host::Pool::instance();
    double best = 0.0;
    for (int r = 0; r < repeats + 1; ++r) {
        ceres_host::Job job;
        job.add(buf, bytes, 0x00);
        const auto t0 = std::chrono::steady_clock::now();
        pool.submit(&job);
        pool.wait(&job);
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (r > 0 && sec > 0.0) best = std::max(best, (double)bytes / sec / 1e9);       // r = 0 touches the pages
    }
    cudaFreeHost(buf);
    *gb_per_s = best;
    return CERES_OK;
}

int ceres_scene_set_tuning(ceres_scene* s, const ceres_tuning* t) {
    if (!s) return fail(CERES_ERR_INVALID, "null scene");
    if (t) s->tune = *t; else s->tune = ceres_tuning{};
    s->shadow_blocks_per_sm = s->shadow_blocks_per_sm_stats = 0;
    for (int k = 0; k < 10; ++k) s->trace_blocks_per_sm[k] = 0;
    return CERES_OK;
}

int ceres_scene_get_step_histogram(ceres_scene* s, uint64_t hist[128]) {
    if (!s || !hist) return fail(CERES_ERR_INVALID, "null argument");
    if (!s->d_hist) return fail(CERES_ERR_INVALID, "ceres_scene_get_step_histogram: statistics mode was never enabled");
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaMemcpy(hist, s->d_hist, 128 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemset(s->d_hist, 0, 128 * sizeof(uint64_t)));
    CUDA_TRY(cudaDeviceSynchronize());
    return CERES_OK;
}

int ceres_scene_set_timeline(ceres_scene* s, int enable) {
    if (!s) return fail(CERES_ERR_INVALID, "null scene");
    DeviceGuard guard(s->device);
    if (enable && !s->d_timeline) {
        s->timeline_warps = (size_t)s->sm_count * 16 * (FRAME_THREADS / 32);
        CUDA_TRY(cudaMalloc((void**)&s->d_timeline, s->timeline_warps * CERES_TIMELINE_WORDS * sizeof(unsigned long long)));
    }
    CUDA_TRY(cudaDeviceSynchronize());
    s->dev.timeline = enable ? s->d_timeline : nullptr;
    return CERES_OK;
}

int ceres_scene_get_timeline(ceres_scene* s, uint64_t* out, size_t cap_warps, size_t* warps) {
    if (!s || !out || !warps) return fail(CERES_ERR_INVALID, "null argument");
    if (!s->d_timeline) return fail(CERES_ERR_INVALID, "ceres_scene_get_timeline: ceres_scene_set_timeline was never enabled");
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaDeviceSynchronize());
    const size_t n = std::min(cap_warps, s->timeline_warps);
    CUDA_TRY(cudaMemcpy(out, s->d_timeline, n * CERES_TIMELINE_WORDS * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    *warps = n;
    return CERES_OK;
}

int ceres_scene_frame_error(ceres_scene* s, uint64_t* bits) {
    if (!s || !bits) return fail(CERES_ERR_INVALID, "null argument");
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaDeviceSynchronize());
    unsigned long long w = 0;
    CUDA_TRY(cudaMemcpy(&w, s->d_out + 10, sizeof w, cudaMemcpyDeviceToHost));
#ifdef CERES_CHECKED
    // protocol invariants between frames: ready[] all-zero, every counter slot re-armed (frame_fused.cuh)
    if (s->d_ready && s->ready_cap) {
        std::vector<uint32_t> r(s->ready_cap);
        CUDA_TRY(cudaMemcpy(r.data(), s->d_ready, r.size() * 4, cudaMemcpyDeviceToHost));
        for (uint32_t x : r) if (x) { w |= 0x10000ull; break; }
    }
    {
        FrameCounters c[CERES_MAX_SUB];
        CUDA_TRY(cudaMemcpy(c, s->d_ctr, sizeof c, cudaMemcpyDeviceToHost));
        for (const auto& k : c)
            if (k.qcount || k.shadow_next || k.occluded || k.done_blocks || k.primary_next || k.primary_done || k.error || k.ext[0] != 0xFFFFFFFFu || k.ext[1] || k.ext[2] != 0xFFFFFFFFu || k.ext[3])
                w |= 0x20000ull;
    }
#endif
    if (w) { CUDA_TRY(cudaMemset(s->d_out + 10, 0, sizeof w)); CUDA_TRY(cudaDeviceSynchronize()); }
    *bits = w;
    return CERES_OK;
}

int ceres_build_is_checked(void) {
#ifdef CERES_CHECKED
    return 1;
#else
    return 0;
#endif
}

int ceres_scene_set_stats(ceres_scene* s, int enable) {
    if (!s) return fail(CERES_ERR_INVALID, "null scene");
    if (enable && !s->d_hist) {
        DeviceGuard guard(s->device);
        CUDA_TRY(cudaMalloc((void**)&s->d_hist, 128 * sizeof(unsigned long long)));
        CUDA_TRY(cudaMemset(s->d_hist, 0, 128 * sizeof(unsigned long long)));
        CUDA_TRY(cudaDeviceSynchronize());
        s->dev.hist = s->d_hist;
    }
    s->stats = enable != 0;
    return CERES_OK;
}

int ceres_scene_set_profiling(ceres_scene* s, int enable) {
    if (!s) return fail(CERES_ERR_INVALID, "null scene");
    s->profile = enable != 0;
    s->prof_frames = 0;
    return CERES_OK;
}

int ceres_scene_get_kernel_ms(ceres_scene* s, double ms[4], uint64_t* frames) {
    if (!s || !ms) return fail(CERES_ERR_INVALID, "null argument");
    DeviceGuard guard(s->device);
    for (int k = 0; k < 4; ++k) ms[k] = 0.0;
    for (size_t fidx = 0; fidx < s->prof_frames; ++fidx) {
        cudaEvent_t* e = &s->prof_events[4 * fidx];
        CUDA_TRY(cudaEventSynchronize(e[3]));
        float a = 0, b = 0, c = 0, t = 0;
        CUDA_TRY(cudaEventElapsedTime(&a, e[0], e[1]));
        CUDA_TRY(cudaEventElapsedTime(&b, e[1], e[2]));
        CUDA_TRY(cudaEventElapsedTime(&c, e[2], e[3]));
        CUDA_TRY(cudaEventElapsedTime(&t, e[0], e[3]));
        if (a + b < 0.002f) c = 0.0f;      // frame rendered as concurrent sub-frames: only the total is meaningful
        ms[0] += a; ms[1] += b; ms[2] += c; ms[3] += t;
    }
    if (frames) *frames = s->prof_frames;
    return CERES_OK;
}

int ceres_scene_get_stats(ceres_scene* s, uint64_t stats[8]) {
    if (!s || !stats) return fail(CERES_ERR_INVALID, "null argument");
    DeviceGuard guard(s->device);
    unsigned long long tmp[10];
    CUDA_TRY(cudaMemcpy(tmp, s->d_out, sizeof tmp, cudaMemcpyDeviceToHost));
    for (int k = 0; k < 8; ++k) stats[k] = tmp[2 + k];
    stats[6] += s->rays_outside;
    return CERES_OK;
}

// host copy of frame_row()
static uint32_t frame_row_host(const Frame& f, uint32_t lrow) {
    if (f.shard_count <= 1) return lrow;
    const uint32_t blk = lrow / f.block_rows;
    return (blk * f.shard_count + f.shard_index) * f.block_rows + (lrow - blk * f.block_rows);
}

// Conservative screen rectangle of the scene: pixels outside it cannot hit anything, so the primary wave never traces
// them (their colour is the 0 the clear pass wrote, render.hpp:116-117).  The root box is inflated by 1e-4 of the
// scene/eye scale -- three orders of magnitude above the rounding error of the binary32 slab test and of the
// normalised ray direction -- its 8 corners are projected in double precision, and the rectangle gets one more pixel
// of margin.  Any corner at or behind the eye plane, or any non-finite value, disables the culling.
static void cull_rectangle(const float bounds[6], Frame& f, uint32_t rows_local) {
    f.cull_i0 = 0; f.cull_i1 = f.W; f.cull_l0 = 0; f.cull_l1 = rows_local;
    double scale = 0.0;
    for (int k = 0; k < 6; ++k) scale = std::max(scale, std::fabs((double)bounds[k]));
    for (int k = 0; k < 3; ++k) scale = std::max(scale, std::fabs((double)f.eye[k]));
    if (!std::isfinite(scale) || !(scale > 0.0)) return;
    const double m = 1e-4 * scale;
    auto dot = [](const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
    double iu[3], iv[3], dir[3];
    for (int k = 0; k < 3; ++k) { iu[k] = f.iu[k]; iv[k] = f.iv[k]; dir[k] = f.dir[k]; }
    const double iu2 = dot(iu, iu), iv2 = dot(iv, iv), d2 = dot(dir, dir);
    if (!(iu2 > 0.0) || !(iv2 > 0.0) || !(d2 > 0.0) || !std::isfinite(iu2 + iv2 + d2)) return;
    double umin = INFINITY, umax = -INFINITY, vmin = INFINITY, vmax = -INFINITY;
    for (int c = 0; c < 8; ++c) {
        double p[3];
        for (int k = 0; k < 3; ++k) p[k] = ((c >> k) & 1 ? (double)bounds[2 * k + 1] + m : (double)bounds[2 * k] - m) - (double)f.eye[k];
        const double z = dot(p, dir) / d2;
        if (!(z > 1e-6 * scale)) return;
        const double u = dot(p, iu) / iu2 / z, v = dot(p, iv) / iv2 / z;
        umin = std::min(umin, u); umax = std::max(umax, u); vmin = std::min(vmin, v); vmax = std::max(vmax, v);
    }
    if (!std::isfinite(umin + umax + vmin + vmax)) return;
    // pixel centre i sits at u = 2 (i + 0.5) / W - 1  (render.hpp:109)
    auto lo_px = [](double u, double n) { double x = std::floor((u + 1.0) * n / 2.0 - 0.5) - 1.0; return x < 0.0 ? 0.0 : (x > n ? n : x); };
    auto hi_px = [](double u, double n) { double x = std::ceil((u + 1.0) * n / 2.0 - 0.5) + 2.0; return x < 0.0 ? 0.0 : (x > n ? n : x); };
    const uint32_t i0 = (uint32_t)lo_px(umin, f.W), i1 = (uint32_t)hi_px(umax, f.W);
    const uint32_t j0 = (uint32_t)lo_px(vmin, f.H), j1 = (uint32_t)hi_px(vmax, f.H);
    f.cull_i0 = i0; f.cull_i1 = i1 > i0 ? i1 : i0;
    // frame rows [j0, j1) -> local rows of this shard (frame_row is increasing in the local row)
    uint32_t l0 = 0;
    while (l0 < rows_local && frame_row_host(f, l0) < j0) ++l0;
    uint32_t l1 = l0;
    while (l1 < rows_local && frame_row_host(f, l1) < j1) ++l1;
    f.cull_l0 = l0; f.cull_l1 = l1;
}

// Fills the per-frame constants (everything but the row range).
static int prepare_frame(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height,
                         unsigned flags, const ceres_shard* shard, float* d_pixels, uint32_t* d_ids, uint64_t* d_counters,
                         Frame& f, uint32_t& rows_local) {
    if (!s || !cam || !sun || !d_pixels) return fail(CERES_ERR_INVALID, "ceres_render: null argument");
    if (!width || !height || width > 65535u * 32u || (uint64_t)width * height >= (1ull << 32))
        return fail(CERES_ERR_INVALID, "ceres_render: bad image size");
    if (shard && shard->count > 1 && (shard->block_rows == 0 || shard->index >= shard->count))
        return fail(CERES_ERR_INVALID, "ceres_render: bad shard");
    f = Frame{};
    for (int k = 0; k < 3; ++k) { f.eye[k] = cam->eye[k]; f.sun[k] = sun[k]; }
    frame_setup_host(cam, width, height, f.dir, f.iu, f.iv);
    f.W = (uint32_t)width; f.H = (uint32_t)height;
    rows_local = shard_rows(height, shard);
    f.rows_local = rows_local;
    const bool sharded = shard && shard->count > 1;
    f.block_rows = sharded ? shard->block_rows : (uint32_t)height;
    f.shard_index = sharded ? shard->index : 0;
    f.shard_count = sharded ? shard->count : 1;
    f.compact = (flags & CERES_COMPACT_ROWS) ? 1u : 0u;
    f.shadows = (flags & CERES_NO_SHADOWS) ? 0u : 1u;
    f.pixels = d_pixels; f.ids = d_ids;
    f.out_counters = d_counters ? (unsigned long long*)d_counters : s->d_out;
    f.out_stats = s->stats ? s->d_out + 2 : nullptr;
    f.ctr = s->d_ctr;
    if (s->tune.variant != 2) cull_rectangle(s->bounds, f, rows_local);
    else { f.cull_i0 = 0; f.cull_i1 = f.W; f.cull_l0 = 0; f.cull_l1 = rows_local; }
    s->last_rect[0] = f.cull_i0; s->last_rect[1] = f.cull_i1; s->last_rect[2] = f.cull_l0; s->last_rect[3] = f.cull_l1;
    return CERES_OK;
}

// What has to be zeroed around the rectangle the frame kernel writes.
enum { CLEAR_NONE = 0,       // nothing: the caller never reads outside the rectangle (host-frame staging), or clears itself
       CLEAR_OUTSIDE = 1 };  // every pixel of the launch's rows outside the rectangle (caller-owned device frames)

// ready[] entries (frame_fused.cuh) a launch over `records` queue slots can touch: its chunks + one terminating entry per warp
static size_t ready_words(const ceres_scene* s, size_t records) { return (records >> READY_MIN_SHIFT) + 2 + (size_t)s->sm_count * 16 * (FRAME_THREADS / 32); }

#ifdef CERES_EXPERIMENTS
static int launch_rows_waves(ceres_scene* s, Frame& f, uint32_t row_begin, uint32_t row_end, cudaStream_t st, cudaEvent_t* pe);
#endif

// Enqueues the kernels for local rows [row_begin, row_end) of the frame on `st`.
static int launch_rows(ceres_scene* s, Frame f, uint32_t row_begin, uint32_t row_end, cudaStream_t st, int sub, int clear_mode) {
    if (row_end <= row_begin) return CERES_OK;
    f.row_begin = row_begin; f.row_end = row_end; f.accumulate = 1u;
    // every launch owns the queue region of its rows and one FrameCounters slot, so that bands of one frame can run
    // concurrently on different streams (the caller has sized the queue for the whole shard)
    if ((size_t)row_end * f.W > s->queue_cap) return fail(CERES_ERR_INVALID, "internal: queue not sized for the frame");
    f.queue = s->d_queue + (size_t)row_begin * f.W;
    f.ready = s->d_ready + (((size_t)row_begin * f.W) >> READY_MIN_SHIFT) + (size_t)sub * ready_words(s, 0);
    f.ctr = s->d_ctr + sub;
    f.out_error = s->d_out + 10;
    f.queue_cap = (row_end - row_begin) * f.W;
    f.ready_cap = (uint32_t)ready_words(s, f.queue_cap);
    f.pix_rows = f.compact ? f.rows_local : f.H;
    if (f.host_ext) f.host_ext += 8 * sub;

    cudaEvent_t* pe = nullptr;
    if (s->profile && s->profile_kernels) {
        if (s->prof_events.size() < 4 * (s->prof_frames + 1)) {
            for (int k = 0; k < 4; ++k) { cudaEvent_t e; CUDA_TRY(cudaEventCreate(&e)); s->prof_events.push_back(e); }
        }
        pe = &s->prof_events[4 * s->prof_frames++];
        CUDA_TRY(cudaEventRecord(pe[0], st));
    }
    // the conservative screen rectangle of the scene within this launch's rows, cut into 8x4-pixel tiles
    {
        const uint32_t r0 = std::max(row_begin, f.cull_l0), r1 = std::min(row_end, f.cull_l1);
        f.rect_x0 = f.cull_i0; f.rect_x1 = f.cull_i1; f.rect_row0 = r0; f.rect_row1 = r1 > r0 ? r1 : r0;
        f.tiles_x = (f.rect_x1 - f.rect_x0 + 7u) / 8u;
        f.tiles_y = (f.rect_row1 - f.rect_row0 + 3u) / 4u;
        const uint64_t tiles = (uint64_t)f.tiles_x * f.tiles_y;
        if (tiles * 32u >= (1ull << 32)) return fail(CERES_ERR_INVALID, "ceres_render: image too large");
        f.work_items = (uint32_t)(tiles * 32u);
    }
    if (s->tune.variant != 0) {
#ifdef CERES_EXPERIMENTS
        // (the wave pipelines always clear their own rows: their misses store nothing)
        return launch_rows_waves(s, f, row_begin, row_end, st, pe);
#else
        return fail(CERES_ERR_UNSUPPORTED, "ceres_tuning.variant != 0 needs the experiments build (libceres_b200_exp.so, -DCERES_EXPERIMENTS)");
#endif
    }
    f.clear_rows = clear_mode == CLEAR_OUTSIDE ? 1u : 0u;        // done by the frame kernel's own warps (phase 0)
    if (pe) CUDA_TRY(cudaEventRecord(pe[1], st));
    {
        const size_t smem = (size_t)s->dev.stack_depth * 4 * FRAME_THREADS;          // <= 32 KB: depth <= 64
        const int order = s->tune.shadow_order == 1 ? ORDER_NEAR : s->tune.shadow_order == 2 ? ORDER_LEFT : FRAME_DEFAULT_SHADOW_ORDER;
        auto go = [&](auto kernel, int slot) -> int {
            int& bps = s->trace_blocks_per_sm[slot];
            if (bps == 0) {
                if (s->tune.ctas_per_sm) bps = (int)std::min(16u, s->tune.ctas_per_sm);
                else CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, FRAME_THREADS, smem));
                if (bps < 1) bps = 1;
                if (bps > 16) bps = 16;             // ready_words() reserves one terminating entry per warp of 16 CTAs per SM
            }
            kernel<<<(unsigned)(s->sm_count * bps), FRAME_THREADS, smem, st>>>(s->dev, f);
            g_launches++;
            CUDA_TRY(cudaGetLastError());
            return CERES_OK;
        };
        // reserved[0]: 0 default, 1 off, 2 on; reserved[2]: period | min_idle << 8 (0 = defaults).  Default: on when the launch is
        // small enough for its tail to matter -- fewer than FRAME_SHARE_AUTO_PX rectangle pixels per resident warp (a rank's share
        // of a frame at N >= 4, small images); a full 4K frame on one GPU is throughput-bound and runs 2 % faster without it
        // (profiles/r2/sweeps.md).  Row bands of one frame overlap each other's tails instead.
        const uint64_t warps_guess = (uint64_t)s->sm_count * 8u * (FRAME_THREADS / 32);
        const bool share_auto = s->cur_nsub <= 1u && f.work_items / warps_guess < FRAME_SHARE_AUTO_PX;
        const bool share = s->tune.reserved[0] ? s->tune.reserved[0] == 2u : share_auto;
        f.share_period = (s->tune.reserved[2] & 0xFFu) ? (s->tune.reserved[2] & 0xFFu) : FRAME_SHARE_PERIOD;
        f.share_min_idle = ((s->tune.reserved[2] >> 8) & 0xFFu) ? ((s->tune.reserved[2] >> 8) & 0xFFu) : FRAME_SHARE_MIN_IDLE;
        if (s->dev.timeline) CUDA_TRY(cudaMemsetAsync(s->d_timeline, 0, s->timeline_warps * CERES_TIMELINE_WORDS * sizeof(unsigned long long), st));
        {   // records per shadow work item: tune.shadow_block = 8 / 16 / 32, 0 = default
            const uint32_t want = s->tune.shadow_block ? s->tune.shadow_block : FRAME_DEFAULT_CHUNK;
            f.chunk_shift = want <= 8u ? 3u : want <= 16u ? 4u : 5u;
        }
        int rc;
        const int near = order == ORDER_NEAR;
        if (s->dev.timeline && !s->stats) {      // the timeline build: default child order only
            rc = share ? go(frame_kernel<false, FRAME_DEFAULT_SHADOW_ORDER, true, true>, 8) : go(frame_kernel<false, FRAME_DEFAULT_SHADOW_ORDER, false, true>, 9);
        } else if (s->stats) {
            rc = share ? (near ? go(frame_kernel<true, ORDER_NEAR, true, false>, 1) : go(frame_kernel<true, ORDER_LEFT, true, false>, 3))
                       : (near ? go(frame_kernel<true, ORDER_NEAR, false, false>, 5) : go(frame_kernel<true, ORDER_LEFT, false, false>, 7));
        } else {
            rc = share ? (near ? go(frame_kernel<false, ORDER_NEAR, true, false>, 0) : go(frame_kernel<false, ORDER_LEFT, true, false>, 2))
                       : (near ? go(frame_kernel<false, ORDER_NEAR, false, false>, 4) : go(frame_kernel<false, ORDER_LEFT, false, false>, 6));
        }
        if (rc) return rc;
    }
    if (pe) { CUDA_TRY(cudaEventRecord(pe[2], st)); CUDA_TRY(cudaEventRecord(pe[3], st)); }
    return CERES_OK;
}

#ifdef CERES_EXPERIMENTS
// The wave pipelines of round 1 (tuning variants 1-6): clear -> primary wave -> shadow wave -> shade pass.
static int launch_rows_waves(ceres_scene* s, Frame& f, uint32_t row_begin, uint32_t row_end, cudaStream_t st, cudaEvent_t* pe) {
    const uint32_t rows = row_end - row_begin;
    const size_t smem_p = (size_t)s->dev.stack_depth * 4 * PRIMARY_THREADS;
    const size_t smem_s = (size_t)s->dev.stack_depth * 4 * SHADOW_THREADS;
    dim3 grid((unsigned)((f.W + 31) / 32), (unsigned)((rows + 7) / 8));
    if (grid.y > 65535u) return fail(CERES_ERR_INVALID, "ceres_render: image too tall");
    if (s->tune.variant == 1) {
        // legacy wave kernels: one thread per pixel, no culling, no refill (kept for A/B measurements)
        if (int rc = smem_opt_in(primary_kernel<true>, smem_p)) return rc;
        if (int rc = smem_opt_in(primary_kernel<false>, smem_p)) return rc;
        if (int rc = smem_opt_in(shadow_kernel<true>, smem_s)) return rc;
        if (int rc = smem_opt_in(shadow_kernel<false>, smem_s)) return rc;
        if (s->stats) primary_kernel<true><<<grid, PRIMARY_THREADS, smem_p, st>>>(s->dev, f);
        else primary_kernel<false><<<grid, PRIMARY_THREADS, smem_p, st>>>(s->dev, f);
        g_launches++;
        if (pe) CUDA_TRY(cudaEventRecord(pe[1], st));
        if (f.shadows) {
            int& bps = s->stats ? s->shadow_blocks_per_sm_stats : s->shadow_blocks_per_sm;
            if (bps == 0) {
                if (s->tune.ctas_per_sm) bps = (int)s->tune.ctas_per_sm;
                else if (s->stats) CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, shadow_kernel<true>, SHADOW_THREADS, smem_s));
                else CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, shadow_kernel<false>, SHADOW_THREADS, smem_s));
                if (bps < 1) bps = 1;
            }
            const unsigned g = (unsigned)(s->sm_count * bps);
            if (s->stats) shadow_kernel<true><<<g, SHADOW_THREADS, smem_s, st>>>(s->dev, f);
            else shadow_kernel<false><<<g, SHADOW_THREADS, smem_s, st>>>(s->dev, f);
            g_launches++;
        }
    } else {
        // clear -> persistent primary wave over the culled rectangle -> persistent shadow wave
        const uint32_t r0 = std::max(row_begin, f.cull_l0), r1 = std::min(row_end, f.cull_l1);
        f.rect_x0 = f.cull_i0; f.rect_x1 = f.cull_i1; f.rect_row0 = r0; f.rect_row1 = r1 > r0 ? r1 : r0;
        f.tiles_x = (f.rect_x1 - f.rect_x0 + 7u) / 8u;
        f.tiles_y = (f.rect_row1 - f.rect_row0 + 3u) / 4u;
        const uint64_t tiles = (uint64_t)f.tiles_x * f.tiles_y;
        if (tiles * 32u >= (1ull << 32)) return fail(CERES_ERR_INVALID, "ceres_render: image too large");
        f.work_items = (uint32_t)(tiles * 32u);
        f.refill_min = s->tune.reserved[0] ? std::min(32u, s->tune.reserved[0]) : REFILL_MIN;
        f.chunk = s->tune.reserved[2] ? (s->tune.reserved[2] + 31u) / 32u * 32u : TRACE_CHUNK;
        if (!f.skip_clear) {
            clear_kernel<<<(unsigned)(s->sm_count * 8), 256, 0, st>>>(f);
            g_launches++;
        }
        // launch helper: persistent grid = SMs x resident CTAs of this instantiation
        auto launch_wave = [&](int wave, auto kernel, int threads) -> int {
            const size_t smem_t = (size_t)s->dev.stack_depth * 4 * threads;
            if (int rc = smem_opt_in(kernel, smem_t)) return rc;
            int& bps = s->trace_blocks_per_sm[wave * 2 + (s->stats ? 1 : 0)];
            if (bps == 0) {
                if (s->tune.ctas_per_sm) bps = (int)s->tune.ctas_per_sm;
                else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, threads, smem_t) != cudaSuccess) bps = 1;
                if (bps < 1) bps = 1;
            }
            kernel<<<(unsigned)(s->sm_count * bps), threads, smem_t, st>>>(s->dev, f);
            g_launches++;
            return 0;
        };
        const bool wide_p = s->tune.primary_block == 256, wide_s = s->tune.shadow_block == 256;
        if (s->tune.variant == 4) {
            // CTA-level repacking of the surviving rays every K steps (trace_repack.cuh)
            auto launch_repack = [&](int wave, auto kernel, int words) {
                const size_t smem_t = ((size_t)s->dev.stack_depth + words) * 4 * 128;
                int& bps = s->trace_blocks_per_sm[wave * 2 + (s->stats ? 1 : 0)];
                if (bps == 0) {
                    if (s->tune.ctas_per_sm) bps = (int)s->tune.ctas_per_sm;
                    else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 128, smem_t) != cudaSuccess) bps = 1;
                    if (bps < 1) bps = 1;
                }
                kernel<<<(unsigned)(s->sm_count * bps), 128, smem_t, st>>>(s->dev, f);
                g_launches++;
            };
            if (s->stats) launch_repack(0, trace_repack_kernel<WAVE_PRIMARY, true, 128>, RepackWords<WAVE_PRIMARY, true>::N);
            else launch_repack(0, trace_repack_kernel<WAVE_PRIMARY, false, 128>, RepackWords<WAVE_PRIMARY, false>::N);
            if (pe) CUDA_TRY(cudaEventRecord(pe[1], st));
            if (f.shadows) {
                if (s->stats) launch_repack(1, trace_repack_kernel<WAVE_SHADOW, true, 128>, RepackWords<WAVE_SHADOW, true>::N);
                else launch_repack(1, trace_repack_kernel<WAVE_SHADOW, false, 128>, RepackWords<WAVE_SHADOW, false>::N);
            }
        } else if (s->tune.variant == 3) {
            // phase-sorted execution (trace_batched.cuh)
            if (s->stats) launch_wave(0, trace_batched_kernel<WAVE_PRIMARY, true, 128>, 128);
            else launch_wave(0, trace_batched_kernel<WAVE_PRIMARY, false, 128>, 128);
            if (pe) CUDA_TRY(cudaEventRecord(pe[1], st));
            if (f.shadows) {
                if (s->stats) launch_wave(1, trace_batched_kernel<WAVE_SHADOW, true, 128>, 128);
                else launch_wave(1, trace_batched_kernel<WAVE_SHADOW, false, 128>, 128);
            }
        } else {
        if (s->stats) { if (wide_p) launch_wave(0, trace_kernel<WAVE_PRIMARY, true, 256>, 256); else launch_wave(0, trace_kernel<WAVE_PRIMARY, true, 128>, 128); }
        else { if (wide_p) launch_wave(0, trace_kernel<WAVE_PRIMARY, false, 256>, 256); else launch_wave(0, trace_kernel<WAVE_PRIMARY, false, 128>, 128); }
        if (pe) CUDA_TRY(cudaEventRecord(pe[1], st));
        if (f.shadows && s->tune.variant == 5) {
            // shadow wave with deferred triangle tests (trace_shadow_spec.cuh); ctas_per_sm picks the register budget
            const size_t smem_t = ((size_t)s->dev.stack_depth + 2 * SPEC_PMAX) * 4 * 128;
            const unsigned want = s->tune.ctas_per_sm ? s->tune.ctas_per_sm : 8u;
            auto go = [&](auto kernel) {
                int bps = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 128, smem_t) != cudaSuccess || bps < 1) bps = 1;
                kernel<<<(unsigned)(s->sm_count * bps), 128, smem_t, st>>>(s->dev, f);
                g_launches++;
            };
            if (s->stats) go(trace_shadow_spec_kernel<true, 128, 8>);
            else if (want >= 12) go(trace_shadow_spec_kernel<false, 128, 12>);
            else if (want >= 10) go(trace_shadow_spec_kernel<false, 128, 10>);
            else go(trace_shadow_spec_kernel<false, 128, 8>);
        } else if (f.shadows) {
            if (s->stats) { if (wide_s) launch_wave(1, trace_kernel<WAVE_SHADOW, true, 256>, 256); else launch_wave(1, trace_kernel<WAVE_SHADOW, true, 128>, 128); }
            else { if (wide_s) launch_wave(1, trace_kernel<WAVE_SHADOW, false, 256>, 256); else launch_wave(1, trace_kernel<WAVE_SHADOW, false, 128>, 128); }
        }
        }
    }
    CUDA_TRY(cudaGetLastError());            // a failed wave launch must not be followed by the shade pass (it re-arms the counters)
    if (pe) CUDA_TRY(cudaEventRecord(pe[2], st));
    shade_kernel<<<(unsigned)(s->sm_count * 4), SHADE_THREADS, 0, st>>>(s->dev, f);
    g_launches++;
    if (pe) CUDA_TRY(cudaEventRecord(pe[3], st));
    CUDA_TRY(cudaGetLastError());
    return CERES_OK;
}
#endif  // CERES_EXPERIMENTS

}  // extern "C" (a C++ template follows)

// defaults from the sweeps in profiles/r1/ and profiles/r2/sweeps.md: device-resident frames stay serial; host frames want
// bands so that the first copy starts early, but every band is a persistent launch of its own with its own tail (a band of
// 1/8 of the rectangle is as latency-bound as one rank of eight): 3-4 bands are the optimum with the frame kernel
// (0.78 ms per 4K float frame against 0.86 ms with 8 bands and 0.98 ms with one)
#define CERES_SUBFRAMES_DEVICE 1u
#define CERES_SUBFRAMES_HOST 4u

// Renders local rows [a, b) of the frame, forked from and joined back to stream `st`.
// The rows are cut into up to CERES_MAX_SUB row bands ("sub-frames"), each enqueued on its own stream with its own
// queue region and counters.  Why: every wave kernel ends with a latency-bound tail (the last warps walk 100+ dependent
// node steps while the rest of the GPU idles, ~30-50 us per wave on the dragon); with concurrent bands the tail of one
// band's wave is filled by the next band's bulk, and in the host path a finished band is copied while others render.
// hook(band, r0, r1, stream) may enqueue follow-up work (the D2H copies) behind a band's kernels.
// After a host-path frame has been synchronised: h_out[10] holds the FRAME_ERR_* bits of its launches.  On error the
// queue bookkeeping may be inconsistent: re-arm it so that the next frame starts clean.
static int frame_error_check(ceres_scene* s) {
    if (!s->h_out[10]) return CERES_OK;
    const unsigned long long bits = s->h_out[10];
    s->h_out[10] = 0;
    cudaDeviceSynchronize();
    cudaMemset(s->d_out + 10, 0, sizeof(unsigned long long));
    if (s->d_ready) cudaMemset(s->d_ready, 0, s->ready_cap * sizeof(uint32_t));
    FrameCounters init[CERES_MAX_SUB] = {};
    for (auto& c : init) { c.ext[0] = 0xFFFFFFFFu; c.ext[2] = 0xFFFFFFFFu; }
    cudaMemcpy(s->d_ctr, init, sizeof init, cudaMemcpyHostToDevice);
    cudaDeviceSynchronize();
    return fail(CERES_ERR_CUDA, std::string("frame kernel: ") + ((bits & 1u) ? "a bounded device-side wait expired" : "an index left its bound (checked build)") +
                                " (error bits " + std::to_string(bits) + ", include/ceres_b200.h: ceres_scene_frame_error)");
}

// Sizes the record queue and its ready[] array for `records` queue slots.
static int size_queue(ceres_scene* s, size_t records) {
    const size_t need_ready = (records >> READY_MIN_SHIFT) + (size_t)CERES_MAX_SUB * ready_words(s, 0) + 2;
    if (records <= s->queue_cap && need_ready <= s->ready_cap) return CERES_OK;
    CUDA_TRY(cudaDeviceSynchronize());      // growing the queue frees the old one: wait for frames still using it
    if (int rc = ensure(s->d_queue, s->queue_cap, records)) return rc;
    if (need_ready > s->ready_cap) {
        if (int rc = ensure(s->d_ready, s->ready_cap, need_ready)) return rc;
        CUDA_TRY(cudaMemset(s->d_ready, 0, s->ready_cap * sizeof(uint32_t)));      // the frame kernel keeps it all-zero between launches
        // the memset runs on the legacy stream, the frames on NON-BLOCKING streams, which it does not order: without this wait a
        // frame kernel could start first and have its ready[] credits wiped under it (seen once as an expired queue wait)
        CUDA_TRY(cudaDeviceSynchronize());
    }
    return CERES_OK;
}

// How local rows [a, b) of a frame are cut into row bands: nsub launches of band_rows rows (the last one may be shorter).
static void band_split(const ceres_scene* s, const Frame& f, uint32_t a, uint32_t b, uint32_t default_sub, uint32_t& nsub, uint32_t& band_rows) {
    const uint32_t rows = b - a;
    nsub = s->tune.reserved[1] ? s->tune.reserved[1] : default_sub;
    if (f.peer) nsub = 1;      // one done flag per rank and frame
    // a band must be worth its launch: at least 256 Ki pixels of the rectangle each
    const uint64_t rect_px = (uint64_t)rows * (f.cull_i1 > f.cull_i0 ? f.cull_i1 - f.cull_i0 : 1u);
    nsub = std::min<uint32_t>(std::min<uint32_t>(nsub, CERES_MAX_SUB), (uint32_t)std::max<uint64_t>(1u, rect_px >> 18));
    if (nsub < 1) nsub = 1;
    band_rows = ((rows + nsub - 1) / nsub + 3u) / 4u * 4u;
}

template <class Hook>
static int render_rows(ceres_scene* s, const Frame& f, uint32_t a, uint32_t b, cudaStream_t st, uint32_t default_sub, int clear_mode, Hook&& hook) {
    if (b <= a) return CERES_OK;
    if (int rc = size_queue(s, (size_t)b * f.W)) return rc;
    // the frames of one scene share the record queue, ready[] and the counter slots: a frame enqueued on another stream
    // than the previous one waits for it
    if (s->last_stream && s->last_stream != st) CUDA_TRY(cudaStreamWaitEvent(st, s->frame_done, 0));
    struct Done { ceres_scene* s; cudaStream_t st; ~Done() { if (cudaEventRecord(s->frame_done, st) == cudaSuccess) s->last_stream = st; else cudaGetLastError(); } } done{s, st};
    CUDA_TRY(cudaMemsetAsync(f.out_counters, 0, 2 * sizeof(unsigned long long), st));
    if (f.out_stats) CUDA_TRY(cudaMemsetAsync(f.out_stats, 0, 8 * sizeof(unsigned long long), st));
    uint32_t nsub = 1, band_rows = b - a;
    band_split(s, f, a, b, default_sub, nsub, band_rows);
    s->cur_nsub = nsub;
    if (nsub <= 1) {
        s->profile_kernels = true;
        if (int rc = launch_rows(s, f, a, b, st, 0, clear_mode)) return rc;
        return hook(0u, a, b, st);
    }
    cudaEvent_t* pe = nullptr;
    if (s->profile) {       // concurrent bands: only the frame total is meaningful
        if (s->prof_events.size() < 4 * (s->prof_frames + 1))
            for (int k = 0; k < 4; ++k) { cudaEvent_t e; CUDA_TRY(cudaEventCreate(&e)); s->prof_events.push_back(e); }
        pe = &s->prof_events[4 * s->prof_frames++];
        for (int k = 0; k < 3; ++k) CUDA_TRY(cudaEventRecord(pe[k], st));
    }
    s->profile_kernels = false;
    // CERES_TRACE_TIMING: device timeline of the bands (kernels done / copies done, relative to the fork)
    static const bool dbg = std::getenv("CERES_TRACE_TIMING") != nullptr;
    static cudaEvent_t dbg_ev[1 + 2 * CERES_MAX_SUB] = {};
    if (dbg && !dbg_ev[0]) for (auto& e : dbg_ev) CUDA_TRY(cudaEventCreate(&e));
    if (dbg) CUDA_TRY(cudaEventRecord(dbg_ev[0], st));
    CUDA_TRY(cudaEventRecord(s->fork_ev, st));
    uint32_t k = 0;
    for (uint32_t r0 = a; r0 < b; r0 += band_rows, ++k) {
        const uint32_t r1 = std::min(b, r0 + band_rows);
        cudaStream_t bs = s->sub_stream[k];
        CUDA_TRY(cudaStreamWaitEvent(bs, s->fork_ev, 0));
        if (int rc = launch_rows(s, f, r0, r1, bs, (int)k, clear_mode)) return rc;
        if (dbg) CUDA_TRY(cudaEventRecord(dbg_ev[1 + 2 * k], bs));
        if (int rc = hook(k, r0, r1, bs)) return rc;
        if (dbg) CUDA_TRY(cudaEventRecord(dbg_ev[2 + 2 * k], bs));
        CUDA_TRY(cudaEventRecord(s->sub_done[k], bs));
        CUDA_TRY(cudaStreamWaitEvent(st, s->sub_done[k], 0));
    }
    s->profile_kernels = true;
    if (pe) CUDA_TRY(cudaEventRecord(pe[3], st));
    if (dbg) {
        CUDA_TRY(cudaStreamSynchronize(st));
        std::fprintf(stderr, "band timeline [us] (kernels done / hook done):");
        for (uint32_t q = 0; q < k; ++q) {
            float t1 = 0, t2 = 0;
            cudaEventElapsedTime(&t1, dbg_ev[0], dbg_ev[1 + 2 * q]); cudaEventElapsedTime(&t2, dbg_ev[0], dbg_ev[2 + 2 * q]);
            std::fprintf(stderr, " %.0f/%.0f", t1 * 1e3, t2 * 1e3);
        }
        std::fprintf(stderr, "\n");
    }
    return CERES_OK;
}

extern "C" {

int ceres_render_device(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height,
                        unsigned flags, const ceres_shard* shard, void* stream_v, float* d_pixels, uint32_t* d_ids,
                        uint64_t* d_counters) {
    Frame f; uint32_t rows_local = 0;
    if (int rc = prepare_frame(s, cam, sun, width, height, flags, shard, d_pixels, d_ids, d_counters, f, rows_local)) return rc;
    DeviceGuard guard(s->device);
    cudaStream_t st = stream_v ? (cudaStream_t)stream_v : s->stream;
    if (rows_local == 0) {
        if (d_counters) CUDA_TRY(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint64_t), st));
        return CERES_OK;
    }
    s->rays_outside = 0;
    return render_rows(s, f, 0, rows_local, st, CERES_SUBFRAMES_DEVICE, CLEAR_OUTSIDE, [](uint32_t, uint32_t, uint32_t, cudaStream_t) { return (int)CERES_OK; });
}

}  // extern "C"

// ---- ceres_render_device_static: the frame kernel over scratch that is never allocated (include/ceres_b200.h) ----
// The frame kernel's scratch -- the record queue, its ready[] credits, one FrameCounters slot, a (rays, hits) pair -- in a
// file-scope device array instead of the scene's cudaMalloc'ed, grow-on-demand buffers.  Every GPU that loads the library
// holds CERES_STATIC_SCRATCH_BYTES of it until exit.  Layout: [control 256 B][ready[] STATIC_READY_WORDS][queue MAX_PIXELS].
#define STATIC_MAX_SMS 256u
#define STATIC_READY_WORDS ((((size_t)CERES_STATIC_MAX_PIXELS >> READY_MIN_SHIFT) + 2 + (size_t)STATIC_MAX_SMS * 16 * (FRAME_THREADS / 32) + 63) / 64 * 64)
#define STATIC_CTL_BYTES 256
#define STATIC_COUNTERS_OFFSET 128
#define STATIC_READY_OFFSET ((size_t)STATIC_CTL_BYTES)
#define STATIC_QUEUE_OFFSET (STATIC_READY_OFFSET + STATIC_READY_WORDS * sizeof(uint32_t))
#define STATIC_SCRATCH_BYTES (STATIC_QUEUE_OFFSET + (size_t)CERES_STATIC_MAX_PIXELS * sizeof(uint4))
static_assert(sizeof(FrameCounters) <= STATIC_COUNTERS_OFFSET, "control block layout");
static_assert(STATIC_SCRATCH_BYTES == CERES_STATIC_SCRATCH_BYTES, "include/ceres_b200.h states the size of the static scratch");
static_assert(STATIC_QUEUE_OFFSET % 256 == 0, "queue alignment");

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// The array keeps the last call's contents (a call that faulted or expired may have left anything): put the words the
// frame kernel expects in their initial state -- ready[] all zero, the counter slot armed, the (rays, hits) pair zero.
// Device code names the array directly; only the host needs its address (static_scratch_setup).
__global__ void static_scratch_arm_kernel(uint32_t ready_words) {
    uint32_t* ready = (uint32_t*)(g_static_scratch + STATIC_READY_OFFSET);
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < ready_words; k += gridDim.x * blockDim.x) ready[k] = 0u;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        FrameCounters c = {};
        c.ext[0] = 0xFFFFFFFFu; c.ext[2] = 0xFFFFFFFFu;
        *(FrameCounters*)g_static_scratch = c;
        unsigned long long* out = (unsigned long long*)(g_static_scratch + STATIC_COUNTERS_OFFSET);
        out[0] = 0ull; out[1] = 0ull;
    }
}

// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, asked once per device with that device current.
struct StaticScratchDev { unsigned char* ptr = nullptr; };
static StaticScratchDev g_static_dev[64];
static std::mutex g_static_mu;

static int static_scratch_setup(int device, unsigned char** out) {       // the caller has made `device` current
    if (device < 0 || device >= 64) return fail(CERES_ERR_INVALID, "ceres_render_device_static: device index out of range");
    std::lock_guard<std::mutex> lock(g_static_mu);
    StaticScratchDev& d = g_static_dev[device];
    if (!d.ptr) {
        int cur = -1;
        CUDA_TRY(cudaGetDevice(&cur));
        if (cur != device) return fail(CERES_ERR_CUDA, "ceres_render_device_static: the scene's device is not current");
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || ((uintptr_t)p & 255u)) return fail(CERES_ERR_CUDA, "ceres_render_device_static: bad address for the static scratch");
        d.ptr = (unsigned char*)p;
    }
    *out = d.ptr;
    return CERES_OK;
}

extern "C" {

int ceres_render_device_static(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height,
                               unsigned flags, void* stream_v, float* d_pixels, uint32_t* d_ids, uint64_t* d_counters) {
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_render_device_static");
    if (width && height && (uint64_t)width * height > (uint64_t)CERES_STATIC_MAX_PIXELS)
        return fail(CERES_ERR_INVALID, "ceres_render_device_static: " + std::to_string(width) + " x " + std::to_string(height) +
                                       " pixels exceed CERES_STATIC_MAX_PIXELS (" + std::to_string(CERES_STATIC_MAX_PIXELS) +
                                       "), the size the static scratch was built for; use ceres_render_device");
    if (flags & ~CERES_NO_SHADOWS) return fail(CERES_ERR_INVALID, "ceres_render_device_static: the only flag it takes is CERES_NO_SHADOWS");
    if (s && (s->stats || s->dev.timeline || s->tune.variant != 0))
        return fail(CERES_ERR_UNSUPPORTED, "ceres_render_device_static: statistics, timeline and tuning variants are ceres_render_device's");
    if (s && s->sm_count > (int)STATIC_MAX_SMS) return fail(CERES_ERR_UNSUPPORTED, "ceres_render_device_static: more SMs than the static scratch was built for");
    Frame f; uint32_t rows_local = 0;
    if (int rc = prepare_frame(s, cam, sun, width, height, flags, nullptr, d_pixels, d_ids, d_counters, f, rows_local)) return rc;
    DeviceGuard guard(s->device);
    if (!guard.ok) return fail(CERES_ERR_CUDA, "ceres_render_device_static: cudaSetDevice failed");
    unsigned char* scratch = nullptr;
    if (int rc = static_scratch_setup(s->device, &scratch)) return rc;
    cudaStream_t st = stream_v ? (cudaStream_t)stream_v : s->stream;

    // one launch over the whole frame (what launch_rows does for band 0 of 1, with the scratch pointers replaced)
    f.row_begin = 0; f.row_end = rows_local; f.accumulate = 1u;
    f.ctr = (FrameCounters*)scratch;
    f.out_counters = d_counters ? (unsigned long long*)d_counters : (unsigned long long*)(scratch + STATIC_COUNTERS_OFFSET);
    f.out_stats = nullptr;
    f.ready = (uint32_t*)(scratch + STATIC_READY_OFFSET);
    f.queue = (uint4*)(scratch + STATIC_QUEUE_OFFSET);
    f.out_error = s->d_out + 10;                       // ceres_scene_frame_error() reads it as for every other frame of the scene
    f.queue_cap = rows_local * f.W;
    f.ready_cap = (uint32_t)ready_words(s, f.queue_cap);
    f.pix_rows = f.H;
    f.rect_x0 = f.cull_i0; f.rect_x1 = f.cull_i1; f.rect_row0 = f.cull_l0; f.rect_row1 = std::max(f.cull_l1, f.cull_l0);
    f.tiles_x = (f.rect_x1 - f.rect_x0 + 7u) / 8u;
    f.tiles_y = (f.rect_row1 - f.rect_row0 + 3u) / 4u;
    f.work_items = f.tiles_x * f.tiles_y * 32u;        // <= 32 * (MAX_PIXELS / 32 + edge tiles): far below 2^32
    f.clear_rows = 1u;                                 // a caller-owned device frame: the kernel zeroes what it does not write
    const uint64_t warps_guess = (uint64_t)s->sm_count * 8u * (FRAME_THREADS / 32);
    const bool share = s->tune.reserved[0] ? s->tune.reserved[0] == 2u : f.work_items / warps_guess < FRAME_SHARE_AUTO_PX;
    f.share_period = (s->tune.reserved[2] & 0xFFu) ? (s->tune.reserved[2] & 0xFFu) : FRAME_SHARE_PERIOD;
    f.share_min_idle = ((s->tune.reserved[2] >> 8) & 0xFFu) ? ((s->tune.reserved[2] >> 8) & 0xFFu) : FRAME_SHARE_MIN_IDLE;
    const uint32_t want = s->tune.shadow_block ? s->tune.shadow_block : FRAME_DEFAULT_CHUNK;
    f.chunk_shift = want <= 8u ? 3u : want <= 16u ? 4u : 5u;
    if (f.ready_cap > STATIC_READY_WORDS) return fail(CERES_ERR_INVALID, "internal: static ready[] too small");

    const size_t smem = (size_t)s->dev.stack_depth * 4 * FRAME_THREADS;
    const int order = s->tune.shadow_order == 1 ? ORDER_NEAR : s->tune.shadow_order == 2 ? ORDER_LEFT : FRAME_DEFAULT_SHADOW_ORDER;
    auto go = [&](auto kernel, int slot) -> int {
        int& bps = s->trace_blocks_per_sm[slot];       // the same cache (and values) as launch_rows
        if (bps == 0) {
            if (s->tune.ctas_per_sm) bps = (int)std::min(16u, s->tune.ctas_per_sm);
            else CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, FRAME_THREADS, smem));
            if (bps < 1) bps = 1;
            if (bps > 16) bps = 16;
        }
        if (d_counters) CUDA_TRY(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint64_t), st));
        static_scratch_arm_kernel<<<(unsigned)std::min<uint32_t>((f.ready_cap + 255u) / 256u, (uint32_t)s->sm_count * 4u), 256, 0, st>>>(f.ready_cap);
        g_launches++;
        CUDA_TRY(cudaGetLastError());
        kernel<<<(unsigned)(s->sm_count * bps), FRAME_THREADS, smem, st>>>(s->dev, f);
        g_launches++;
        CUDA_TRY(cudaGetLastError());
        return CERES_OK;
    };
    s->rays_outside = 0;
    const bool near = order == ORDER_NEAR;
    return share ? (near ? go(frame_kernel<false, ORDER_NEAR, true, false>, 0) : go(frame_kernel<false, ORDER_LEFT, true, false>, 2))
                 : (near ? go(frame_kernel<false, ORDER_NEAR, false, false>, 4) : go(frame_kernel<false, ORDER_LEFT, false, false>, 6));
}

// ---- peer frame (include/ceres_b200.h: "one frame on N GPUs, fused with its gather over NVLink peer memory") ----
struct ceres_peer_frame {
    void* base = nullptr;          // the allocation (owner) / the IPC mapping (opened) / the owner's base (alias)
    bool owner = false, mapped = false;
    int device = 0, world = 1, rank = 0;
    size_t W = 0, H = 0;
    bool with_ids = false;
    uint32_t epoch = 0;
    PeerCtl* ctl = nullptr;
    float* pixels = nullptr;
    uint32_t* ids = nullptr;
    size_t n4 = 0, ni4 = 0;
    // Dirty-rectangle bookkeeping of THIS rank's rows.  Invariant between frames: the rank's rows hold the last frame's
    // values inside `prev` (columns [i0, i1) x local rows [l0, l1)) and zeros / CERES_MISS everywhere else.
    uint32_t prev_i0 = 0, prev_i1 = 0, prev_l0 = 0, prev_l1 = 0;
    uint32_t* d_gate = nullptr;    // ranks > 0: two words in THIS rank's memory, the local relay of rank 0's epoch flag (Frame::gate)
    bool dirty_all = false;        // somebody else wrote into the frame (ceres_peer_frame_invalidate): clear all my rows once
};

static size_t peer_layout(ceres_peer_frame* pf) {
    const size_t ctl_bytes = (sizeof(PeerCtl) + 4095) / 4096 * 4096;
    pf->n4 = (pf->W * pf->H * 3 + 3) / 4;
    pf->ni4 = pf->with_ids ? (pf->W * pf->H + 3) / 4 : 0;
    char* b = (char*)pf->base;
    pf->ctl = (PeerCtl*)b;
    pf->pixels = (float*)(b + ctl_bytes);
    pf->ids = pf->with_ids ? (uint32_t*)(b + ctl_bytes + pf->n4 * 16) : nullptr;
    return ctl_bytes + pf->n4 * 16 + pf->ni4 * 16;
}

__global__ void peer_publish_kernel(PeerCtl* ctl, uint32_t rank, uint32_t epoch) {
    ctl->counters[rank][0] = 0; ctl->counters[rank][1] = 0;
    __threadfence_system();
    st_release_sys(&ctl->done_epoch[rank], epoch);
}

// The record queue of the scene is sized when the frame handle is made: growing it later would need a device-wide
// synchronisation in the middle of a frame that other ranks (possibly on this very device) are waiting on.
static int peer_size_queue(ceres_scene* s, size_t width, size_t height, int world, int rank) {
    ceres_shard sh{CERES_TILE_BLOCK_ROWS, (uint32_t)rank, (uint32_t)world};
    const size_t rows = shard_rows(height, world > 1 ? &sh : nullptr);
    DeviceGuard guard(s->device);
    return size_queue(s, rows * width);
}

static int peer_args_ok(ceres_scene* s, size_t width, size_t height, int world, int rank) {
    if (!s) return fail(CERES_ERR_INVALID, "ceres_peer_frame: null scene");
    if (!width || !height || (uint64_t)width * height >= (1ull << 32)) return fail(CERES_ERR_INVALID, "ceres_peer_frame: bad image size");
    if (world < 1 || world > CERES_PEER_MAX_RANKS || rank < 0 || rank >= world) return fail(CERES_ERR_INVALID, "ceres_peer_frame: bad world/rank");
    return CERES_OK;
}

int ceres_peer_frame_create(ceres_scene* s, size_t width, size_t height, int want_ids, int world, ceres_peer_frame** out,
                            unsigned char* handle_out) {
    if (int rc = peer_args_ok(s, width, height, world, 0)) return rc;
    if (!out) return fail(CERES_ERR_INVALID, "ceres_peer_frame_create: null out");
    static_assert(sizeof(cudaIpcMemHandle_t) <= CERES_PEER_HANDLE_BYTES, "handle size");
    if (int rc = peer_size_queue(s, width, height, world, 0)) return rc;
    DeviceGuard guard(s->device);
    ceres_peer_frame* pf = new (std::nothrow) ceres_peer_frame();
    if (!pf) return fail(CERES_ERR_INVALID, "out of host memory");
    pf->owner = true; pf->device = s->device; pf->world = world; pf->rank = 0; pf->W = width; pf->H = height; pf->with_ids = want_ids != 0;
    const size_t bytes = peer_layout(pf);      // base is null: only the size is meaningful here
    cudaError_t e = cudaMalloc(&pf->base, bytes);
    if (e == cudaSuccess) e = cudaMemset(pf->base, 0, bytes);                       // pixels: zeros
    if (e == cudaSuccess && pf->with_ids) { peer_layout(pf); e = cudaMemset(pf->ids, 0xFF, pf->ni4 * 16); }    // ids: CERES_MISS
    if (e == cudaSuccess && handle_out) {
        cudaIpcMemHandle_t h;
        e = cudaIpcGetMemHandle(&h, pf->base);
        if (e == cudaSuccess) { std::memset(handle_out, 0, CERES_PEER_HANDLE_BYTES); std::memcpy(handle_out, &h, sizeof(h)); }
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();      // legacy-stream memsets are complete before any rank's (non-blocking) stream stores
    if (e != cudaSuccess) { cudaFree(pf->base); delete pf; return fail(CERES_ERR_CUDA, std::string("ceres_peer_frame_create: ") + cudaGetErrorString(e)); }
    peer_layout(pf);
    *out = pf;
    return CERES_OK;
}

int ceres_peer_frame_open(ceres_scene* s, const unsigned char* handle, size_t width, size_t height, int want_ids, int world,
                          int rank, ceres_peer_frame** out) {
    if (int rc = peer_args_ok(s, width, height, world, rank)) return rc;
    if (!out || !handle || rank == 0) return fail(CERES_ERR_INVALID, "ceres_peer_frame_open: bad argument (rank 0 creates the frame)");
    if (int rc = peer_size_queue(s, width, height, world, rank)) return rc;
    DeviceGuard guard(s->device);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* base = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    ceres_peer_frame* pf = new (std::nothrow) ceres_peer_frame();
    if (!pf) { cudaIpcCloseMemHandle(base); return fail(CERES_ERR_INVALID, "out of host memory"); }
    pf->base = base; pf->mapped = true; pf->device = s->device; pf->world = world; pf->rank = rank;
    pf->W = width; pf->H = height; pf->with_ids = want_ids != 0;
    peer_layout(pf);
    if (cudaMalloc((void**)&pf->d_gate, 2 * sizeof(uint32_t)) != cudaSuccess || cudaMemset(pf->d_gate, 0, 2 * sizeof(uint32_t)) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        cudaIpcCloseMemHandle(base); cudaFree(pf->d_gate); delete pf;
        return fail(CERES_ERR_CUDA, std::string("ceres_peer_frame_open: ") + cudaGetErrorString(cudaGetLastError()));
    }
    *out = pf;
    return CERES_OK;
}

int ceres_peer_frame_alias(ceres_peer_frame* owner, ceres_scene* s, int rank, ceres_peer_frame** out) {
    if (!owner || !owner->owner || !out) return fail(CERES_ERR_INVALID, "ceres_peer_frame_alias: needs the owning frame");
    if (int rc = peer_args_ok(s, owner->W, owner->H, owner->world, rank)) return rc;
    if (rank == 0) return fail(CERES_ERR_INVALID, "ceres_peer_frame_alias: rank 0 is the owner");
    if (int rc = peer_size_queue(s, owner->W, owner->H, owner->world, rank)) return rc;
    if (s->device != owner->device) {
        DeviceGuard guard(s->device);
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, s->device, owner->device));
        if (!can) return fail(CERES_ERR_UNSUPPORTED, "ceres_peer_frame_alias: no peer access between the two GPUs");
        cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(CERES_ERR_CUDA, cudaGetErrorString(e));
        cudaGetLastError();
    }
    ceres_peer_frame* pf = new (std::nothrow) ceres_peer_frame(*owner);
    if (!pf) return fail(CERES_ERR_INVALID, "out of host memory");
    pf->owner = false; pf->mapped = false; pf->device = s->device; pf->rank = rank; pf->epoch = owner->epoch;
    pf->d_gate = nullptr;
    {
        DeviceGuard guard(s->device);
        // the relay starts at the owner's current epoch: everything up to it has been announced
        if (cudaMalloc((void**)&pf->d_gate, 2 * sizeof(uint32_t)) != cudaSuccess ||
            cudaMemcpy(pf->d_gate, &pf->epoch, sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemcpy(pf->d_gate + 1, &pf->epoch, sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaDeviceSynchronize() != cudaSuccess) {
            cudaFree(pf->d_gate); delete pf;
            return fail(CERES_ERR_CUDA, std::string("ceres_peer_frame_alias: ") + cudaGetErrorString(cudaGetLastError()));
        }
    }
    *out = pf;
    return CERES_OK;
}

void ceres_peer_frame_destroy(ceres_peer_frame* pf) {
    if (!pf) return;
    DeviceGuard guard(pf->device);
    cudaDeviceSynchronize();
    if (pf->owner) cudaFree(pf->base);
    else if (pf->mapped) cudaIpcCloseMemHandle(pf->base);
    cudaFree(pf->d_gate);
    delete pf;
}

float* ceres_peer_frame_pixels(ceres_peer_frame* pf) { return pf ? pf->pixels : nullptr; }
uint32_t* ceres_peer_frame_ids(ceres_peer_frame* pf) { return pf ? pf->ids : nullptr; }

int ceres_peer_frame_error(ceres_peer_frame* pf) {
    if (!pf) return -1;
    DeviceGuard guard(pf->device);
    uint32_t e = 0;
    if (cudaMemcpy(&e, &pf->ctl->error, sizeof(e), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int)e;
}

int ceres_peer_frame_invalidate(ceres_peer_frame* pf) {
    if (!pf) return fail(CERES_ERR_INVALID, "ceres_peer_frame_invalidate: null frame");
    pf->dirty_all = true;
    return CERES_OK;
}

int ceres_render_peer(ceres_scene* s, const ceres_camera* cam, const float sun[3], unsigned flags, void* stream_v,
                      ceres_peer_frame* pf, uint64_t* d_counters) {
    if (!s || !pf) return fail(CERES_ERR_INVALID, "ceres_render_peer: null argument");
    if (s->device != pf->device) return fail(CERES_ERR_INVALID, "ceres_render_peer: scene and frame handle belong to different devices");
    if (s->tune.variant != 0) return fail(CERES_ERR_UNSUPPORTED, "ceres_render_peer: needs the frame kernel (tuning variant 0)");
    ceres_shard sh{CERES_TILE_BLOCK_ROWS, (uint32_t)pf->rank, (uint32_t)pf->world};
    Frame f; uint32_t rows_local = 0;
    if (int rc = prepare_frame(s, cam, sun, pf->W, pf->H, flags & ~CERES_COMPACT_ROWS, pf->world > 1 ? &sh : nullptr, pf->pixels,
                               pf->ids, nullptr, f, rows_local)) return rc;
    DeviceGuard guard(s->device);
    cudaStream_t st = stream_v ? (cudaStream_t)stream_v : s->stream;
    const uint32_t epoch = ++pf->epoch;
    f.peer = pf->ctl; f.peer_epoch = epoch; f.peer_rank = (uint32_t)pf->rank; f.gate = pf->d_gate;
    if (pf->rank == 0) { f.peer_wait_world = (uint32_t)pf->world; f.peer_out = (unsigned long long*)d_counters; }
    s->rays_outside = 0;
    if (rows_local == 0) {
        peer_publish_kernel<<<1, 1, 0, st>>>(pf->ctl, (uint32_t)pf->rank, epoch);
        g_launches++;
        CUDA_TRY(cudaGetLastError());
        return CERES_OK;
    }
    // Dirty-rectangle clear of MY rows: the frame kernel writes every pixel of the current rectangle; outside it my rows
    // already hold zeros except where the previous frame's rectangle sticks out (a moving camera), or everywhere after
    // ceres_peer_frame_invalidate.  Nothing to do for a repeated or growing rectangle -- no clear pass, no cross-rank wait.
    const uint32_t l0 = std::min(f.cull_l0, rows_local), l1 = std::max(l0, std::min(f.cull_l1, rows_local));
    const uint32_t i0 = std::min(f.cull_i0, f.W), i1 = std::max(i0, std::min(f.cull_i1, f.W));
    const bool cur_empty = l1 == l0 || i1 == i0;
    ClearRegion q{0, 0, 0, 0, l0, l1, i0, i1};
    if (pf->dirty_all) { q.row_a = 0; q.row_b = rows_local; q.x_a = 0; q.x_b = f.W; }
    else if (pf->prev_l1 > pf->prev_l0 && pf->prev_i1 > pf->prev_i0 &&
             (cur_empty || pf->prev_l0 < l0 || pf->prev_l1 > l1 || pf->prev_i0 < i0 || pf->prev_i1 > i1)) {
        q.row_a = pf->prev_l0; q.row_b = pf->prev_l1; q.x_a = pf->prev_i0; q.x_b = pf->prev_i1;
    }
    if (q.row_b > q.row_a) {
        const unsigned grid = (unsigned)std::min<uint32_t>(q.row_b - q.row_a, (uint32_t)s->sm_count * 8u);
        clear_outside_kernel<<<grid, 256, 0, st>>>(f, q);
        g_launches++;
        CUDA_TRY(cudaGetLastError());
    }
    pf->dirty_all = false;
    pf->prev_l0 = l0; pf->prev_l1 = l1; pf->prev_i0 = i0; pf->prev_i1 = i1;
    return render_rows(s, f, 0, rows_local, st, 1u, CLEAR_NONE, [](uint32_t, uint32_t, uint32_t, cudaStream_t) { return (int)CERES_OK; });
}

// Host-framebuffer render (the drop-in for render.hpp:86-156).
//   * Only the conservative screen rectangle of the scene is rendered and copied: its rows are traced in bands on the
//     scene stream, and each finished band is DMA-copied (columns of the rectangle only, cudaMemcpy2DAsync) on a
//     second stream while the next band renders.
//   * Everything outside the rectangle is 0 by construction (render.hpp:116-117; ids: CERES_MISS) and is written by
//     the host-fill pool (host_fill.hpp) concurrently with the GPU work -- zeros never cross PCIe.
int ceres_render_shard(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height,
                       unsigned flags, const ceres_shard* shard, float* pixels, uint32_t* ids, uint64_t counters[2]) {
    if (!s || !pixels) return fail(CERES_ERR_INVALID, "ceres_render: null argument");
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_render");
    DeviceGuard guard(s->device);
    const size_t npix = width * height;
    // device staging is written in COMPACT form (the shard's rows packed); rows are scattered to their frame
    // position by the copies below, so a shard moves only its own bytes over PCIe.
    const bool sharded = shard && shard->count > 1;
    const unsigned dev_flags = flags | (sharded ? CERES_COMPACT_ROWS : 0u);
    Frame f; uint32_t rows = 0;
    {
        float dummy; // prepare_frame validates before we size the staging buffers
        if (int rc = prepare_frame(s, cam, sun, width, height, dev_flags, shard, &dummy, nullptr, nullptr, f, rows)) return rc;
    }
    if (int rc = ensure(s->d_pixels, s->pixels_cap, npix * 3)) return rc;
    if (ids) if (int rc = ensure(s->d_ids, s->ids_cap, npix)) return rc;
    f.pixels = s->d_pixels; f.ids = ids ? s->d_ids : nullptr;
    if (rows == 0) { if (counters) counters[0] = counters[1] = 0; return CERES_OK; }

    const bool want_compact = (flags & CERES_COMPACT_ROWS) != 0;
    const bool scatter = sharded && !want_compact;
    auto out_row = [&](uint32_t lrow) -> size_t { return scatter ? (size_t)frame_row_host(f, lrow) : (size_t)lrow; };
    // rectangle in local rows / columns (whole shard when culling is off)
    const uint32_t l0 = std::min(f.cull_l0, rows), l1 = std::max(l0, std::min(f.cull_l1, rows));
    const uint32_t i0 = std::min<uint32_t>(f.cull_i0, (uint32_t)width), i1 = std::max(i0, std::min<uint32_t>(f.cull_i1, (uint32_t)width));
    const bool empty = l1 == l0 || i1 == i0;

    // ---- which path: tight band copies with in-kernel bands, and the CERES_FRAME_RETAINED contract on top of them -----------
    static const bool tight_off = std::getenv("CERES_NO_TIGHT_COPY") != nullptr;
    const bool tight = !tight_off && s->tune.variant != 1;
    uint32_t nb = 1, trpb = 0;                               // band layout of the tight path: nb launches of trpb rows
    if (tight && !empty) band_split(s, f, l0, l1, CERES_SUBFRAMES_HOST, nb, trpb);
    const ceres_scene::Retained old = s->kept;
    s->kept.valid = false;                                   // re-established at the end of a successful retained call
    const bool retained_req = (flags & CERES_FRAME_RETAINED) != 0 && tight;
    const uint32_t sh3[3] = {sharded ? shard->block_rows : 0u, sharded ? shard->index : 0u, sharded ? shard->count : 0u};
    const bool kept_ok = retained_req && old.valid && old.px == pixels && old.ids == ids && old.W == width && old.H == height &&
                         old.compact == want_compact && old.sh[0] == sh3[0] && old.sh[1] == sh3[1] && old.sh[2] == sh3[2];
    const bool same_layout = kept_ok && !empty && old.nb == nb && old.trpb == trpb && old.i0 == i0 && old.i1 == i1 && old.l0 == l0 && old.l1 == l1;

    // ---- host fill of everything outside the rectangle ---------------------------------------------------------
    ceres_host::Job job;
    if (kept_ok) {
        // the buffers still hold the previous frame: zeros / CERES_MISS everywhere but inside its bands' rectangles.  Same band
        // layout: every band zeroes what its old rectangle leaves uncovered once its new one is known (below).  Otherwise all
        // old rectangles are wiped here, before the first copy of the new frame lands.
        if (!same_layout)
            for (uint32_t b = 0; b < old.nb; ++b) {
                const uint32_t* e = old.ext[b];
                if (e[0] > e[1]) continue;
                for (uint32_t r = e[2]; r <= e[3]; ++r) {
                    const size_t o = out_row(r) * width + e[0];
                    job.segs.push_back({reinterpret_cast<unsigned char*>(pixels + 3 * o), (size_t)(e[1] - e[0] + 1) * 12, 0x00});
                    if (ids) job.segs.push_back({reinterpret_cast<unsigned char*>(ids + o), (size_t)(e[1] - e[0] + 1) * 4, 0xFF});
                }
            }
    } else {
        auto fill_rows = [&](uint32_t a, uint32_t b) {       // local rows [a, b), full width; merge rows contiguous in the output
            uint32_t r = a;
            while (r < b) {
                uint32_t e = r + 1;
                while (e < b && out_row(e) == out_row(e - 1) + 1) ++e;
                job.add(pixels + 3 * out_row(r) * width, (size_t)(e - r) * width * 12, 0x00);
                if (ids) job.add(ids + out_row(r) * width, (size_t)(e - r) * width * 4, 0xFF);
                r = e;
            }
        };
        if (empty) fill_rows(0, rows);
        else {
            fill_rows(0, l0);
            fill_rows(l1, rows);
            for (uint32_t r = l0; r < l1 && (i0 > 0 || i1 < width); ++r) {
                const size_t o = out_row(r) * width;
                if (i0 > 0) { job.add(pixels + 3 * o, (size_t)i0 * 12, 0x00); if (ids) job.add(ids + o, (size_t)i0 * 4, 0xFF); }
                if (i1 < width) { job.add(pixels + 3 * (o + i1), (width - i1) * 12, 0x00); if (ids) job.add(ids + o + i1, (width - i1) * 4, 0xFF); }
            }
        }
    }
    static const bool trace_timing = std::getenv("CERES_TRACE_TIMING") != nullptr;
    auto now_us = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = trace_timing ? now_us() : 0.0;
    static const bool no_fill = std::getenv("CERES_DEBUG_NO_FILL") != nullptr;     // timing experiments only
    if (no_fill) job.segs.clear();
    ceres_host::Pool& pool = ceres_host::Pool::instance();
    pool.submit(&job);
    const double t_submit = trace_timing ? now_us() : 0.0;
    struct Waiter { ceres_host::Pool& p; ceres_host::Job* j; ~Waiter() { p.wait(j); } } waiter{pool, &job};   // also on error paths
    // ... and no return -- error paths included -- leaves kernels or DMA copies in flight that still write the caller's buffers
    struct Drain {
        ceres_scene* s; bool armed = true;
        ~Drain() {
            if (!armed) return;
            cudaStreamSynchronize(s->copy_stream);
            for (int k = 0; k < CERES_MAX_SUB; ++k) cudaStreamSynchronize(s->sub_stream[k]);
            cudaStreamSynchronize(s->stream);
            cudaGetLastError();
        }
    } drain{s};

    const uint64_t prim_outside = (uint64_t)width * (rows - (empty ? 0 : l1 - l0));
    s->last_d2h_bytes = empty ? 0 : (uint64_t)(l1 - l0) * (i1 - i0) * (ids ? 16u : 12u) + 2 * sizeof(unsigned long long);
    if (empty) {
        if (counters) { counters[0] = (uint64_t)width * rows; counters[1] = 0; }
        pool.wait(&job);
        if (retained_req) {
            s->kept = ceres_scene::Retained{};
            s->kept.valid = true; s->kept.px = pixels; s->kept.ids = ids; s->kept.W = width; s->kept.H = height; s->kept.compact = want_compact;
            for (int k = 0; k < 3; ++k) s->kept.sh[k] = sh3[k];
        }
        return CERES_OK;
    }

    // ---- GPU: concurrent row bands of the rectangle; each band is copied as soon as its own kernels are done ---------
    // local rows [a, b) x columns [x0, x1), rows contiguous in the caller's frame: one pitched DMA
    auto copy_run = [&](uint32_t a, uint32_t b, uint32_t x0, uint32_t x1, cudaStream_t bs) -> int {
        if (b <= a || x1 <= x0) return CERES_OK;
        const size_t src = (size_t)a * width + x0, dst = out_row(a) * width + x0;
        CUDA_TRY(cudaMemcpy2DAsync(pixels + 3 * dst, width * 12, s->d_pixels + 3 * src, width * 12, (size_t)(x1 - x0) * 12, b - a,
                                   cudaMemcpyDeviceToHost, bs));
        if (ids) CUDA_TRY(cudaMemcpy2DAsync(ids + dst, width * 4, s->d_ids + src, width * 4, (size_t)(x1 - x0) * 4, b - a,
                                            cudaMemcpyDeviceToHost, bs));
        return CERES_OK;
    };
    // local rows [r0, r1) x columns [x0, x1) to their place in the caller's frame
    auto copy_rect = [&](uint32_t r0, uint32_t r1, uint32_t x0, uint32_t x1, cudaStream_t bs) -> int {
        if (!scatter) return copy_run(r0, r1, x0, x1, bs);
        // scattered shard: local block lb (block_rows packed rows) lands at frame block lb * count + index.  The whole
        // blocks of the range go in ONE strided 3-D DMA (slice = one block; source slices are adjacent, destination
        // slices are `count` blocks apart); a partial block at either end of the range is a 2-D copy.
        const uint32_t br = f.block_rows;
        const uint32_t b0 = (r0 + br - 1) / br, b1 = r1 / br;              // whole local blocks [b0, b1)
        if (b1 <= b0) {                                                   // the range lies inside one or two blocks
            const uint32_t mid = std::min(r1, (r0 / br + 1) * br);
            if (int rc = copy_run(r0, mid, x0, x1, bs)) return rc;
            return copy_run(mid, r1, x0, x1, bs);
        }
        if (int rc = copy_run(r0, b0 * br, x0, x1, bs)) return rc;
        if (x1 > x0) {
            const size_t src = (size_t)b0 * br * width + x0, dst = out_row(b0 * br) * width + x0;
            cudaMemcpy3DParms p3 = {};
            p3.srcPtr = make_cudaPitchedPtr(s->d_pixels + 3 * src, width * 12, width * 12, br);
            p3.dstPtr = make_cudaPitchedPtr(pixels + 3 * dst, width * 12, width * 12, (size_t)br * f.shard_count);
            p3.extent = make_cudaExtent((size_t)(x1 - x0) * 12, br, b1 - b0);
            p3.kind = cudaMemcpyDeviceToHost;
            CUDA_TRY(cudaMemcpy3DAsync(&p3, bs));
            if (ids) {
                cudaMemcpy3DParms q3 = {};
                q3.srcPtr = make_cudaPitchedPtr(s->d_ids + src, width * 4, width * 4, br);
                q3.dstPtr = make_cudaPitchedPtr(ids + dst, width * 4, width * 4, (size_t)br * f.shard_count);
                q3.extent = make_cudaExtent((size_t)(x1 - x0) * 4, br, b1 - b0);
                q3.kind = cudaMemcpyDeviceToHost;
                CUDA_TRY(cudaMemcpy3DAsync(&q3, bs));
            }
        }
        return copy_run(b1 * br, r1, x0, x1, bs);
    };
    auto copy_band = [&](uint32_t, uint32_t r0, uint32_t r1, cudaStream_t bs) -> int { return copy_rect(r0, r1, i0, i1, bs); };
    s->rays_outside = prim_outside;
    // Tight band copies: the band's last CTA publishes the extents of
    // its hit pixels into mapped host memory; this thread, which has nothing else to do once the frame is enqueued,
    // polls the flags band by band, copies only that sub-rectangle (everything else inside the conservative rectangle is
    // zero / CERES_MISS by construction) and hands the strips around it to the fill pool.  For the dragon the bands'
    // own rectangles hold ~2/3 of the conservative rectangle's bytes, and the copy engine is the serial resource.
    struct Band { uint32_t k, r0, r1; cudaStream_t bs; };
    std::vector<Band> bands;
    std::vector<std::unique_ptr<ceres_host::Job>> gap_jobs;
    struct GapWaiter { ceres_host::Pool& p; std::vector<std::unique_ptr<ceres_host::Job>>& v; ~GapWaiter() { for (auto& j : v) p.wait(j.get()); } } gap_waiter{pool, gap_jobs};
    if (tight) {
        f.host_ext = s->h_ext;
        f.ext_epoch = ++s->ext_epoch;
        if (f.ext_epoch == 0u) f.ext_epoch = ++s->ext_epoch;      // 0 is the value of a never-written flag
        auto note_band = [&](uint32_t k, uint32_t r0, uint32_t r1, cudaStream_t bs) -> int { bands.push_back({k, r0, r1, bs}); return CERES_OK; };
        if (int rc = render_rows(s, f, l0, l1, s->stream, CERES_SUBFRAMES_HOST, CLEAR_NONE, note_band)) return rc;
    } else {
        if (int rc = render_rows(s, f, l0, l1, s->stream, CERES_SUBFRAMES_HOST, CLEAR_NONE, copy_band)) return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(s->h_out, s->d_out, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->h_out + 10, s->d_out + 10, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
    const double t_launched = trace_timing ? now_us() : 0.0;
    if (tight) {
        uint64_t copied = 0;
        for (const Band& b : bands) {
            volatile uint32_t* e = s->h_ext + 8 * b.k;
            // wait for the band; the copy engine is the serial resource, so this thread does nothing else meanwhile (the
            // pool fills without it) -- and a failed launch must not hang the caller
            for (uint64_t spins = 0; e[4] != f.ext_epoch; ++spins) {
                _mm_pause();
                if ((spins & 0x3FFu) == 0x3FFu) {
                    const cudaError_t q = cudaStreamQuery(b.bs);
                    if (q != cudaErrorNotReady && e[4] != f.ext_epoch)
                        return fail(CERES_ERR_CUDA, std::string("ceres_render: band did not complete: ") + cudaGetErrorString(q == cudaSuccess ? cudaErrorUnknown : q));
                }
            }
            std::atomic_thread_fence(std::memory_order_acquire);
            if (trace_timing) std::fprintf(stderr, "  band %u flag at %.0f us\n", b.k, now_us() - t_begin);
            const uint32_t ci0 = e[0], ci1 = e[1], cr0 = e[2], cr1 = e[3];
            const bool any = ci0 <= ci1 && cr0 <= cr1 && ci0 >= i0 && ci1 < i1 && cr0 >= b.r0 && cr1 < b.r1;
            if (!(ci0 > ci1) && !any) return fail(CERES_ERR_CUDA, "ceres_render: band extents outside the band (internal error)");
            auto gaps = std::make_unique<ceres_host::Job>();
            auto strip = [&](uint32_t row, uint32_t a, uint32_t c) {      // columns [a, c) of local row `row`
                if (c <= a) return;
                const size_t o = out_row(row) * width + a;
                gaps->segs.push_back({reinterpret_cast<unsigned char*>(pixels + 3 * o), (size_t)(c - a) * 12, 0x00});
                if (ids) gaps->segs.push_back({reinterpret_cast<unsigned char*>(ids + o), (size_t)(c - a) * 4, 0xFF});
            };
            // what this band has to zero around its new rectangle: everything else of the band's part of the conservative
            // rectangle -- or, under CERES_FRAME_RETAINED with an unchanged band layout, only what the OLD rectangle of the band
            // leaves uncovered (the rest still holds the zeros of the previous call)
            uint32_t zi0 = i0, zi1 = i1, zr0 = b.r0, zr1 = b.r1;        // region to reconcile: columns [zi0, zi1) x rows [zr0, zr1)
            if (same_layout) {
                const uint32_t* o = old.ext[b.k];
                if (o[0] > o[1]) { zr1 = zr0; }                          // the old frame left nothing in this band
                else { zi0 = o[0]; zi1 = o[1] + 1u; zr0 = o[2]; zr1 = o[3] + 1u; }
            } else if (kept_ok && !job.segs.empty()) {
                pool.wait(&job);                                         // every old rectangle is wiped before a new one lands
                job.segs.clear();
            }
            if (kept_ok && !same_layout) zr1 = zr0;                      // already wiped above: nothing to reconcile per band
            if (any) {
                if (int rc = copy_rect(cr0, cr1 + 1u, ci0, ci1 + 1u, s->copy_stream)) return rc;
                copied += (uint64_t)(ci1 - ci0 + 1u) * (cr1 - cr0 + 1u) * (ids ? 16u : 12u);
            }
            for (uint32_t r = zr0; r < zr1; ++r) {
                if (!any || r < cr0 || r > cr1) strip(r, zi0, zi1);
                else { strip(r, zi0, std::min(zi1, ci0)); strip(r, std::max(zi0, ci1 + 1u), zi1); }
            }
            if (retained_req) {
                uint32_t* k = s->kept.ext[b.k];
                if (any) { k[0] = ci0; k[1] = ci1; k[2] = cr0; k[3] = cr1; } else { k[0] = 1u; k[1] = 0u; k[2] = 1u; k[3] = 0u; }
            }
            pool.submit(gaps.get());
            gap_jobs.push_back(std::move(gaps));
        }
        s->last_d2h_bytes = copied + 2 * sizeof(unsigned long long);
    }
    pool.wait(&job);                                      // this thread helps with the fill while the GPU works
    if (tight) {
        for (auto& j : gap_jobs) pool.wait(j.get());
        CUDA_TRY(cudaStreamSynchronize(s->copy_stream));
    }
    const double t_filled = trace_timing ? now_us() : 0.0;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    const double t_kernels = trace_timing ? now_us() : 0.0;
    if (trace_timing)
        std::fprintf(stderr, "ceres_render timing [us]: segments %zu bands %zu | job built+submitted %.0f, launches issued %.0f, fill done %.0f, "
                     "kernels done %.0f, copies done %.0f\n", job.segs.size(), bands.size(), t_submit - t_begin, t_launched - t_begin,
                     t_filled - t_begin, t_kernels - t_begin, now_us() - t_begin);
    drain.armed = false;       // everything was synchronised above
    if (int rc = frame_error_check(s)) return rc;
    if (counters) { counters[0] = s->h_out[0] + prim_outside; counters[1] = s->h_out[1]; }
    if (retained_req) {       // the bands' rectangles were noted above
        s->kept.valid = true; s->kept.px = pixels; s->kept.ids = ids; s->kept.W = width; s->kept.H = height; s->kept.compact = want_compact;
        for (int k = 0; k < 3; ++k) s->kept.sh[k] = sh3[k];
        s->kept.i0 = i0; s->kept.i1 = i1; s->kept.l0 = l0; s->kept.l1 = l1; s->kept.nb = nb; s->kept.trpb = trpb;
    }
    return CERES_OK;
}

// The frame as the payload of the reference's P6 file (static.cpp:135-147): render, quantise on the device, and move
// only the 8-bit screen rectangle over PCIe (a quarter of the float path's bytes); the rest of the image is zero-filled
// by the host pool.  rgb8: HOST 3*width*height bytes, row 0 = frame row height-1.
}  // extern "C"

// The host-frame path for an already quantised, vertically flipped P6 payload with BPC bytes per channel.  With a shard only
// the shard's rows of the full-size payload are written (output row height-1-j for frame row j): the ranks of a box fill one
// shared payload, each over its own PCIe link.
template <int BPC>
static int render_quantized(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height, unsigned flags,
                            const ceres_shard* shard, uint8_t* rgb8, uint64_t counters[2]) {
    if (!s || !rgb8) return fail(CERES_ERR_INVALID, "ceres_render_ppm8/16: null argument");
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_render_ppm8/16");
    DeviceGuard guard(s->device);
    const size_t npix = width * height;
    const bool sharded = shard && shard->count > 1;
    Frame f; uint32_t rows = 0;
    {
        float dummy;
        const unsigned dev_flags = (flags & ~CERES_COMPACT_ROWS) | (sharded ? CERES_COMPACT_ROWS : 0u);      // packed float staging
        if (int rc = prepare_frame(s, cam, sun, width, height, dev_flags, shard, &dummy, nullptr, nullptr, f, rows)) return rc;
    }
    if (rows == 0) { if (counters) counters[0] = counters[1] = 0; return CERES_OK; }
    if (int rc = ensure(s->d_pixels, s->pixels_cap, npix * 3)) return rc;
    if (int rc = ensure(s->d_rgb8, s->rgb8_cap, npix * 3 * BPC)) return rc;
    f.pixels = s->d_pixels; f.ids = nullptr;
    const uint32_t l0 = std::min(f.cull_l0, rows), l1 = std::max(l0, std::min(f.cull_l1, rows));
    const uint32_t i0 = std::min<uint32_t>(f.cull_i0, (uint32_t)width), i1 = std::max(i0, std::min<uint32_t>(f.cull_i1, (uint32_t)width));
    const bool empty = l1 == l0 || i1 == i0;
    const size_t rowb = width * 3 * BPC;
    const bool flip = (flags & CERES_PAYLOAD_NO_FLIP) == 0;
    f.no_flip = flip ? 0u : 1u;
    auto out_row = [&](uint32_t lrow) -> size_t { const size_t j = frame_row_host(f, lrow); return flip ? height - 1 - j : j; };
    // the first (lowest) output row of local rows [a, b), which are contiguous in the output
    auto low_row = [&](uint32_t a, uint32_t b) -> size_t { return flip ? out_row(b - 1) : out_row(a); };
    // host fill of everything outside the rectangle: local rows are ascending frame rows = DESCENDING output rows when flipped
    ceres_host::Job job;
    {
        auto fill_rows = [&](uint32_t a, uint32_t b) {       // local rows [a, b), full width; merge rows contiguous in the output
            uint32_t r = a;
            while (r < b) {
                uint32_t e = r + 1;
                while (e < b && (flip ? out_row(e) + 1 == out_row(e - 1) : out_row(e) == out_row(e - 1) + 1)) ++e;
                job.add(rgb8 + rowb * low_row(r, e), rowb * (e - r), 0x00);
                r = e;
            }
        };
        if (empty) fill_rows(0, rows);
        else {
            fill_rows(0, l0);
            fill_rows(l1, rows);
            for (uint32_t r = l0; r < l1 && (i0 > 0 || i1 < width); ++r) {
                if (i0 > 0) job.add(rgb8 + rowb * out_row(r), (size_t)i0 * 3 * BPC, 0x00);
                if (i1 < width) job.add(rgb8 + rowb * out_row(r) + (size_t)i1 * 3 * BPC, (width - i1) * 3 * BPC, 0x00);
            }
        }
    }
    ceres_host::Pool& pool = ceres_host::Pool::instance();
    pool.submit(&job);
    struct Waiter { ceres_host::Pool& p; ceres_host::Job* j; ~Waiter() { p.wait(j); } } waiter{pool, &job};
    const uint64_t prim_outside = (uint64_t)width * (rows - (empty ? 0 : l1 - l0));
    s->last_d2h_bytes = empty ? 0 : (uint64_t)(l1 - l0) * (i1 - i0) * 3u * BPC + 2 * sizeof(unsigned long long);
    if (empty) {
        if (counters) { counters[0] = (uint64_t)width * rows; counters[1] = 0; }
        return CERES_OK;
    }
    s->rays_outside = prim_outside;
    // device payload and host payload have the same geometry: a run of local rows that is contiguous in the output is one 2-D copy
    auto copy_rows = [&](uint32_t a, uint32_t b, cudaStream_t bs) -> int {      // local rows [a, b), contiguous in the output
        if (b <= a) return CERES_OK;
        const size_t off = (low_row(a, b) * width + i0) * 3 * BPC;
        CUDA_TRY(cudaMemcpy2DAsync(rgb8 + off, rowb, s->d_rgb8 + off, rowb, (size_t)(i1 - i0) * 3 * BPC, b - a, cudaMemcpyDeviceToHost, bs));
        return CERES_OK;
    };
    auto quantize_band = [&](uint32_t, uint32_t r0, uint32_t r1, cudaStream_t bs) -> int {
        quantize_rect_kernel<BPC><<<(unsigned)(s->sm_count * 4), 256, 0, bs>>>(s->d_pixels, f, i0, i1, r0, r1, s->d_rgb8);
        g_launches++;
        CUDA_TRY(cudaGetLastError());
        if (!sharded) return copy_rows(r0, r1, bs);
        // scattered: the whole blocks of the band in ONE strided 3-D copy, enumerated from the LAST local block backwards
        // (output rows grow as frame rows fall); partial blocks at the band's ends as 2-D copies
        const uint32_t br = f.block_rows;
        const uint32_t b0 = (r0 + br - 1) / br, b1 = r1 / br;
        if (b1 <= b0) {
            const uint32_t mid = std::min(r1, (r0 / br + 1) * br);
            if (int rc = copy_rows(r0, mid, bs)) return rc;
            return copy_rows(mid, r1, bs);
        }
        if (int rc = copy_rows(r0, b0 * br, bs)) return rc;
        {
            // flipped: the LAST whole block has the lowest output rows; unflipped: the first one
            const size_t off = ((flip ? out_row(b1 * br - 1) : out_row(b0 * br)) * width + i0) * 3 * BPC;
            cudaMemcpy3DParms p3 = {};
            p3.srcPtr = make_cudaPitchedPtr(s->d_rgb8 + off, rowb, rowb, (size_t)br * f.shard_count);
            p3.dstPtr = make_cudaPitchedPtr(rgb8 + off, rowb, rowb, (size_t)br * f.shard_count);
            p3.extent = make_cudaExtent((size_t)(i1 - i0) * 3 * BPC, br, b1 - b0);
            p3.kind = cudaMemcpyDeviceToHost;
            CUDA_TRY(cudaMemcpy3DAsync(&p3, bs));
        }
        return copy_rows(b1 * br, r1, bs);
    };
    if (int rc = render_rows(s, f, l0, l1, s->stream, CERES_SUBFRAMES_HOST, CLEAR_NONE, quantize_band)) return rc;
    CUDA_TRY(cudaMemcpyAsync(s->h_out, s->d_out, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->h_out + 10, s->d_out + 10, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
    pool.wait(&job);
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (int rc = frame_error_check(s)) return rc;
    if (counters) { counters[0] = s->h_out[0] + prim_outside; counters[1] = s->h_out[1]; }
    return CERES_OK;
}

extern "C" {

int ceres_render_ppm8(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height, unsigned flags,
                      uint8_t* rgb8, uint64_t counters[2]) {
    return render_quantized<1>(s, cam, sun, width, height, flags, nullptr, rgb8, counters);
}

int ceres_render_ppm16(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height, unsigned flags,
                       uint8_t* rgb16, uint64_t counters[2]) {
    return render_quantized<2>(s, cam, sun, width, height, flags, nullptr, rgb16, counters);
}

int ceres_render_ppm8_shard(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height, unsigned flags,
                            const ceres_shard* shard, uint8_t* rgb8, uint64_t counters[2]) {
    return render_quantized<1>(s, cam, sun, width, height, flags, shard, rgb8, counters);
}

int ceres_render_ppm16_shard(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height, unsigned flags,
                             const ceres_shard* shard, uint8_t* rgb16, uint64_t counters[2]) {
    return render_quantized<2>(s, cam, sun, width, height, flags, shard, rgb16, counters);
}

int ceres_render(ceres_scene* s, const ceres_camera* cam, const float sun[3], size_t width, size_t height,
                 unsigned flags, float* pixels, uint32_t* ids, uint64_t counters[2]) {
    return ceres_render_shard(s, cam, sun, width, height, flags & ~CERES_COMPACT_ROWS, nullptr, pixels, ids, counters);
}

// One frame on several GPUs of the box from ONE process: scene k renders the row blocks b = k (mod n) and writes
// them by DMA straight into their place in the caller's host frame -- every GPU drains over its own PCIe link, so
// there is no gather step and no collective.  One host thread per GPU.
int ceres_render_multi(ceres_scene* const* scenes, int n, const ceres_camera* cam, const float sun[3], size_t width,
                       size_t height, unsigned flags, float* pixels, uint32_t* ids, uint64_t counters[2]) {
    if (!scenes || n < 1) return fail(CERES_ERR_INVALID, "ceres_render_multi: no scenes");
    for (int k = 0; k < n; ++k) if (!scenes[k]) return fail(CERES_ERR_INVALID, "ceres_render_multi: null scene");
    if (n == 1) return ceres_render(scenes[0], cam, sun, width, height, flags, pixels, ids, counters);
    std::vector<int> rc(n, CERES_OK);
    std::vector<std::string> err(n);
    std::vector<uint64_t> cnt(2 * (size_t)n, 0);
    auto work = [&](int k) {
        ceres_shard sh{CERES_TILE_BLOCK_ROWS, (uint32_t)k, (uint32_t)n};
        rc[k] = ceres_render_shard(scenes[k], cam, sun, width, height, flags & ~CERES_COMPACT_ROWS, &sh, pixels, ids, &cnt[2 * k]);
        if (rc[k]) err[k] = g_err;      // g_err is thread-local
    };
    std::vector<std::thread> threads;
    for (int k = 1; k < n; ++k) threads.emplace_back(work, k);
    work(0);
    for (auto& t : threads) t.join();
    uint64_t rays = 0, hits = 0;
    for (int k = 0; k < n; ++k) {
        if (rc[k]) return fail(rc[k], "GPU shard " + std::to_string(k) + ": " + err[k]);
        rays += cnt[2 * k]; hits += cnt[2 * k + 1];
    }
    if (counters) { counters[0] = rays; counters[1] = hits; }
    return CERES_OK;
}

static int render_sequence_impl(ceres_scene* s, size_t n_frames, const float* cams, size_t width, size_t height, unsigned flags,
                                int bytes_per_channel, void* out_all, uint64_t* counters_all);

int ceres_render_sequence(ceres_scene* s, size_t n_frames, const float* cams, size_t width, size_t height,
                          unsigned flags, float* pixels_all, uint64_t* counters_all) {
    return render_sequence_impl(s, n_frames, cams, width, height, flags, 0, pixels_all, counters_all);
}

int ceres_render_sequence_payload(ceres_scene* s, size_t n_frames, const float* cams, size_t width, size_t height, unsigned flags,
                                  int bytes_per_channel, uint8_t* payload_all, uint64_t* counters_all) {
    if (bytes_per_channel != 1 && bytes_per_channel != 2) return fail(CERES_ERR_INVALID, "ceres_render_sequence_payload: bytes_per_channel is 1 or 2");
    return render_sequence_impl(s, n_frames, cams, width, height, flags, bytes_per_channel, payload_all, counters_all);
}

static int render_sequence_impl(ceres_scene* s, size_t n_frames, const float* cams, size_t width, size_t height, unsigned flags,
                                int bpc, void* out_all, uint64_t* counters_all) {
    float* pixels_all = static_cast<float*>(out_all);
    uint8_t* payload_all = static_cast<uint8_t*>(out_all);
    if (!s || !cams || !out_all) return fail(CERES_ERR_INVALID, "ceres_render_sequence: null argument");
    // Every frame goes through the host-frame path of ceres_render: only the rectangles that hold hit pixels cross PCIe
    // (row bands copied while later bands still render) and the zeros around them are written by the host-fill pool.
    // Two frames are in flight at a time: a second set of per-frame resources (streams, record queue, staging frame,
    // band flags) over the SAME resident geometry renders the odd frames from a second host thread, so that one frame's
    // drain (copies + zero fill) overlaps the other frame's kernels.  Frames are independent (anim.cpp:82-105).
    const size_t npix = width * height;
    auto frame = [&](ceres_scene* v, size_t k) -> int {
        const float* c = cams + 13 * k;
        ceres_camera cam;
        for (int q = 0; q < 3; ++q) { cam.eye[q] = c[q]; cam.dir[q] = c[3 + q]; cam.up[q] = c[6 + q]; }
        cam.fov = c[9];
        uint64_t* cnt = counters_all ? counters_all + 2 * k : nullptr;
        if (bpc == 1) return ceres_render_ppm8(v, &cam, c + 10, width, height, flags, payload_all + k * npix * 3, cnt);
        if (bpc == 2) return ceres_render_ppm16(v, &cam, c + 10, width, height, flags, payload_all + k * npix * 6, cnt);
        return ceres_render(v, &cam, c + 10, width, height, flags, pixels_all + k * npix * 3, nullptr, cnt);
    };
    static const bool serial = std::getenv("CERES_SEQUENCE_SERIAL") != nullptr;
    if (n_frames < 2 || s->stats || s->profile || serial) {
        for (size_t k = 0; k < n_frames; ++k) if (int rc = frame(s, k)) return rc;
        return CERES_OK;
    }
    if (!s->seq_view) {
        DeviceGuard guard(s->device);
        if (int rc = scene_new(s->device, s->sm_count, &s->seq_view)) return rc;
        s->seq_view->borrowed = true;
    }
    CERES_REJECT_DURING_SEQUENCE(s, "ceres_render_sequence");
    struct InSequence {
        ceres_scene* s;
        explicit InSequence(ceres_scene* q) : s(q) { s->in_sequence.store(1, std::memory_order_release); ++tl_sequence_depth; }
        ~InSequence() { --tl_sequence_depth; s->in_sequence.store(0, std::memory_order_release); }
    } in_sequence(s);
    ceres_scene* v = s->seq_view;       // refreshed every call: a refit moves the boxes, set_tuning the knobs
    { unsigned long long* own = v->dev.check; v->dev = s->dev; v->dev.check = own; v->dev.timeline = nullptr; }
    v->d_pairs = s->d_pairs; v->d_tris = s->d_tris; v->d_norms = s->d_norms; v->d_slot_id = s->d_slot_id;
    v->scene_bytes = s->scene_bytes; v->tri_count = s->tri_count; v->node_count = s->node_count; v->n_pairs = s->n_pairs;
    v->tune = s->tune; v->shadow_blocks_per_sm = s->shadow_blocks_per_sm; v->shadow_blocks_per_sm_stats = s->shadow_blocks_per_sm_stats;
    for (int k = 0; k < 10; ++k) v->trace_blocks_per_sm[k] = s->trace_blocks_per_sm[k];
    for (int k = 0; k < 6; ++k) v->bounds[k] = s->bounds[k];
    int rc_odd = CERES_OK; std::string err_odd;
    std::thread odd([&] {
        for (size_t k = 1; k < n_frames; k += 2) if ((rc_odd = frame(v, k)) != CERES_OK) { err_odd = g_err; break; }
    });
    int rc_even = CERES_OK;
    for (size_t k = 0; k < n_frames; k += 2) if ((rc_even = frame(s, k)) != CERES_OK) break;
    odd.join();
    s->last_d2h_bytes = std::max(s->last_d2h_bytes, v->last_d2h_bytes);
    if (rc_even) return rc_even;
    if (rc_odd) return fail(rc_odd, err_odd);
    return CERES_OK;
}

int ceres_trace_rays(ceres_scene* s, const float* rays, size_t n, uint32_t* out_id, float* out_tuv) {
    if (!s || !rays || !out_id || !out_tuv) return fail(CERES_ERR_INVALID, "ceres_trace_rays: null argument");
    if (n >= (1ull << 31)) return fail(CERES_ERR_INVALID, "ceres_trace_rays: too many rays");
    if (!n) return CERES_OK;
    for (size_t k = 0; k < n; ++k)
        if (std::isnan(rays[8 * k + 6]) || std::isnan(rays[8 * k + 7]))
            return fail(CERES_ERR_INVALID, "ceres_trace_rays: tmin/tmax must not be NaN");
    DeviceGuard guard(s->device);
    float* d_rays = nullptr; uint32_t* d_id = nullptr; float* d_tuv = nullptr;
    auto cleanup = [&]() { cudaFree(d_rays); cudaFree(d_id); cudaFree(d_tuv); };
#define TRY_T(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); cleanup(); return CERES_ERR_CUDA; } } while (0)
    TRY_T(cudaMalloc((void**)&d_rays, n * 8 * sizeof(float)));
    TRY_T(cudaMalloc((void**)&d_id, n * sizeof(uint32_t)));
    TRY_T(cudaMalloc((void**)&d_tuv, n * 3 * sizeof(float)));
    TRY_T(cudaMemcpyAsync(d_rays, rays, n * 8 * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    {
        // a 64-deep stack x 256 threads = 64 KB of dynamic shared memory: above the 48 KB default, needs the opt-in
        const size_t smem = (size_t)s->dev.stack_depth * 4 * 256;
        if (int rc = smem_opt_in(trace_rays_kernel, smem)) { cleanup(); return rc; }
        trace_rays_kernel<<<(unsigned)((n + 255) / 256), 256, smem, s->stream>>>(s->dev, d_rays, (uint32_t)n, d_id, d_tuv);
    }
    g_launches++;
    TRY_T(cudaGetLastError());
    TRY_T(cudaMemcpyAsync(out_id, d_id, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    TRY_T(cudaMemcpyAsync(out_tuv, d_tuv, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    TRY_T(cudaStreamSynchronize(s->stream));
#undef TRY_T
    cleanup();
    return CERES_OK;
}

int ceres_quantize_device(ceres_scene* s, const float* d_pixels, size_t width, size_t height, void* stream_v, uint8_t* d_rgb8) {
    if (!s || !d_pixels || !d_rgb8 || !width || !height) return fail(CERES_ERR_INVALID, "ceres_quantize_device: null/zero argument");
    DeviceGuard guard(s->device);
    cudaStream_t st = stream_v ? (cudaStream_t)stream_v : s->stream;
    quantize_kernel<<<(unsigned)(s->sm_count * 8), 256, 0, st>>>(d_pixels, (uint32_t)width, (uint32_t)height, d_rgb8);
    g_launches++;
    CUDA_TRY(cudaGetLastError());
    return CERES_OK;
}

}  // extern "C"

#include "render_f64.cuh"
