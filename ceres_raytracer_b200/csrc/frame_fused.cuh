// frame_fused.cuh -- the frame kernel: ONE persistent launch renders a row range of a frame end to end (included by
// ceres_b200.cu after ray_step.cuh).
//
// Every warp of the `SMs x 8` resident CTAs runs the same loop:
//   phase 1  PRIMARY   take one 8x4-pixel tile of the conservative screen rectangle (one atomicAdd per tile), generate
//                      the 32 camera rays (render.hpp:105-113) and walk each through the reference's ordered
//                      closest-hit traversal (single_ray_traverser.hpp:65-126).  Misses store their zeros at once
//                      (render.hpp:116-117; ids: CERES_MISS); hits are ballot-compacted into the 16-byte record queue
//                      and the 32-record queue chunks they land in are credited in ready[] (a per-chunk counter).
//   phase 2  SHADOW    once the tile cursor has run out, take queue chunk c (one atomicAdd), wait until ready[c] says
//            + SHADE   its 32 records are written, rebuild the shadow rays (render.hpp:120-136), run the any-hit
//                      traversal, and shade + store the 32 pixels (render.hpp:56-84,140-150) from registers.
//            SHARE: while the 32 shadow rays of a chunk are walked, lanes whose ray has ended take over the OLDEST pending
//            subtree (the bottom entry of the traversal stack) of a lane that is still busy and walk it with a copy of
//            that ray; a hit found by any lane settles the record for all of them.  Parity-safe for shadow rays only:
//            `bool(hit)` (render.hpp:137-140) is "some leaf that passes all slab tests above it holds a triangle hit in
//            [0, FLT_MAX]" -- tmax never shrinks before the first accepted triangle, so the set of reachable leaves does
//            not depend on who walks which subtree in which order.  It turns the idle lanes of the shadow phase (14 of
//            32 active per instruction without it) into helpers and cuts the longest chains -- the tail of a frame.
// So warps that run out of tiles during the primary tail start on shadow rays while the last primary warps are still
// walking their long rays: one latency tail per frame instead of three (primary wave, shadow wave, shade pass), one
// launch instead of four, and the only pass over the framebuffer is the one that writes final values.
//
// ready[] protocol (all gpu-scope): a producer writes its records, fences, and lane 0 adds the number of records it
// put into chunk c to ready[c] (a retire spans at most two chunks).  The warp that retires the LAST tile knows the
// final record count q: it ORs READY_FINAL into the partial chunk q/32 (if any) and stores READY_FINAL into the next
// `warps` entries -- one terminating claim per warp.  A consumer claims c, polls ready[c] until it reads 32 or
// READY_FINAL|n, zeroes the entry (so the array is all-zero again when the kernel ends) and processes n records;
// n == 0 ends the warp.  A consumer only ever waits for tiles that running warps hold, so the scheme needs no
// co-residency of the whole grid (row bands of one frame run concurrently on several streams).  Waits are bounded.

#define FRAME_THREADS 128
#ifndef FRAME_MIN_CTAS
#define FRAME_MIN_CTAS 8                    // registers capped at 64 per thread: 32 warps per SM
#endif
#define READY_FINAL 0x80000000u
#define READY_COUNT 0x0000FFFFu
#define FRAME_WAIT_NS 2000000000ull
#define FRAME_ERR_READY_TIMEOUT 1u

// Zeroes columns [a, b) of local row lr (pixels: 0, ids: CERES_MISS) with the 32 lanes of one warp, 16-byte stores where the
// alignment allows.  Used by the frame kernel for the pixels of its rows that lie outside the traced rectangle.
__device__ __forceinline__ void warp_zero_cols(const Frame& f, uint32_t lr, uint32_t a, uint32_t b, uint32_t lane) {
    if (b <= a) return;
    const size_t orow = f.compact ? lr : frame_row(f, lr);
    {
        float* p = f.pixels + (orow * f.W + a) * 3ull;
        const size_t n = 3ull * (b - a);
        const size_t head = min(n, (size_t)((16u - ((uintptr_t)p & 15u)) / 4u % 4u));
        const size_t nv = (n - head) / 4;
        float4* v = reinterpret_cast<float4*>(p + head);
        for (size_t k = lane; k < nv; k += 32u) v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < head) p[lane] = 0.0f;
        const size_t t0 = head + nv * 4;
        if (t0 + lane < n) p[t0 + lane] = 0.0f;
    }
    if (f.ids) {
        uint32_t* p = f.ids + orow * f.W + a;
        const size_t n = b - a;
        const size_t head = min(n, (size_t)((16u - ((uintptr_t)p & 15u)) / 4u % 4u));
        const size_t nv = (n - head) / 4;
        uint4* v = reinterpret_cast<uint4*>(p + head);
        for (size_t k = lane; k < nv; k += 32u) v[k] = make_uint4(CERES_MISS, CERES_MISS, CERES_MISS, CERES_MISS);
        if (lane < head) p[lane] = CERES_MISS;
        const size_t t0 = head + nv * 4;
        if (t0 + lane < n) p[t0 + lane] = CERES_MISS;
    }
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// smooth_shading of render.hpp:56-84 for one lit record.  sun_line = the shadow ray's direction (render.hpp:133).
__device__ __forceinline__ void shade_lit(const DevScene& sc, const Frame& f, V3 sun_line, uint32_t i, uint32_t j, uint32_t slot,
                                          float u, float v, float& c0, float& c1, float& c2) {
    const V3 view = primary_dir(f, i, j);
    const V3 mview = vscale(view, -1.0f);                      // render.hpp:66
    const V3 h = vnormalize(vadd(sun_line, mview));            // render.hpp:53
    const float4* np = sc.norms + 3ull * slot;
    const float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2);
    const V3 nk[3] = {mk(a.x, a.y, a.z), mk(a.w, b.x, b.y), mk(b.z, b.w, c.x)};
    const float wk[3] = {u, v, sub(sub(1.0f, u), v)};
    const float amb = 0.2f;
    c0 = c1 = c2 = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // lambertian (render.hpp:46-49): abs of the left-to-right sum
        const float diffuse = mul(0.5f, fabsf(add(add(mul(sun_line.x, nk[k].x), mul(sun_line.y, nk[k].y)), mul(sun_line.z, nk[k].z))));
        const float specular = mul(0.8f, spec_pow24(vdot(nk[k], h)));
        const float ad = add(amb, diffuse);
        c0 = add(c0, mul(wk[k], clamp01(add(mul(ad, 0.5f), specular))));
        c1 = add(c1, mul(wk[k], clamp01(add(mul(ad, 0.0f), specular))));
        c2 = add(c2, mul(wk[k], clamp01(add(mul(ad, 0.8f), specular))));
    }
}

// Last CTA of a frame launch: frame counters (render.hpp:115,119,138,148,155), extents for the host path, peer flags,
// and the per-frame state re-armed for the next launch on this counter slot.
__device__ __forceinline__ void publish_frame(const Frame& f, unsigned long long hits) {
    FrameCounters* ctr = f.ctr;
    const unsigned long long occ = *(volatile uint32_t*)&ctr->occluded;
    const unsigned long long prim = (unsigned long long)f.W * (f.row_end - f.row_begin);
    const unsigned long long rays = f.shadows ? prim + hits : prim;                          // render.hpp:115,138
    const unsigned long long nh = hits + occ;                                                 // render.hpp:119,148
    // several launches (row bands, possibly on concurrent streams) add into the same frame counters; the host
    // zeroes them before the first launch of a frame
    if (f.out_counters) { atomicAdd(&f.out_counters[0], rays); atomicAdd(&f.out_counters[1], nh); }
    if (f.out_stats) {
        for (int k = 0; k < 6; ++k) atomicAdd(&f.out_stats[k], *(volatile unsigned long long*)&ctr->stats[k]);
        atomicAdd(&f.out_stats[6], rays); atomicAdd(&f.out_stats[7], nh);
    }
    const uint32_t err = *(volatile uint32_t*)&ctr->error;
    if (err && f.out_error) atomicOr(f.out_error, (unsigned long long)err);
    ctr->qcount = 0; ctr->shadow_next = 0; ctr->occluded = 0; ctr->done_blocks = 0; ctr->primary_next = 0;
    ctr->primary_done = 0; ctr->error = 0;
    for (int k = 0; k < 6; ++k) ctr->stats[k] = 0;
    if (f.host_ext) {
        // every CTA fenced its pixel stores and its extent atomics before it was counted: tell the host which
        // rectangle of this band holds anything but zeros, then raise the flag it is polling
        volatile uint32_t* e = ctr->ext;
        for (int k = 0; k < 4; ++k) f.host_ext[k] = e[k];
        __threadfence_system();
        st_release_sys(f.host_ext + 4, f.ext_epoch);
    }
    ctr->ext[0] = 0xFFFFFFFFu; ctr->ext[1] = 0u; ctr->ext[2] = 0xFFFFFFFFu; ctr->ext[3] = 0u;
    __threadfence();
    if (f.peer) {
        // every CTA fenced its stores at system scope before it was counted; publish counters, then the flag
        f.peer->counters[f.peer_rank][0] = rays;
        f.peer->counters[f.peer_rank][1] = nh;
        __threadfence_system();
        st_release_sys(&f.peer->done_epoch[f.peer_rank], f.peer_epoch);
    }
}

// Peer frames: rank 0 owns the frame and announces "epoch e has begun" when its own kernels of that epoch start --
// stream order puts that after whatever its process did with frame e-1.  The other ranks hold back their first STORE
// of epoch e until then (they trace meanwhile), so a rank that runs ahead never writes into a frame the owner is still
// reading.  Only ONE thread per GPU polls rank 0's flag over NVLink: the first warp that needs the gate elects itself
// (f.gate[1]), waits for the remote flag and relays it into f.gate[0] in local memory, where everybody else looks.
// (Every CTA polling the remote word -- 7 x 1184 pollers on one address at N = 8 -- delayed the flag by 20-90 us.)
__device__ __forceinline__ void peer_announce(const Frame& f) {          // rank 0, once per kernel
    if (blockIdx.x == 0 && threadIdx.x == 0 && (int32_t)(ld_acquire_sys(&f.peer->cleared_epoch) - f.peer_epoch) < 0)
        st_release_sys(&f.peer->cleared_epoch, f.peer_epoch);
}
__device__ __forceinline__ void peer_gate_wait(const Frame& f) {         // ranks > 0, one thread on behalf of its warp / CTA
    const uint32_t e = f.peer_epoch;
    if ((int32_t)(ld_acquire_gpu(&f.gate[0]) - e) >= 0) return;
    uint32_t seen = *(volatile uint32_t*)&f.gate[1];
    bool poller = false;
    while ((int32_t)(seen - e) < 0) {
        const uint32_t old = atomicCAS(&f.gate[1], seen, e);
        if (old == seen) { poller = true; break; }
        seen = old;
    }
    if (poller) {
        if (!peer_wait_flag(&f.peer->cleared_epoch, e)) f.peer->error = 1u;
        asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(&f.gate[0]), "r"(e) : "memory");
        return;
    }
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_gpu(&f.gate[0]) - e) < 0) {
        __nanosleep(100);
        if (global_ns() - t0 > PEER_WAIT_NS + 1000000000ull) { f.peer->error = 1u; return; }
    }
}

#ifndef FRAME_DEFAULT_CHUNK
#define FRAME_DEFAULT_CHUNK 32u
#endif
#define READY_MIN_SHIFT 3               // ready[] is sized for the smallest work item (8 records)
#ifndef FRAME_SHARE_AUTO_PX
#define FRAME_SHARE_AUTO_PX 200u        // shadow-subtree sharing is on by default below this many rectangle pixels per resident warp
#endif
#ifndef FRAME_SHARE_PERIOD
#define FRAME_SHARE_PERIOD 4u
#endif
#ifndef FRAME_SHARE_MIN_IDLE
#define FRAME_SHARE_MIN_IDLE 6u
#endif
#ifndef FRAME_DEFAULT_SHADOW_ORDER
#define FRAME_DEFAULT_SHADOW_ORDER ORDER_NEAR
#endif

template <bool STATS, int ORDER, bool SHARE, bool TL>
__global__ void __launch_bounds__(FRAME_THREADS, FRAME_MIN_CTAS) frame_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    FrameCounters* const ctr = f.ctr;
    const float4* __restrict__ pairs = sc.pairs;
    const float4* __restrict__ tris = sc.tris;
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);
    const V3 eye = mk(f.eye[0], f.eye[1], f.eye[2]);
    unsigned long long st[6] = {0, 0, 0, 0, 0, 0};   // STATS: S_miss,T_miss,S_hit,T_hit,S_shadow,T_shadow
    unsigned long long tl_begin = 0, tl_primary = 0;  // sc.timeline: per-warp timestamps
    unsigned long long tl_tile_max = 0, tl_tile_last = 0, tl_chunk_max = 0, tl_wait = 0, tl_t = 0;
    uint32_t tl_tile_max_base = 0, tl_chunk_max_id = 0;
    uint32_t my_chunks = 0;
    if (TL) tl_begin = global_ns();
    unsigned long long tl_gate = 0;
    bool gated = f.peer != nullptr && f.peer_rank != 0u;      // a peer's first store of the epoch waits for rank 0's announcement
    if (f.peer && f.peer_rank == 0u) peer_announce(f);

    // =================================================================== phase 0: zeros outside the rectangle
    // (render.hpp:116-117 for the pixels that are never traced.)  Quarter rows are dealt out to the warps of the grid; the stores
    // are fire-and-forget, so the memory system drains them while the same warps start tracing.
    if (f.clear_rows) {
        const uint32_t n_warps = gridDim.x * (FRAME_THREADS / 32), w = blockIdx.x * (FRAME_THREADS / 32) + (tid >> 5);
        const uint32_t n_items = (f.row_end - f.row_begin) * 4u, q = (f.W + 3u) / 4u;
        for (uint32_t it = w; it < n_items; it += n_warps) {
            const uint32_t lr = f.row_begin + (it >> 2), a = (it & 3u) * q, b = min(f.W, a + q);
            if (lr >= f.rect_row0 && lr < f.rect_row1 && f.rect_x0 < f.rect_x1) {
                warp_zero_cols(f, lr, a, min(b, max(a, f.rect_x0)), lane);
                warp_zero_cols(f, lr, max(a, min(b, f.rect_x1)), b, lane);
            } else warp_zero_cols(f, lr, a, b, lane);
        }
    }

    // =================================================================== phase 1: primary tiles
    const uint32_t total = f.work_items;            // 32 per tile
    uint32_t my_tiles = 0;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&ctr->primary_next, 32u);
        base = __shfl_sync(FULL, base, 0);
        if (base >= total) break;
        ++my_tiles;
        if (TL) tl_t = tl_tile_last = global_ns();
        uint32_t i, lrow, n_steps = 0, n_tests = 0;
        const bool have = item_pixel(f, base + lane, i, lrow);
        const uint32_t j = frame_row(f, lrow);
        RaySlot r;
        r.found = false; r.slot = 0; r.u = r.v = 0.0f;
        if (have) {
            ray_begin(r, eye, primary_dir(f, i, j), tid);
            if (sc.root_leaf_count) {
                // the root itself is a leaf (single_ray_traverser.hpp:72-73): its triangles are the whole scene
                const V3 o = mk(r.ox, r.oy, r.oz), d = mk(r.dx, r.dy, r.dz);
                for (uint32_t s = 0; s < sc.root_leaf_count; ++s) {
                    if (STATS) n_tests++;
                    float t, u, v;
                    if (tri_test(load_tri(tris, s), o, d, 0.0f, r.tmax, t, u, v)) { r.slot = s; r.u = u; r.v = v; r.tmax = t; r.found = true; }
                }
            } else {
                bool active = true;
                while (active) active = ray_step<false, STATS, FRAME_THREADS, ORDER_NEAR>(pairs, tris, r, smem_stack, tid, n_steps, n_tests, sc);
            }
        }
        __syncwarp();
        const bool found = have && r.found;
        const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
        if (gated) {
            if (lane == 0) peer_gate_wait(f);
            __syncwarp();
            gated = false;
            if (TL) tl_gate = global_ns();
        }
        if (have && !found) {   // render.hpp:116-117
            CERES_CK((f.compact ? lrow : j) < f.pix_rows && i < f.W, CHECK_PIXEL_INDEX, sc);
            float* p = f.pixels + 3 * opix;
            p[0] = 0.0f; p[1] = 0.0f; p[2] = 0.0f;
            if (f.ids) f.ids[opix] = CERES_MISS;
        }
        const unsigned m = __ballot_sync(FULL, found);
        if (m) {
            const uint32_t n = (uint32_t)__popc(m);
            uint32_t qb = 0;
            if (lane == 0) qb = atomicAdd(&ctr->qcount, n);
            qb = __shfl_sync(FULL, qb, 0);
            if (found) {
                // pixel is stored as (local row, column): the shade step rebuilds both the frame row and the output position
                uint32_t at = qb + __popc(m & lt_mask);
                CERES_CLAMP(at, f.queue_cap, CHECK_QUEUE_INDEX, sc);
                f.queue[at] = make_uint4(lrow * f.W + i, r.slot, __float_as_uint(r.u), __float_as_uint(r.v));
                __threadfence();
            }
            __syncwarp();
            {   // credit the work items (1 << chunk_shift records each) these n records fall into: one lane per item
                const uint32_t sh = f.chunk_shift, e = (qb >> sh) + lane;
                const uint32_t lo = max(qb, e << sh), hi = min(qb + n, (e + 1u) << sh);
                if (hi > lo) { CERES_CK(e < f.ready_cap, CHECK_READY_INDEX, sc); atomicAdd(&f.ready[e], hi - lo); }
            }
        }
        if (TL) { const unsigned long long d = global_ns() - tl_t; if (d > tl_tile_max) { tl_tile_max = d; tl_tile_max_base = base; } }
        if (STATS && have) {
            st[found ? 2 : 0] += n_steps; st[found ? 3 : 1] += n_tests;
            if (sc.hist) atomicAdd(&sc.hist[min(n_steps >> 2, 63u)], 1ull);
        }
    }
    {
        // the warp that retires the last tile closes the queue
        const uint32_t n_tiles = total >> 5;
        bool fin = false;
        if (lane == 0) {
            if (my_tiles) { __threadfence(); fin = atomicAdd(&ctr->primary_done, my_tiles) + my_tiles == n_tiles; }
            else if (n_tiles == 0u) fin = blockIdx.x == 0 && tid == 0;
        }
        fin = __shfl_sync(FULL, fin, 0);
        if (fin) {
            __threadfence();
            uint32_t q = 0;
            if (lane == 0) q = atomicAdd(&ctr->qcount, 0u);
            q = __shfl_sync(FULL, q, 0);
            const uint32_t c_full = q >> f.chunk_shift, rem = q & ((1u << f.chunk_shift) - 1u), c_tot = c_full + (rem ? 1u : 0u);
            const uint32_t n_warps = gridDim.x * (FRAME_THREADS / 32);
            CERES_CK(c_tot + n_warps <= f.ready_cap, CHECK_READY_INDEX, sc);
            if (rem && lane == 0) atomicOr(&f.ready[c_full], READY_FINAL);
            for (uint32_t k = lane; k < n_warps; k += 32u) atomicExch(&f.ready[c_tot + k], READY_FINAL);
        }
    }

    if (TL) tl_primary = global_ns();

    // =================================================================== phase 2: shadow rays + shading, 32 records at a time
    uint32_t occl = 0;
    uint32_t e_i0 = 0xFFFFFFFFu, e_i1 = 0u, e_r0 = 0xFFFFFFFFu, e_r1 = 0u;
    for (;;) {
        uint32_t c = 0, v = 0;
        if (TL) tl_t = global_ns();
        if (lane == 0) {
            c = atomicAdd(&ctr->shadow_next, 1u);
            CERES_CLAMP(c, f.ready_cap, CHECK_READY_INDEX, sc);
            unsigned ns = 32;
            unsigned long long t0 = 0;
            for (;;) {
                v = ld_acquire_gpu(&f.ready[c]);
                if (v == (1u << f.chunk_shift) || (v & READY_FINAL)) break;
                __nanosleep(ns);
                if (ns < 1024u) ns <<= 1;
                else if (t0 == 0) t0 = global_ns();
                else if (global_ns() - t0 > FRAME_WAIT_NS) { atomicOr(&ctr->error, FRAME_ERR_READY_TIMEOUT); v = READY_FINAL; break; }
            }
            *(volatile uint32_t*)&f.ready[c] = 0u;
        }
        c = __shfl_sync(FULL, c, 0);
        v = __shfl_sync(FULL, v, 0);
        const uint32_t n = v & READY_COUNT;
        if (TL) { const unsigned long long t = global_ns(); tl_wait += t - tl_t; tl_t = t; }
        if (n == 0u) break;
        if (gated) {
            if (lane == 0) peer_gate_wait(f);
            __syncwarp();
            gated = false;
            if (TL) tl_gate = global_ns();
        }
        ++my_chunks;
        if (SHARE && f.shadows && !sc.root_leaf_count) {
            // ---- the chunk's shadow rays, walked by all 32 lanes together (see SHARE above)
            uint32_t my_rec = lane;             // the record whose ray this lane is walking
            uint32_t occ_mask = 0;              // warp-uniform: records known to be occluded
            uint32_t n_steps = 0, n_tests = 0;
            RaySlot r;
            bool active = lane < n;
            if (active) {
                CERES_CK(((size_t)c << f.chunk_shift) + lane < f.queue_cap, CHECK_QUEUE_INDEX, sc);
                const uint4 rec = __ldcg(f.queue + ((size_t)c << f.chunk_shift) + lane);
                CERES_CK(rec.y < sc.n_slots, CHECK_SLOT_INDEX, sc);
                V3 P, sun_line;
                shadow_setup(load_tri(tris, rec.y), __uint_as_float(rec.z), __uint_as_float(rec.w), sun, P, sun_line);
                ray_begin(r, P, sun_line, tid);
            } else {
                r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.ix = r.iy = r.iz = r.sx = r.sy = r.sz = 0.0f;
                r.tmax = FLT_MAX; r.u = r.v = 0.0f; r.slot = 0; r.cur = 0; r.sp = r.sb = tid; r.found = false;
            }
            const uint32_t period = f.share_period, min_idle = f.share_min_idle;
            for (;;) {
                for (uint32_t k = 0; k < period && active; ++k)
                    active = ray_step<true, STATS, FRAME_THREADS, ORDER>(pairs, tris, r, smem_stack, tid, n_steps, n_tests, sc);
                // a lane that ended on a hit settles its record for every lane working on it
                occ_mask |= __reduce_or_sync(FULL, (!active && r.found) ? 1u << my_rec : 0u);
                if ((occ_mask >> my_rec) & 1u) active = false;
                const unsigned act = __ballot_sync(FULL, active);
                if (!act) break;
                if ((uint32_t)__popc(~act) < min_idle) continue;
                const unsigned donors = __ballot_sync(FULL, active && r.sp != r.sb);
                if (!donors) continue;
                // k-th idle lane <- oldest pending subtree of the k-th donor
                const uint32_t n_idle = (uint32_t)__popc(~act), n_don = (uint32_t)__popc(donors);
                const bool thief = !active && (uint32_t)__popc(~act & lt_mask) < n_don;
                const bool donor = active && r.sp != r.sb && (uint32_t)__popc(donors & lt_mask) < n_idle;
                uint32_t give = 0;
                if (donor) { CERES_CK(r.sb >= tid && r.sb < r.sp, CHECK_STACK_UNDER, sc); give = smem_stack[r.sb]; r.sb += FRAME_THREADS; }
                const uint32_t src = thief ? __fns(donors, 0u, __popc(~act & lt_mask) + 1) : lane;
                give = __shfl_sync(FULL, give, src);
                const uint32_t rec_src = __shfl_sync(FULL, my_rec, src);
                const float ox = __shfl_sync(FULL, r.ox, src), oy = __shfl_sync(FULL, r.oy, src), oz = __shfl_sync(FULL, r.oz, src);
                const float dx = __shfl_sync(FULL, r.dx, src), dy = __shfl_sync(FULL, r.dy, src), dz = __shfl_sync(FULL, r.dz, src);
                const float ix = __shfl_sync(FULL, r.ix, src), iy = __shfl_sync(FULL, r.iy, src), iz = __shfl_sync(FULL, r.iz, src);
                const float sx = __shfl_sync(FULL, r.sx, src), sy = __shfl_sync(FULL, r.sy, src), sz = __shfl_sync(FULL, r.sz, src);
                if (thief) {
                    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz;
                    r.ix = ix; r.iy = iy; r.iz = iz; r.sx = sx; r.sy = sy; r.sz = sz;
                    r.cur = give; r.sp = r.sb = tid; r.found = false; r.tmax = FLT_MAX;
                    my_rec = rec_src; active = true;
                }
            }
            if (STATS) { st[4] += n_steps; st[5] += n_tests; }
            if (lane < n) {
                const uint4 rec = __ldcg(f.queue + ((size_t)c << f.chunk_shift) + lane);
                const uint32_t lrow = rec.x / f.W, i = rec.x - lrow * f.W;
                const uint32_t j = frame_row(f, lrow);
                const uint32_t slot = rec.y;
                float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
                if ((occ_mask >> lane) & 1u) occl++;
                else {
                    const float hu = __uint_as_float(rec.z), hv = __uint_as_float(rec.w);
                    V3 P, sun_line;
                    shadow_setup(load_tri(tris, slot), hu, hv, sun, P, sun_line);
                    shade_lit(sc, f, sun_line, i, j, slot, hu, hv, c0, c1, c2);
                }
                const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
                CERES_CK((f.compact ? lrow : j) < f.pix_rows && i < f.W && slot < sc.n_slots, CHECK_PIXEL_INDEX, sc);
                float* p = f.pixels + 3 * opix;
                p[0] = c0; p[1] = c1; p[2] = c2;
                if (f.ids) f.ids[opix] = __ldg(sc.slot_id + slot);
                e_i0 = min(e_i0, i); e_i1 = max(e_i1, i); e_r0 = min(e_r0, lrow); e_r1 = max(e_r1, lrow);
            }
        } else if (lane < n) {
            const uint4 rec = __ldcg(f.queue + ((size_t)c << f.chunk_shift) + lane);
            const uint32_t lrow = rec.x / f.W, i = rec.x - lrow * f.W;
            const uint32_t j = frame_row(f, lrow);
            const uint32_t slot = rec.y;
            const float hu = __uint_as_float(rec.z), hv = __uint_as_float(rec.w);
            V3 P, sun_line;
            shadow_setup(load_tri(tris, slot), hu, hv, sun, P, sun_line);
            bool occluded = false;
            if (f.shadows) {
                uint32_t n_steps = 0, n_tests = 0;
                RaySlot r;
                ray_begin(r, P, sun_line, tid);
                if (sc.root_leaf_count) {
                    for (uint32_t s = 0; s < sc.root_leaf_count && !r.found; ++s) {
                        if (STATS) n_tests++;
                        float t, u, w;
                        if (tri_test(load_tri(tris, s), P, sun_line, 0.0f, r.tmax, t, u, w)) r.found = true;
                    }
                } else {
                    bool active = true;
                    while (active) active = ray_step<true, STATS, FRAME_THREADS, ORDER>(pairs, tris, r, smem_stack, tid, n_steps, n_tests, sc);
                }
                occluded = r.found;
                if (occluded) occl++;
                if (STATS) {
                    st[4] += n_steps; st[5] += n_tests;
                    if (sc.hist) atomicAdd(&sc.hist[64u + min(n_steps >> 2, 63u)], 1ull);
                }
            }
            float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
            if (!occluded) shade_lit(sc, f, sun_line, i, j, slot, hu, hv, c0, c1, c2);
            const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
            CERES_CK((f.compact ? lrow : j) < f.pix_rows && i < f.W && slot < sc.n_slots, CHECK_PIXEL_INDEX, sc);
            float* p = f.pixels + 3 * opix;
            p[0] = c0; p[1] = c1; p[2] = c2;
            if (f.ids) f.ids[opix] = __ldg(sc.slot_id + slot);
            e_i0 = min(e_i0, i); e_i1 = max(e_i1, i); e_r0 = min(e_r0, lrow); e_r1 = max(e_r1, lrow);
        }
        __syncwarp();
        if (TL) { const unsigned long long d = global_ns() - tl_t; if (d > tl_chunk_max) { tl_chunk_max = d; tl_chunk_max_id = c; } }
        // no early exit after the partial last chunk: every warp consumes exactly one terminating entry, which is what
        // leaves ready[] all-zero for the next launch
    }

    // =================================================================== per-warp sums, last CTA publishes the frame
    for (int o = 16; o; o >>= 1) occl += __shfl_xor_sync(FULL, occl, o);
    if (lane == 0 && occl) atomicAdd(&ctr->occluded, occl);
    if (STATS) {
        for (int k = 0; k < 6; ++k)
            for (int o = 16; o; o >>= 1) st[k] += __shfl_xor_sync(FULL, st[k], o);
        if (lane == 0) for (int k = 0; k < 6; ++k) if (st[k]) atomicAdd(&ctr->stats[k], st[k]);
    }
    if (TL && lane == 0) {
        unsigned long long* t = sc.timeline + (unsigned long long)CERES_TIMELINE_WORDS * (blockIdx.x * (FRAME_THREADS / 32) + (tid >> 5));
        t[0] = tl_begin; t[1] = tl_primary; t[2] = global_ns(); t[3] = my_tiles; t[4] = my_chunks;
        t[5] = tl_tile_max; t[6] = tl_tile_max_base; t[7] = tl_tile_last; t[8] = tl_chunk_max; t[9] = tl_chunk_max_id; t[10] = tl_wait; t[11] = tl_gate;
    }
    if (f.host_ext) {        // uniform branch: extents of the hit pixels this warp wrote
        e_i0 = __reduce_min_sync(FULL, e_i0); e_i1 = __reduce_max_sync(FULL, e_i1);
        e_r0 = __reduce_min_sync(FULL, e_r0); e_r1 = __reduce_max_sync(FULL, e_r1);
        if (lane == 0 && e_i0 <= e_i1) {
            atomicMin(&ctr->ext[0], e_i0); atomicMax(&ctr->ext[1], e_i1); atomicMin(&ctr->ext[2], e_r0); atomicMax(&ctr->ext[3], e_r1);
        }
    }
    __shared__ bool last;
    if (f.peer) __threadfence_system();     // my pixel stores (NVLink) are performed before the CTA is counted done
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        last = atomicAdd(&ctr->done_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    if (tid == 0) {
        __threadfence();
        publish_frame(f, *(volatile uint32_t*)&ctr->qcount);
    }
    if (f.peer && f.peer_wait_world) {
        // rank 0 of a peer frame: the frame is complete when every rank's done flag has reached this epoch
        __syncthreads();
        if (tid < 32u) {
            unsigned long long rays = 0, hits = 0;
            if (tid < f.peer_wait_world) {
                if (!peer_wait_flag(&f.peer->done_epoch[tid], f.peer_epoch)) f.peer->error = 2u;
                const volatile unsigned long long* pc = f.peer->counters[tid];
                rays = pc[0]; hits = pc[1];
            }
            for (int o = 16; o; o >>= 1) { rays += __shfl_xor_sync(FULL, rays, o); hits += __shfl_xor_sync(FULL, hits, o); }
            if (tid == 0 && f.peer_out) { f.peer_out[0] = rays; f.peer_out[1] = hits; }
        }
    }
}

// Zeroes the pixels of local rows [row_a, row_b) x columns [x_a, x_b) that lie OUTSIDE the hole (local rows
// [hole_r0, hole_r1) x columns [hole_x0, hole_x1)); ids get CERES_MISS.  The hole is the rectangle the frame kernel
// writes completely.  One CTA per row at a time, 16-byte stores where the alignment allows.
struct ClearRegion { uint32_t row_a, row_b, x_a, x_b, hole_r0, hole_r1, hole_x0, hole_x1; };

__device__ __forceinline__ void clear_span(float* __restrict__ p, size_t n) {       // n floats from p, whole CTA
    const size_t head = min(n, (size_t)((16u - ((uintptr_t)p & 15u)) / 4u % 4u));
    const size_t nv = (n - head) / 4;
    float4* v = reinterpret_cast<float4*>(p + head);
    for (size_t k = threadIdx.x; k < nv; k += blockDim.x) v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t k = threadIdx.x; k < head; k += blockDim.x) p[k] = 0.0f;
    for (size_t k = head + nv * 4 + threadIdx.x; k < n; k += blockDim.x) p[k] = 0.0f;
}

__global__ void __launch_bounds__(256) clear_outside_kernel(const Frame f, const ClearRegion q) {
    if (f.peer) {
        if (f.peer_rank == 0u) peer_announce(f);
        else { if (threadIdx.x == 0) peer_gate_wait(f); __syncthreads(); }
    }
    for (uint32_t lr = q.row_a + blockIdx.x; lr < q.row_b; lr += gridDim.x) {
        const size_t orow = f.compact ? lr : frame_row(f, lr);
        float* px = f.pixels + orow * 3ull * f.W;
        uint32_t* id = f.ids ? f.ids + orow * (size_t)f.W : nullptr;
        const bool holed = lr >= q.hole_r0 && lr < q.hole_r1 && q.hole_x0 < q.hole_x1;
        // up to two column spans of this row: [x_a, s0) and [s1, x_b)
        const uint32_t s0 = holed ? min(max(q.hole_x0, q.x_a), q.x_b) : q.x_b;
        const uint32_t s1 = holed ? max(min(q.hole_x1, q.x_b), q.x_a) : q.x_b;
        if (s0 > q.x_a) {
            clear_span(px + 3ull * q.x_a, 3ull * (s0 - q.x_a));
            if (id) for (uint32_t k = q.x_a + threadIdx.x; k < s0; k += blockDim.x) id[k] = CERES_MISS;
        }
        if (q.x_b > s1) {
            clear_span(px + 3ull * s1, 3ull * (q.x_b - s1));
            if (id) for (uint32_t k = s1 + threadIdx.x; k < q.x_b; k += blockDim.x) id[k] = CERES_MISS;
        }
    }
}
