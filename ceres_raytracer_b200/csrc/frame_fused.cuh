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
// stream order puts that after whatever its process did with frame e-1.  The other ranks hold back their first store
// of epoch e until then, so a rank that runs ahead never writes into a frame the owner is still reading.
__device__ __forceinline__ void peer_begin_gate(const Frame& f) {
    if (f.peer_rank == 0u) {
        if (blockIdx.x == 0 && threadIdx.x == 0 && (int32_t)(ld_acquire_sys(&f.peer->cleared_epoch) - f.peer_epoch) < 0)
            st_release_sys(&f.peer->cleared_epoch, f.peer_epoch);
    } else {
        if (threadIdx.x == 0 && !peer_wait_flag(&f.peer->cleared_epoch, f.peer_epoch)) f.peer->error = 1u;
        __syncthreads();
    }
}

#ifndef FRAME_DEFAULT_SHADOW_ORDER
#define FRAME_DEFAULT_SHADOW_ORDER ORDER_NEAR
#endif

template <bool STATS, int ORDER>
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
    if (f.peer) peer_begin_gate(f);

    // =================================================================== phase 1: primary tiles
    const uint32_t total = f.work_items;            // 32 per tile
    uint32_t my_tiles = 0;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&ctr->primary_next, 32u);
        base = __shfl_sync(FULL, base, 0);
        if (base >= total) break;
        ++my_tiles;
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
                while (active) active = ray_step<false, STATS, FRAME_THREADS, ORDER_NEAR>(pairs, tris, r, smem_stack, tid, n_steps, n_tests);
            }
        }
        __syncwarp();
        const bool found = have && r.found;
        const size_t opix = (size_t)(f.compact ? lrow : j) * f.W + i;
        if (have && !found) {   // render.hpp:116-117
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
                f.queue[qb + __popc(m & lt_mask)] = make_uint4(lrow * f.W + i, r.slot, __float_as_uint(r.u), __float_as_uint(r.v));
                __threadfence();
            }
            __syncwarp();
            if (lane == 0) {
                const uint32_t n0 = min(n, 32u - (qb & 31u));
                atomicAdd(&f.ready[qb >> 5], n0);
                if (n > n0) atomicAdd(&f.ready[(qb >> 5) + 1u], n - n0);
            }
        }
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
            const uint32_t c_full = q >> 5, rem = q & 31u, c_tot = c_full + (rem ? 1u : 0u);
            const uint32_t n_warps = gridDim.x * (FRAME_THREADS / 32);
            if (rem && lane == 0) atomicOr(&f.ready[c_full], READY_FINAL);
            for (uint32_t k = lane; k < n_warps; k += 32u) atomicExch(&f.ready[c_tot + k], READY_FINAL);
        }
    }

    // =================================================================== phase 2: shadow rays + shading, 32 records at a time
    uint32_t occl = 0;
    uint32_t e_i0 = 0xFFFFFFFFu, e_i1 = 0u, e_r0 = 0xFFFFFFFFu, e_r1 = 0u;
    for (;;) {
        uint32_t c = 0, v = 0;
        if (lane == 0) {
            c = atomicAdd(&ctr->shadow_next, 1u);
            unsigned ns = 32;
            unsigned long long t0 = 0;
            for (;;) {
                v = ld_acquire_gpu(&f.ready[c]);
                if (v == 32u || (v & READY_FINAL)) break;
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
        if (n == 0u) break;
        if (lane < n) {
            const uint4 rec = __ldcg(f.queue + (size_t)c * 32u + lane);
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
                    while (active) active = ray_step<true, STATS, FRAME_THREADS, ORDER>(pairs, tris, r, smem_stack, tid, n_steps, n_tests);
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
            float* p = f.pixels + 3 * opix;
            p[0] = c0; p[1] = c1; p[2] = c2;
            if (f.ids) f.ids[opix] = __ldg(sc.slot_id + slot);
            e_i0 = min(e_i0, i); e_i1 = max(e_i1, i); e_r0 = min(e_r0, lrow); e_r1 = max(e_r1, lrow);
        }
        __syncwarp();
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
    if (f.peer) peer_begin_gate(f);
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
