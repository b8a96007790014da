// trace_waves.cuh (experiments; the round-1 default, now tuning variant 6) -- persistent-warp, lane-refilling traversal kernel (included by ceres_b200.cu after the
// device math helpers).
//
// One kernel template serves both ray waves of a frame:
//   WAVE_PRIMARY  work item k = pixel k of the culled tile rectangle (8x4-pixel tiles, 32 consecutive items = one
//                 tile, so a freshly filled warp is as coherent as a screen tile); closest-hit traversal; hits are
//                 ballot-compacted into the record queue.  Misses write NOTHING: the framebuffer was cleared by
//                 clear_kernel, and pixels outside the conservative screen rectangle of the scene are never traced.
//   WAVE_SHADOW   work item k = queue record k; the shadow ray of render.hpp:120-136, any-hit traversal; an occluded
//                 record gets its OCCLUDED bit set.
// Scheduling: `SMs x occupancy` CTAs stay resident; every warp owns 32 ray slots.  When at least REFILL_MIN slots of
// a warp are idle it retires the finished rays (one atomic per warp for the queue append), grabs as many new work
// items as it has idle lanes with ONE atomicAdd on the global cursor, and sets the new rays up in those lanes while
// the other lanes keep their rays.  SIMD lanes therefore stay busy although ray lengths differ wildly (1 to 127
// node-pair steps on the dragon).  The per-ray semantics are untouched: every lane walks its own ray through
// exactly the reference's sequence of slab tests and triangle tests (see traverse() in ceres_b200.cu for the
// citations), so the hit ids stay bit-exact.
//
// The traversal stack lives in shared memory, one column per thread (word index = depth * blockDim + tid: bank =
// lane, conflict-free), addressed through a running word offset so that a push/pop is one STS/LDS + one add.

#define WAVE_PRIMARY 0
#define WAVE_SHADOW 1
#define TRACE_THREADS 128
#define TRACE_CHUNK 32u
#ifndef TRACE_MIN_CTAS
#define TRACE_MIN_CTAS(T) (1024 / (T))      // registers capped at 64 per thread: 32 warps per SM
#endif
#define REFILL_MIN 32u       // default: refill a warp only when all 32 ray slots are idle (coherence wins)

// RaySlot / ray_begin / ray_step / item_pixel: ../ray_step.cuh

template <int WAVE, bool STATS, int THREADS>
__global__ void __launch_bounds__(THREADS, TRACE_MIN_CTAS(THREADS)) trace_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t total = WAVE == WAVE_PRIMARY ? f.work_items : f.ctr->qcount;
    uint32_t* cursor = WAVE == WAVE_PRIMARY ? &f.ctr->primary_next : &f.ctr->shadow_next;
    const float4* __restrict__ pairs = sc.pairs;
    const float4* __restrict__ tris = sc.tris;
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);
    const V3 eye = mk(f.eye[0], f.eye[1], f.eye[2]);

    RaySlot r;
    r.found = false; r.item = 0; r.sp = tid; r.sb = tid; r.cur = 0; r.tmax = 0.0f; r.u = r.v = 0.0f; r.slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.ix = r.iy = r.iz = r.sx = r.sy = r.sz = 0.0f;
    bool active = false;        // this lane holds a ray in flight
    bool retire = false;        // this lane holds a finished ray whose result is not yet published
    bool exhausted = false;     // warp-uniform: the cursor ran past the last work item
    uint32_t chunk_next = 0, chunk_end = 0;     // warp-uniform: this warp's private range of work items
    uint32_t n_steps = 0, n_tests = 0, occl = 0;
    unsigned long long st[4] = {0, 0, 0, 0};   // STATS: S_miss,T_miss,S_hit,T_hit (primary) / S,T in [0],[1] (shadow)

    for (;;) {
        // ---- retire finished rays, refill idle lanes --------------------------------------------------------
        if (WAVE == WAVE_PRIMARY) {
            const bool out = retire && r.found;
            const unsigned m = __ballot_sync(0xFFFFFFFFu, out);
            if (m) {
                const int leader = __ffs(m) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(&f.ctr->qcount, (uint32_t)__popc(m));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (out) {
                    uint32_t i, lrow;
                    item_pixel(f, r.item, i, lrow);
                    f.queue[base + __popc(m & lt_mask)] = make_uint4(lrow * f.W + i, r.slot, __float_as_uint(r.u), __float_as_uint(r.v));
                }
            }
        }
        if (STATS && retire) {
            if (WAVE == WAVE_PRIMARY) { st[r.found ? 2 : 0] += n_steps; st[r.found ? 3 : 1] += n_tests; }
            else { st[0] += n_steps; st[1] += n_tests; }
            if (sc.hist) atomicAdd(&sc.hist[(WAVE == WAVE_SHADOW ? 64u : 0u) + min(n_steps >> 2, 63u)], 1ull);
            n_steps = n_tests = 0;
        }
        retire = false;
        if (!exhausted) {
            // The warp owns a private chunk [chunk_next, chunk_end) of consecutive work items (neighbouring pixels /
            // records) and refills idle lanes from it, so that replacement rays stay coherent with the rays still in
            // flight; a new chunk costs one atomicAdd on the global cursor.
            const unsigned idle = ~__ballot_sync(0xFFFFFFFFu, active);
            if (chunk_next >= chunk_end) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(cursor, f.chunk);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                chunk_next = base < total ? base : total;
                chunk_end = base + f.chunk < total ? base + f.chunk : total;
            }
            const uint32_t avail = chunk_end - chunk_next;
            const uint32_t rank = __popc(idle & lt_mask);
            const uint32_t item = chunk_next + rank;
            const bool take = !active && rank < avail;
            chunk_next += min(avail, (uint32_t)__popc(idle));
            exhausted = chunk_next >= total;       // own chunk empty and the cursor is past the end
            if (take) {
                r.item = item;
                bool have = true;
                if (WAVE == WAVE_PRIMARY) {
                    uint32_t i, lrow;
                    have = item_pixel(f, item, i, lrow);
                    if (have) ray_begin(r, eye, primary_dir(f, i, frame_row(f, lrow)), tid);
                } else {
                    const uint4 rec = f.queue[item];
                    const Tri tr = load_tri(tris, rec.y);
                    V3 P, sun_line;
                    shadow_setup(tr, __uint_as_float(rec.z), __uint_as_float(rec.w), sun, P, sun_line);
                    ray_begin(r, P, sun_line, tid);
                }
                if (have) {
                    if (sc.root_leaf_count) {
                        // the root itself is a leaf (single_ray_traverser.hpp:72-73): its triangles are the whole scene
                        const V3 o = mk(r.ox, r.oy, r.oz), d = mk(r.dx, r.dy, r.dz);
                        for (uint32_t s = 0; s < sc.root_leaf_count; ++s) {
                            if (STATS) n_tests++;
                            float t, u, v;
                            if (tri_test(load_tri(tris, s), o, d, 0.0f, r.tmax, t, u, v)) {
                                r.slot = s; r.u = u; r.v = v; r.tmax = t; r.found = true;
                                if (WAVE == WAVE_SHADOW) break;
                            }
                        }
                        retire = true;
                        if (WAVE == WAVE_SHADOW && r.found) { atomicOr(&f.queue[item].y, OCCLUDED_BIT); occl++; }
                    } else {
                        active = true;
                    }
                }
            }
        }
        const unsigned act = __ballot_sync(0xFFFFFFFFu, active);
        if (!act) {
            if (exhausted && !__any_sync(0xFFFFFFFFu, retire)) break;
            continue;
        }
        // ---- trace until enough lanes have gone idle to make a refill worthwhile -----------------------------------
        if (f.refill_min >= 32u || exhausted) {
            // whole-warp refill: every lane runs its ray to the end, the warp reconverges after the loop
            const bool had_ray = active;
            while (active) {
                active = ray_step<WAVE == WAVE_SHADOW, STATS, THREADS>(pairs, tris, r, smem_stack, tid, n_steps, n_tests, sc);
            }
            if (had_ray) {
                retire = true;
                if (WAVE == WAVE_SHADOW && r.found) { atomicOr(&f.queue[r.item].y, OCCLUDED_BIT); occl++; }
            }
        } else {
            const uint32_t keep_going = 32u - f.refill_min;    // leave the loop when active lanes <= this
            for (;;) {
                if (active) {
                    active = ray_step<WAVE == WAVE_SHADOW, STATS, THREADS>(pairs, tris, r, smem_stack, tid, n_steps, n_tests, sc);
                    if (!active) {
                        retire = true;
                        if (WAVE == WAVE_SHADOW && r.found) { atomicOr(&f.queue[r.item].y, OCCLUDED_BIT); occl++; }
                    }
                }
                const unsigned a = __ballot_sync(0xFFFFFFFFu, active);
                if ((uint32_t)__popc(a) <= keep_going) break;
            }
        }
    }

    if (WAVE == WAVE_SHADOW) {
        for (int o = 16; o; o >>= 1) occl += __shfl_xor_sync(0xFFFFFFFFu, occl, o);
        if (lane == 0 && occl) atomicAdd(&f.ctr->occluded, occl);
    }
    if (STATS) {
        for (int k = 0; k < 4; ++k)
            for (int o = 16; o; o >>= 1) st[k] += __shfl_xor_sync(0xFFFFFFFFu, st[k], o);
        if (lane == 0) {
            if (WAVE == WAVE_PRIMARY) { for (int k = 0; k < 4; ++k) atomicAdd(&f.ctr->stats[k], st[k]); }
            else { atomicAdd(&f.ctr->stats[4], st[0]); atomicAdd(&f.ctr->stats[5], st[1]); }
        }
    }
}

// Zeroes the launch's rows of the framebuffer (render.hpp:116-117,149: every pixel that is not a lit hit is 0) and
// marks the ids CERES_MISS; shade_kernel then overwrites the hit pixels.  16-byte stores where the row start allows.
__global__ void __launch_bounds__(256) clear_kernel(const Frame f) {
    const uint32_t rows = f.row_end - f.row_begin;
    const size_t row_floats = 3ull * f.W;
    if (f.compact || f.shard_count <= 1) {
        // the launch's rows are contiguous in the output
        float* base = f.pixels + (size_t)f.row_begin * row_floats;
        const size_t n = (size_t)rows * row_floats;
        const size_t head = (16u - ((uintptr_t)base & 15u)) / 4u % 4u;      // floats up to the first 16-B boundary
        const size_t nv = n > head ? (n - head) / 4 : 0;
        float4* v = reinterpret_cast<float4*>(base + head);
        for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < nv; k += (size_t)gridDim.x * 256u) v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (blockIdx.x == 0) {
            for (size_t k = threadIdx.x; k < head && k < n; k += 256u) base[k] = 0.0f;
            for (size_t k = head + nv * 4 + threadIdx.x; k < n; k += 256u) base[k] = 0.0f;
        }
        if (f.ids) {
            uint32_t* ib = f.ids + (size_t)f.row_begin * f.W;
            const size_t ni = (size_t)rows * f.W;
            for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < ni; k += (size_t)gridDim.x * 256u) ib[k] = CERES_MISS;
        }
    } else {
        for (uint32_t lr = f.row_begin + blockIdx.x; lr < f.row_end; lr += gridDim.x) {
            const size_t j = frame_row(f, lr);
            float* p = f.pixels + j * row_floats;
            for (size_t k = threadIdx.x; k < row_floats; k += 256u) p[k] = 0.0f;
            if (f.ids) for (uint32_t k = threadIdx.x; k < f.W; k += 256u) f.ids[j * f.W + k] = CERES_MISS;
        }
    }
}
