// trace_repack.cuh -- persistent CTAs that REPACK their surviving rays every K node-pair steps (tuning variant 4;
// included after trace_persistent.cuh, whose RaySlot / ray_begin / ray_step / item_pixel it reuses).
//
// Why (profiles/r1/sweeps.md, "Steps per ray"): with whole-warp refill a warp costs the maximum of its 32 rays; the
// dragon's shadow wave runs with 14 of 32 lanes because its rays need 8 to 124 steps (mean 24).  Here the traversal is
// cut into slices of K steps.  At the end of a slice the CTA's warps meet at a barrier, the rays that are still alive are
// packed densely into the first lanes of the CTA (their state travels through shared memory; the traversal stack stays
// where it is, every ray carries the index of its stack column), and the freed lanes take fresh work items.  Rays that
// share a warp after a repack have all survived the same number of slices, so their remaining lengths are closer, and
// idle lanes never wait longer than K steps.  A warp whose rays have ALL finished inside a slice refills itself at
// once, as the default kernel does.
// Per-ray semantics are untouched: a ray executes exactly the same sequence of ray_step() calls, only on varying lanes.

#define REPACK_K 16u

// words of ray state that travel at a repack
template <int WAVE, bool STATS> struct RepackWords { static constexpr int N = (WAVE == WAVE_PRIMARY ? 21 : 16) + (STATS ? 2 : 0); };

template <int WAVE, bool STATS, int THREADS>
__global__ void __launch_bounds__(THREADS, TRACE_MIN_CTAS(THREADS)) trace_repack_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_dyn[];
    constexpr int NW = THREADS / 32, XW = RepackWords<WAVE, STATS>::N;
    uint32_t* stack = smem_dyn;                                   // stack_depth x THREADS, column = ray's `col`
    uint32_t* xchg = smem_dyn + sc.stack_depth * THREADS;         // XW x THREADS: state exchange, word k of slot j at k*THREADS+j
    __shared__ uint32_t s_cnt[NW], s_free[NW], s_base, s_busy[THREADS];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
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
    uint32_t col = tid;             // stack column owned by this lane's ray (or free for the next ray of this lane)
    bool live = false, retire = false;
    uint32_t n_steps = 0, n_tests = 0, occl = 0;
    unsigned long long st[4] = {0, 0, 0, 0};

    // publishes the finished rays of this warp (queue append / statistics); called with the whole warp converged
    auto flush = [&]() {
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
    };
    // starts work item `item` in this lane (stack column `col`); leaves live = false for padding pixels / root leaves
    auto begin = [&](uint32_t item) {
        r.item = item;
        bool have = true;
        if (WAVE == WAVE_PRIMARY) {
            uint32_t i, lrow;
            have = item_pixel(f, item, i, lrow);
            if (have) ray_begin(r, eye, primary_dir(f, i, frame_row(f, lrow)), col);
        } else {
            const uint4 rec = f.queue[item];
            const Tri tr = load_tri(tris, rec.y);
            V3 P, sun_line;
            shadow_setup(tr, __uint_as_float(rec.z), __uint_as_float(rec.w), sun, P, sun_line);
            ray_begin(r, P, sun_line, col);
        }
        if (!have) return;
        if (sc.root_leaf_count) {
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
            live = true;
        }
    };

    for (;;) {
        // ---- one slice: K step iterations per warp; a warp whose rays have all finished refills itself at once ---------
        for (uint32_t it = 0; it < REPACK_K; ++it) {
            if (live) {
                live = ray_step<WAVE == WAVE_SHADOW, STATS, THREADS>(pairs, tris, r, stack, col, n_steps, n_tests, sc);
                if (!live) {
                    retire = true;
                    if (WAVE == WAVE_SHADOW && r.found) { atomicOr(&f.queue[r.item].y, OCCLUDED_BIT); occl++; }
                }
            }
            if (!__any_sync(0xFFFFFFFFu, live)) {
                flush();
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(cursor, 32u);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (base >= total) break;                   // no work left: wait for the others at the barrier
                if (base + lane < total) begin(base + lane);
            }
        }
        flush();

        // ---- repack: survivors to the front of the CTA, fresh work items into the freed lanes ---------------------------
        const unsigned m_live = __ballot_sync(0xFFFFFFFFu, live);
        if (lane == 0) s_cnt[warp] = (uint32_t)__popc(m_live);
        s_busy[tid] = 0u;
        __syncthreads();
        uint32_t before = 0, S = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) { const uint32_t c = s_cnt[w]; if ((uint32_t)w < warp) before += c; S += c; }
        if (live) {
            uint32_t* x = xchg + before + (uint32_t)__popc(m_live & lt_mask);
            int k = 0;
            x[THREADS * k++] = r.item; x[THREADS * k++] = r.cur; x[THREADS * k++] = r.sp; x[THREADS * k++] = col;
            x[THREADS * k++] = __float_as_uint(r.ox); x[THREADS * k++] = __float_as_uint(r.oy); x[THREADS * k++] = __float_as_uint(r.oz);
            x[THREADS * k++] = __float_as_uint(r.dx); x[THREADS * k++] = __float_as_uint(r.dy); x[THREADS * k++] = __float_as_uint(r.dz);
            x[THREADS * k++] = __float_as_uint(r.ix); x[THREADS * k++] = __float_as_uint(r.iy); x[THREADS * k++] = __float_as_uint(r.iz);
            x[THREADS * k++] = __float_as_uint(r.sx); x[THREADS * k++] = __float_as_uint(r.sy); x[THREADS * k++] = __float_as_uint(r.sz);
            if (WAVE == WAVE_PRIMARY) {
                x[THREADS * k++] = __float_as_uint(r.tmax); x[THREADS * k++] = __float_as_uint(r.u); x[THREADS * k++] = __float_as_uint(r.v);
                x[THREADS * k++] = r.slot; x[THREADS * k++] = r.found ? 1u : 0u;
            }
            if (STATS) { x[THREADS * k++] = n_steps; x[THREADS * k++] = n_tests; }
            s_busy[col] = 1u;
        }
        __syncthreads();
        const unsigned m_free = __ballot_sync(0xFFFFFFFFu, s_busy[tid] == 0u);      // free stack columns tid
        if (lane == 0) s_free[warp] = m_free;
        if (tid == 0) s_base = S < (uint32_t)THREADS ? atomicAdd(cursor, (uint32_t)THREADS - S) : total;
        live = tid < S;
        if (live) {
            const uint32_t* x = xchg + tid;
            int k = 0;
            r.item = x[THREADS * k++]; r.cur = x[THREADS * k++]; r.sp = x[THREADS * k++]; col = x[THREADS * k++]; r.sb = col;
            r.ox = __uint_as_float(x[THREADS * k++]); r.oy = __uint_as_float(x[THREADS * k++]); r.oz = __uint_as_float(x[THREADS * k++]);
            r.dx = __uint_as_float(x[THREADS * k++]); r.dy = __uint_as_float(x[THREADS * k++]); r.dz = __uint_as_float(x[THREADS * k++]);
            r.ix = __uint_as_float(x[THREADS * k++]); r.iy = __uint_as_float(x[THREADS * k++]); r.iz = __uint_as_float(x[THREADS * k++]);
            r.sx = __uint_as_float(x[THREADS * k++]); r.sy = __uint_as_float(x[THREADS * k++]); r.sz = __uint_as_float(x[THREADS * k++]);
            if (WAVE == WAVE_PRIMARY) {
                r.tmax = __uint_as_float(x[THREADS * k++]); r.u = __uint_as_float(x[THREADS * k++]); r.v = __uint_as_float(x[THREADS * k++]);
                r.slot = x[THREADS * k++]; r.found = x[THREADS * k++] != 0u;
            } else {
                r.tmax = FLT_MAX; r.found = false;
            }
            if (STATS) { n_steps = x[THREADS * k++]; n_tests = x[THREADS * k++]; }
        }
        __syncthreads();
        const uint32_t base = s_base;
        if (S == 0u && base >= total) break;                     // CTA-uniform: nothing alive, nothing left to fetch
        if (!live) {
            // my stack column: the (tid - S)-th free one
            uint32_t want = tid - S;
            uint32_t c = 0;
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                const uint32_t m = s_free[w], n = (uint32_t)__popc(m);
                if (want < n) { c = 32u * w + __fns(m, 0, want + 1); want = 0xFFFFFFFFu; }
                else if (want != 0xFFFFFFFFu) want -= n;
            }
            col = c;
            if (STATS) n_steps = n_tests = 0;
            const uint32_t item = base + (tid - S);
            if (base < total && item < total) begin(item);
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
