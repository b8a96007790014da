// trace_shadow_spec.cuh -- shadow wave with DEFERRED triangle tests (tuning variant 5; included after trace_batched.cuh,
// whose batch_node_step() it reuses).
//
// ncu, source page of the default shadow wave (profiles/r1/sweeps.md): 46 % of its issue slots go to the triangle block,
// which runs for ~6 of 32 lanes -- in almost every round a couple of lanes have reached a leaf.  Parking those lanes
// until enough of them wait (variant 3) does not help: a parked lane does not step, so the node rounds lose the lanes the
// triangle rounds win.  For SHADOW rays the two can be decoupled: only bool(hit) is consumed (render.hpp:137-139), tmax
// stays FLT_MAX until the first accepted triangle, and that triangle ends the ray -- so no node step ever depends on the
// outcome of an earlier triangle test.  A lane therefore queues the leaf ranges it meets (a small per-thread list in shared
// memory) and KEEPS STEPPING; when three quarters of the live lanes have triangles queued (or nobody can step) the warp
// runs a triangle round, one triangle per lane.  The only cost is speculative node steps of rays whose queued triangles
// will turn out to be hits.  The set of triangles a ray can be tested against is the same as in the ordered traversal
// minus nothing (no pruning happens before the first hit), so the verdict is identical; S'/T' statistics of the shadow
// wave differ from variant 0 (a little more node work, the same or less triangle work).

#define SPEC_PMAX 6u        // queued leaf ranges per thread (two words each)

template <bool STATS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) trace_shadow_spec_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_dyn[];
    uint32_t* stack = smem_dyn;                                     // stack_depth x THREADS
    uint32_t* pend = smem_dyn + sc.stack_depth * THREADS;           // 2*SPEC_PMAX x THREADS: (slot, count) entries
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t total = f.ctr->qcount;
    uint32_t* cursor = &f.ctr->shadow_next;
    const float4* __restrict__ pairs = sc.pairs;
    const float4* __restrict__ tris = sc.tris;
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);

    RaySlot r;
    r.found = false; r.item = 0; r.sp = tid; r.cur = BATCH_ENDED; r.tmax = FLT_MAX; r.u = r.v = 0.0f; r.slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.ix = r.iy = r.iz = r.sx = r.sy = r.sz = 0.0f;
    bool live = false;
    uint32_t pn = 0;                    // entries in my pending list
    uint32_t c_slot = 0, c_n = 0;       // the range being tested
    uint32_t n_steps = 0, n_tests = 0, occl = 0;
    unsigned long long st0 = 0, st1 = 0;

    for (;;) {
        // ---- whole-warp refill (all rays of the warp are finished here) ------------------------------------------------
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= total) break;
        const uint32_t item = base + lane;
        if (item < total) {
            r.item = item;
            const uint4 rec = f.queue[item];
            const Tri tr = load_tri(tris, rec.y);
            V3 P, sun_line;
            shadow_setup(tr, __uint_as_float(rec.z), __uint_as_float(rec.w), sun, P, sun_line);
            ray_begin(r, P, sun_line, tid);
            r.found = false;
            live = true; pn = 0; c_n = 0;
            if (sc.root_leaf_count) { c_slot = 0u; c_n = sc.root_leaf_count; r.cur = BATCH_ENDED; }
        }
        // ---- rounds ---------------------------------------------------------------------------------------------------
        for (;;) {
            const bool can_step = live && r.cur != BATCH_ENDED && pn + 2u <= SPEC_PMAX;
            const bool has_tri = live && (c_n | pn) != 0u;
            const unsigned ms = __ballot_sync(0xFFFFFFFFu, can_step), mp = __ballot_sync(0xFFFFFFFFu, has_tri);
            if (!(ms | mp)) break;
            const uint32_t n_live = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, live));
            if (ms != 0u && (uint32_t)__popc(mp) * 4u < n_live * 3u) {
                if (can_step) {
                    uint32_t a_slot, a_n, b_slot, b_n;
                    batch_node_step<STATS, THREADS>(pairs, r, stack, tid, a_slot, a_n, b_slot, b_n, n_steps);
                    if (a_n) { pend[(2u * pn) * THREADS + tid] = a_slot; pend[(2u * pn + 1u) * THREADS + tid] = a_n; ++pn; }
                    if (b_n) { pend[(2u * pn) * THREADS + tid] = b_slot; pend[(2u * pn + 1u) * THREADS + tid] = b_n; ++pn; }
                }
            } else if (has_tri) {
                if (c_n == 0u) { --pn; c_slot = pend[(2u * pn) * THREADS + tid]; c_n = pend[(2u * pn + 1u) * THREADS + tid]; }
                if (STATS) n_tests++;
                const Tri tr = load_tri(tris, c_slot);
                float t, u, v;
                if (tri_test(tr, mk(r.ox, r.oy, r.oz), mk(r.dx, r.dy, r.dz), 0.0f, FLT_MAX, t, u, v)) {
                    r.found = true; live = false;                                   // occluded: the ray is over
                    atomicOr(&f.queue[r.item].y, OCCLUDED_BIT);
                    occl++;
                }
                ++c_slot; --c_n;
            }
            if (live && r.cur == BATCH_ENDED && (c_n | pn) == 0u) live = false;      // walked everything: lit
        }
        if (STATS) { st0 += n_steps; st1 += n_tests; n_steps = n_tests = 0; }
    }

    for (int o = 16; o; o >>= 1) occl += __shfl_xor_sync(0xFFFFFFFFu, occl, o);
    if (lane == 0 && occl) atomicAdd(&f.ctr->occluded, occl);
    if (STATS) {
        for (int o = 16; o; o >>= 1) { st0 += __shfl_xor_sync(0xFFFFFFFFu, st0, o); st1 += __shfl_xor_sync(0xFFFFFFFFu, st1, o); }
        if (lane == 0) { atomicAdd(&f.ctr->stats[4], st0); atomicAdd(&f.ctr->stats[5], st1); }
    }
}
