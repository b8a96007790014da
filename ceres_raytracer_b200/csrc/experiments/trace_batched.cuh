// trace_batched.cuh -- persistent-warp traversal with PHASE-SORTED execution (tuning variant 3; included after
// trace_persistent.cuh, whose RaySlot / ray_begin / item_pixel it reuses).
//
// Why: ncu of the whole-warp-refill kernel (profiles/r1/ncu_full_*): the limiter is SIMD efficiency -- 23 of 32 lanes
// on the dragon, 14.7 (primary) and 10.5 (shadow) of 32 on the 10 M-triangle grid -- because (a) a warp waits for its
// longest ray and (b) the ~90-instruction triangle block runs for the few lanes whose step happened to reach a leaf.
// Refilling single lanes fixes (a) but makes (b) worse (profiles/r1/sweeps.md).  Here both are addressed:
//   * a lane whose node step reached leaves does NOT test them at once: it parks the (at most two) slot ranges;
//   * every iteration the warp votes: if more lanes are ready for a node step than are parked, the steppers step,
//     otherwise every parked lane tests ONE triangle.  Either block therefore runs with the majority of the live
//     lanes, whatever phase the individual rays are in;
//   * finished lanes are refilled from the warp's private chunk as soon as REFILL lanes are idle.
// Per-ray semantics are untouched (single_ray_traverser.hpp:65-126): the slab tests of a step use the tmax of the
// step's start; the step's descend/pop decision does not depend on its leaf tests (:109-123), so it is taken at once;
// the parked triangles (left range then right range, slots ascending, :89-107) are all tested before the lane's next
// node step sees tmax.  Only the interleaving ACROSS lanes changes, so hit ids stay bit-exact.

#define BATCH_ENDED 0xFFFFFFFFu      // r.cur: the stack ran empty, nothing left to step

// Node part of one step.  Parks the leaf ranges in (p_slot, p_n, q_slot, q_n); leaves r.cur = next pair or BATCH_ENDED.
template <bool STATS, int STRIDE>
__device__ __forceinline__ void batch_node_step(const float4* __restrict__ pairs, RaySlot& r, uint32_t* __restrict__ stack,
                                                uint32_t tid, uint32_t& p_slot, uint32_t& p_n, uint32_t& q_slot, uint32_t& q_n,
                                                uint32_t& n_steps) {
    if (STATS) n_steps++;
    const float4* pp = pairs + 4ull * r.cur;
    const float4 q0 = __ldg(pp), q1 = __ldg(pp + 1), q2 = __ldg(pp + 2), q3 = __ldg(pp + 3);
    const bool nx = __float_as_int(r.ix) < 0, ny = __float_as_int(r.iy) < 0, nz = __float_as_int(r.iz) < 0;
    const float tmax = r.tmax;
    float le, lx, re, rx;
    {
        const float e0 = __fmaf_rn(nx ? q0.y : q0.x, r.ix, r.sx), x0 = __fmaf_rn(nx ? q0.x : q0.y, r.ix, r.sx);
        const float e1 = __fmaf_rn(ny ? q0.w : q0.z, r.iy, r.sy), x1 = __fmaf_rn(ny ? q0.z : q0.w, r.iy, r.sy);
        const float e2 = __fmaf_rn(nz ? q1.y : q1.x, r.iz, r.sz), x2 = __fmaf_rn(nz ? q1.x : q1.y, r.iz, r.sz);
        le = fmaxf(e0, fmaxf(e1, fmaxf(e2, 0.0f)));
        lx = fminf(x0, fminf(x1, fminf(x2, tmax)));
    }
    {
        const float e0 = __fmaf_rn(nx ? q2.y : q2.x, r.ix, r.sx), x0 = __fmaf_rn(nx ? q2.x : q2.y, r.ix, r.sx);
        const float e1 = __fmaf_rn(ny ? q2.w : q2.z, r.iy, r.sy), x1 = __fmaf_rn(ny ? q2.z : q2.w, r.iy, r.sy);
        const float e2 = __fmaf_rn(nz ? q3.y : q3.x, r.iz, r.sz), x2 = __fmaf_rn(nz ? q3.x : q3.y, r.iz, r.sz);
        re = fmaxf(e0, fmaxf(e1, fmaxf(e2, 0.0f)));
        rx = fminf(x0, fminf(x1, fminf(x2, tmax)));
    }
    const uint32_t cL = __float_as_uint(q1.z), fL = __float_as_uint(q1.w);
    const uint32_t cR = __float_as_uint(q3.z), fR = __float_as_uint(q3.w);
    const bool hitL = le <= lx, hitR = re <= rx;
    // parked leaf ranges: the left one first (if any), then the right one
    const uint32_t nL = hitL ? cL : 0u, nR = hitR ? cR : 0u;
    if (nL) { p_slot = fL; p_n = nL; q_slot = fR; q_n = nR; }
    else    { p_slot = fR; p_n = nR; q_slot = 0u; q_n = 0u; }
    const bool goL = hitL && cL == 0u, goR = hitR && cR == 0u;
    if (goL || goR) {
        const bool right_first = goR && (!goL || le > re);
        if (goL && goR) { stack[r.sp] = right_first ? fL : fR; r.sp += STRIDE; }
        r.cur = right_first ? fR : fL;
    } else if (r.sp == tid) {
        r.cur = BATCH_ENDED;
    } else {
        r.sp -= STRIDE;
        r.cur = stack[r.sp];
    }
}

template <int WAVE, bool STATS, int THREADS>
__global__ void __launch_bounds__(THREADS) trace_batched_kernel(const DevScene sc, const Frame f) {
    extern __shared__ uint32_t smem_stack[];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t total = WAVE == WAVE_PRIMARY ? f.work_items : f.ctr->qcount;
    uint32_t* cursor = WAVE == WAVE_PRIMARY ? &f.ctr->primary_next : &f.ctr->shadow_next;
    const float4* __restrict__ pairs = sc.pairs;
    const float4* __restrict__ tris = sc.tris;
    const V3 sun = mk(f.sun[0], f.sun[1], f.sun[2]);
    const V3 eye = mk(f.eye[0], f.eye[1], f.eye[2]);
    const uint32_t refill_min = f.refill_min;

    RaySlot r;
    r.found = false; r.item = 0; r.sp = tid; r.cur = BATCH_ENDED; r.tmax = 0.0f; r.u = r.v = 0.0f; r.slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.ix = r.iy = r.iz = r.sx = r.sy = r.sz = 0.0f;
    bool live = false;                          // this lane holds a ray that has not been retired yet
    uint32_t p_slot = 0, p_n = 0, q_slot = 0, q_n = 0;      // parked leaf ranges (tested before the next node step)
    bool exhausted = false;                     // warp-uniform: no work items left to fetch
    uint32_t chunk_next = 0, chunk_end = 0;     // warp-uniform: this warp's private range of work items
    uint32_t n_steps = 0, n_tests = 0, occl = 0;
    unsigned long long st[4] = {0, 0, 0, 0};

    for (;;) {
        // ---- retire the finished rays ----------------------------------------------------------------------------
        const bool fin = live && p_n == 0u && r.cur == BATCH_ENDED;
        if (WAVE == WAVE_PRIMARY) {
            const bool out = fin && r.found;
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
        } else if (fin && r.found) {
            atomicOr(&f.queue[r.item].y, OCCLUDED_BIT);
            occl++;
        }
        if (STATS && fin) {
            if (WAVE == WAVE_PRIMARY) { st[r.found ? 2 : 0] += n_steps; st[r.found ? 3 : 1] += n_tests; }
            else { st[0] += n_steps; st[1] += n_tests; }
            n_steps = n_tests = 0;
        }
        if (fin) live = false;
        // ---- refill idle lanes from the warp's private chunk ------------------------------------------------------
        if (!exhausted) {
            const unsigned idle = ~__ballot_sync(0xFFFFFFFFu, live);
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
            const bool take = !live && rank < avail;
            chunk_next += min(avail, (uint32_t)__popc(idle));
            exhausted = chunk_next >= total;
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
                    live = true;
                    p_n = q_n = 0u;
                    if (sc.root_leaf_count) {       // the root itself is a leaf (single_ray_traverser.hpp:72-73)
                        p_slot = 0u; p_n = sc.root_leaf_count; r.cur = BATCH_ENDED;
                    }
                }
            }
        }
        if (!__ballot_sync(0xFFFFFFFFu, live)) {
            if (exhausted) break;
            continue;
        }
        // ---- phase-sorted execution ---------------------------------------------------------------------------------
        for (;;) {
            const bool can_step = live && p_n == 0u && r.cur != BATCH_ENDED;
            const bool parked = live && p_n != 0u;
            const unsigned ms = __ballot_sync(0xFFFFFFFFu, can_step), mp = __ballot_sync(0xFFFFFFFFu, parked);
            if (!(ms | mp)) break;
            if (!exhausted && 32u - (uint32_t)__popc(ms | mp) >= refill_min) break;     // enough idle lanes: go refill
            if (__popc(ms) >= __popc(mp)) {
                if (can_step) batch_node_step<STATS, THREADS>(pairs, r, smem_stack, tid, p_slot, p_n, q_slot, q_n, n_steps);
            } else if (parked) {
                if (STATS) n_tests++;
                const Tri tr = load_tri(tris, p_slot);
                float t, u, v;
                const bool hit = tri_test(tr, mk(r.ox, r.oy, r.oz), mk(r.dx, r.dy, r.dz), 0.0f, r.tmax, t, u, v);
                if (hit) { r.slot = p_slot; r.u = u; r.v = v; r.tmax = t; r.found = true; }
                ++p_slot; --p_n;
                if (p_n == 0u) { p_slot = q_slot; p_n = q_n; q_n = 0u; }
                if (WAVE == WAVE_SHADOW && hit) { p_n = q_n = 0u; r.cur = BATCH_ENDED; }      // any hit will do
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
