// ray_step.cuh -- per-lane ray state and ONE node-pair step of the reference's traversal (included by ceres_b200.cu
// after the device math helpers; used by the frame kernel and by the experiment kernels).
//
// The traversal stack lives in shared memory, one column per thread (word index = depth * blockDim + tid: bank =
// lane, conflict-free), addressed through a running word offset so that a push/pop is one STS/LDS + one add.

struct RaySlot {
    float ox, oy, oz, dx, dy, dz;      // origin, direction
    float ix, iy, iz, sx, sy, sz;      // safe inverse direction, scaled origin (node_intersectors.hpp:89-94)
    float tmax;                         // ray.tmax, shrinks with every accepted hit (single_ray_traverser.hpp:58)
    float u, v;                         // best hit so far (its t is tmax)
    uint32_t slot;                      // leaf slot of the best hit
    uint32_t cur;                       // node pair to test next
    uint32_t sp;                        // word offset of the next free stack entry in this thread's column
    uint32_t sb;                        // word offset of the column's oldest live entry (= tid unless entries were given away)
    uint32_t item;                      // work item being traced
    bool found;
};

__device__ __forceinline__ void ray_begin(RaySlot& r, V3 o, V3 d, uint32_t tid) {
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z;
    const float eps = FLT_EPSILON;                                   // vector.hpp:69-74
    r.ix = rcp(fabsf(d.x) < eps ? copysignf(eps, d.x) : d.x);
    r.iy = rcp(fabsf(d.y) < eps ? copysignf(eps, d.y) : d.y);
    r.iz = rcp(fabsf(d.z) < eps ? copysignf(eps, d.z) : d.z);
    r.sx = mul(-o.x, r.ix); r.sy = mul(-o.y, r.iy); r.sz = mul(-o.z, r.iz);
    r.tmax = FLT_MAX; r.u = 0.0f; r.v = 0.0f; r.slot = 0; r.cur = 0; r.sp = tid; r.sb = tid; r.found = false;
}

// Child order of a step whose two children are both inner nodes and both hit.
//   ORDER_NEAR   the reference's rule (single_ray_traverser.hpp:109-115): nearer entry first, ties keep the left child.
//                Mandatory for closest-hit rays: it decides which triangle wins an exact-t tie.
//   ORDER_LEFT   always the left child first.  Only for any-hit rays, whose boolean does not depend on the order (a
//                leaf is reached iff its own box and all boxes above it pass the slab test, whatever the order; tmax
//                never shrinks before the first accepted triangle, and that one ends the ray).
#define ORDER_NEAR 0
#define ORDER_LEFT 1

// One node-pair step of single_ray_traverser.hpp:82-123 for this lane's ray, including the eager leaf tests.
// Returns false when the traversal is over (stack empty, or first hit in ANY mode).
template <bool ANY, bool STATS, int STRIDE, int ORDER = ORDER_NEAR>
__device__ __forceinline__ bool ray_step(const float4* __restrict__ pairs, const float4* __restrict__ tris, RaySlot& r,
                                         uint32_t* __restrict__ stack, uint32_t tid, uint32_t& n_steps, uint32_t& n_tests,
                                         const DevScene& sc) {
    if (STATS) n_steps++;
    CERES_CLAMP(r.cur, sc.n_pairs, CHECK_PAIR_INDEX, sc);
    const float4* pp = pairs + 4ull * r.cur;
    const float4 q0 = __ldg(pp), q1 = __ldg(pp + 1), q2 = __ldg(pp + 2), q3 = __ldg(pp + 3);
    // octant = signbit(direction) = signbit(inverse) (safe_inverse keeps the sign, -0 included)
    const bool nx = __float_as_int(r.ix) < 0, ny = __float_as_int(r.iy) < 0, nz = __float_as_int(r.iz) < 0;
    const float tmax = r.tmax;
    float le, lx, re, rx;
    {
        const float e0 = __fmaf_rn(nx ? q0.y : q0.x, r.ix, r.sx), x0 = __fmaf_rn(nx ? q0.x : q0.y, r.ix, r.sx);
        const float e1 = __fmaf_rn(ny ? q0.w : q0.z, r.iy, r.sy), x1 = __fmaf_rn(ny ? q0.z : q0.w, r.iy, r.sy);
        const float e2 = __fmaf_rn(nz ? q1.y : q1.x, r.iz, r.sz), x2 = __fmaf_rn(nz ? q1.x : q1.y, r.iz, r.sz);
        le = fmaxf(e0, fmaxf(e1, fmaxf(e2, 0.0f)));          // tmin = 0 for every ray of render() (ray.hpp:19)
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

    // leaves: the left range, then the right range, slots ascending (single_ray_traverser.hpp:89-107,51-62)
    const uint32_t nL = hitL ? cL : 0u, nR = hitR ? cR : 0u;
    const uint32_t nT = nL + nR;
    if (nT) {
        const V3 o = mk(r.ox, r.oy, r.oz), d = mk(r.dx, r.dy, r.dz);
        for (uint32_t k = 0; k < nT; ++k) {
            uint32_t s = k < nL ? fL + k : fR + (k - nL);
            CERES_CLAMP(s, sc.n_slots, CHECK_SLOT_INDEX, sc);
            if (STATS) n_tests++;
            const Tri tr = load_tri(tris, s);
            float t, u, v;
            if (tri_test(tr, o, d, 0.0f, r.tmax, t, u, v)) {
                r.slot = s; r.u = u; r.v = v; r.tmax = t; r.found = true;
                if (ANY) return false;
            }
        }
    }

    const bool goL = hitL && cL == 0u, goR = hitR && cR == 0u;
    if (goL || goR) {
        // ties keep the left child first (:111-112)
        const bool right_first = ORDER == ORDER_NEAR ? goR && (!goL || le > re) : !goL;
        if (goL && goR) {
            const uint32_t far_pair = right_first ? fL : fR;
            CERES_CK(r.sp < tid + sc.stack_depth * STRIDE, CHECK_STACK_DEPTH, sc);
            stack[r.sp] = far_pair;
            r.sp += STRIDE;
        }
        r.cur = right_first ? fR : fL;
        return true;
    }
    if (r.sp == r.sb) return false;
    CERES_CK(r.sp >= r.sb + STRIDE && r.sb >= tid, CHECK_STACK_UNDER, sc);
    r.sp -= STRIDE;
    r.cur = stack[r.sp];
    return true;
}

// work item -> pixel of the culled tile rectangle.  Returns false for the padding pixels of edge tiles.
__device__ __forceinline__ bool item_pixel(const Frame& f, uint32_t item, uint32_t& i, uint32_t& lrow) {
    const uint32_t tile = item >> 5, p = item & 31u;
    const uint32_t ty = tile / f.tiles_x, tx = tile - ty * f.tiles_x;
    i = f.rect_x0 + tx * 8u + (p & 7u);
    lrow = f.rect_row0 + ty * 4u + (p >> 3);
    return i < f.rect_x1 && lrow < f.rect_row1;
}
