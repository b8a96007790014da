// scene_prep.cuh -- scene preparation on the device, included by ceres_b200.cu after the device
// math helpers.  The step BEFORE the render path: moving the object of an animation without a host rebuild.
//
//   prep_triangles_kernel   rotate_triangles<Axis> (render.hpp:24-44) and transform_triangles (transform.hpp:106-129):
//                           every Triangle is rebuilt from its transformed p0, p1(), p2() (triangle.hpp:30-37).
//   refit_links_kernel      parent link and number of inner children of every node pair (bottom_up_algorithm.hpp:32-43).
//   refit_kernel            HierarchyRefitter::refit (hierarchy_refitter.hpp:17-32) over the pair array: bottom-up, one
//                           thread starts at every pair whose children are both leaves and climbs while it is the last
//                           child to arrive (bottom_up_algorithm.hpp:62-84: Karras' atomic-flag walk).
// Arithmetic: the same binary32 operations in the same order as the reference (non-contracting *_rn intrinsics);
// min/max are std::min/std::max (b < a ? b : a), so the boxes are bit-identical to the reference's, signed zeros
// included.  HBM-bound passes: 48 B read + 48 B written per triangle; 48 B per triangle + 64 B r/w per pair.

#define PREP_ROTATE_X 0
#define PREP_ROTATE_Y 1
#define PREP_ROTATE_Z 2
#define PREP_AFFINE 3

struct PrepParams {
    int mode;
    float c, s;          // rotate: cos, sin (computed on the host like the reference, render.hpp:27-28)
    float a[9], v[3];    // affine: Transform::a (row-major), Transform::v
};

__device__ __forceinline__ V3 prep_point(const PrepParams& q, V3 p) {
    switch (q.mode) {
    case PREP_ROTATE_X: return mk(p.x, sub(mul(p.y, q.c), mul(p.z, q.s)), add(mul(p.y, q.s), mul(p.z, q.c)));
    case PREP_ROTATE_Y: return mk(add(mul(p.x, q.c), mul(p.z, q.s)), p.y, add(mul(-p.x, q.s), mul(p.z, q.c)));
    case PREP_ROTATE_Z: return mk(sub(mul(p.x, q.c), mul(p.y, q.s)), add(mul(p.x, q.s), mul(p.y, q.c)), p.z);
    default:            // transform.hpp:106-112: a[r][0]*p[0] + a[r][1]*p[1] + a[r][2]*p[2] + v[r], left to right
        return mk(add(add(add(mul(q.a[0], p.x), mul(q.a[1], p.y)), mul(q.a[2], p.z)), q.v[0]),
                  add(add(add(mul(q.a[3], p.x), mul(q.a[4], p.y)), mul(q.a[5], p.z)), q.v[1]),
                  add(add(add(mul(q.a[6], p.x), mul(q.a[7], p.y)), mul(q.a[8], p.z)), q.v[2]));
    }
}

__global__ void __launch_bounds__(256) prep_triangles_kernel(float4* __restrict__ tris, uint32_t n, const PrepParams q) {
    for (uint32_t s = blockIdx.x * 256u + threadIdx.x; s < n; s += gridDim.x * 256u) {
        float4* p = tris + 3ull * s;
        const float4 a = p[0], b = p[1], c = p[2];
        const V3 p0 = mk(a.x, a.y, a.z), e1 = mk(a.w, b.x, b.y), e2 = mk(b.z, b.w, c.x);
        const V3 q0 = prep_point(q, p0), q1 = prep_point(q, vsub(p0, e1)), q2 = prep_point(q, vadd(p0, e2));
        const V3 f1 = vsub(q0, q1), f2 = vsub(q2, q0), nn = vcross(f1, f2);       // Triangle(p0, p1, p2), left-handed normal
        p[0] = make_float4(q0.x, q0.y, q0.z, f1.x);
        p[1] = make_float4(f1.y, f1.z, f2.x, f2.y);
        p[2] = make_float4(f2.z, nn.x, nn.y, nn.z);
    }
}

#define REFIT_ROOT 0xFFFFFFFFu

// parent[p] = parent pair * 2 + side of the inner child that owns pair p; need[p] = inner children of pair p.
__global__ void __launch_bounds__(256) refit_links_kernel(const float4* __restrict__ pairs, uint32_t n_pairs,
                                                          uint32_t* __restrict__ parent, uint32_t* __restrict__ need) {
    const uint32_t p = blockIdx.x * 256u + threadIdx.x;
    if (p >= n_pairs) return;
    if (p == 0) parent[0] = REFIT_ROOT;
    const float4 l = pairs[4ull * p + 1], r = pairs[4ull * p + 3];
    uint32_t k = 0;
    if (__float_as_uint(l.z) == 0u) { parent[__float_as_uint(l.w)] = 2u * p; ++k; }
    if (__float_as_uint(r.z) == 0u) { parent[__float_as_uint(r.w)] = 2u * p + 1u; ++k; }
    need[p] = k;
}

__device__ __forceinline__ float min_std(float a, float b) { return b < a ? b : a; }     // std::min(a, b)
__device__ __forceinline__ float max_std(float a, float b) { return a < b ? b : a; }     // std::max(a, b)

// Box of a leaf: BoundingBox::empty() extended by every primitive's bounding_box(), slots ascending.
__device__ __forceinline__ void leaf_box(const float4* __restrict__ tris, uint32_t first, uint32_t count, float* bx) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t k = 0; k < count; ++k) {
        const float4* p = tris + 3ull * (first + k);
        const float4 a = __ldcg(p), b = __ldcg(p + 1), c = __ldcg(p + 2);
        const V3 p0 = mk(a.x, a.y, a.z), p1 = vsub(p0, mk(a.w, b.x, b.y)), p2 = vadd(p0, mk(b.z, b.w, c.x));
        const float v0[3] = {p0.x, p0.y, p0.z}, v1[3] = {p1.x, p1.y, p1.z}, v2[3] = {p2.x, p2.y, p2.z};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const float mn = min_std(min_std(v0[d], v1[d]), v2[d]), mx = max_std(max_std(v0[d], v1[d]), v2[d]);
            lo[d] = min_std(lo[d], mn); hi[d] = max_std(hi[d], mx);
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) { bx[2 * d] = lo[d]; bx[2 * d + 1] = hi[d]; }
}

__global__ void __launch_bounds__(256) refit_kernel(float* pairs, const float4* __restrict__ tris,
                                                    const uint32_t* __restrict__ parent, const uint32_t* __restrict__ need,
                                                    uint32_t* arrived, uint32_t n_pairs, float* root_bounds) {
    uint32_t p = blockIdx.x * 256u + threadIdx.x;
    if (p >= n_pairs || need[p] != 0u) return;
    for (;;) {
        float* rec = pairs + 16ull * p;
        float box[2][6];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const uint32_t cnt = __float_as_uint(__ldcg(rec + 8 * c + 6)), first = __float_as_uint(__ldcg(rec + 8 * c + 7));
            if (cnt) {
                leaf_box(tris, first, cnt, box[c]);
#pragma unroll
                for (int k = 0; k < 6; ++k) rec[8 * c + k] = box[c][k];
            } else {
                // written by the thread that finished that child, before its fence + atomic
#pragma unroll
                for (int k = 0; k < 6; ++k) box[c][k] = __ldcg(rec + 8 * c + k);
            }
        }
        float u[6];       // left.extend(right): min = std::min(L.min, R.min), max = std::max(L.max, R.max)
#pragma unroll
        for (int d = 0; d < 3; ++d) { u[2 * d] = min_std(box[0][2 * d], box[1][2 * d]); u[2 * d + 1] = max_std(box[0][2 * d + 1], box[1][2 * d + 1]); }
        const uint32_t par = parent[p];
        if (par == REFIT_ROOT) {
#pragma unroll
            for (int k = 0; k < 6; ++k) root_bounds[k] = u[k];
            return;
        }
        const uint32_t pp = par >> 1, side = par & 1u;
        float* dst = pairs + 16ull * pp + 8u * side;
#pragma unroll
        for (int k = 0; k < 6; ++k) __stcg(dst + k, u[k]);
        __threadfence();
        const uint32_t old = atomicAdd(&arrived[pp], 1u);
        if (old + 1u != need[pp]) return;
        arrived[pp] = 0u;           // re-armed for the next refit
        __threadfence();
        p = pp;
    }
}
