// lbvh_build.cuh -- BVH construction on the GPU, included by ceres_b200.cu after scene_prep.cuh.
//
// The reference builds its trees on the host (SweepSahBuilder, sweep_sah_builder.hpp:43-252); for the 10 M-triangle
// configuration that build is the wall-clock bottleneck of scene preparation.  This is the GPU-friendly alternative the
// reference itself sketches in lib/bvh/linear_bvh_builder.hpp:118-185: Morton codes of the primitive centres, a radix
// sort, and a radix-tree hierarchy -- here built in one pass with Karras' construction ("Maximizing Parallelism in the
// Construction of BVHs, Octrees, and k-d Trees", 2012) instead of the reference's level-by-level merging.
//
// Output = the REFERENCE's own data structure (bvh.hpp:25-30,93-96): node 0 is the root, the two children of an inner
// node are adjacent, a leaf holds `primitive_count` slots of primitive_indices starting at `first_child_or_primitive`.
// So the result can be handed to ceres_scene_create, or to the reference's traverser alike.  It is NOT the
// tree SweepSah would build: which triangle wins an exact-t tie depends on the tree, so parity for
// frames rendered from this tree is checked against the reference traversal walking the SAME tree (tests/test_gpu_lbvh.py).
//
//   lbvh_boxes_kernel    per triangle: bounding_box() (triangle.hpp:39-44) and its centre; scene box by atomics
//   lbvh_codes_kernel    30-bit Morton code of the centre in the scene box (10 bits per axis)
//   cub::DeviceRadixSort sorts (code, triangle) pairs
//   lbvh_tree_kernel     inner node i of the radix tree over the keys (code << 32 | position): range, split, children
//   lbvh_emit_kernel     writes the reference nodes of both children of inner node i at 1 + 2 i, 2 + 2 i
//   lbvh_fit_kernel      bottom-up boxes: one thread per leaf climbs while it is the second child to arrive
// One triangle per leaf; boxes are exact unions (std::min/std::max semantics), so HierarchyRefitter::refit reproduces
// them bit for bit -- that is the structural test.
#include <cub/device/device_radix_sort.cuh>

#define LBVH_LEAF 0x80000000u

__device__ __forceinline__ unsigned int lbvh_ord(float f) {            // order-preserving float -> uint
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float lbvh_unord(unsigned int u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

// box[6*k..] = minx,maxx,miny,maxy,minz,maxz of triangle k; scene[0..5] = ordered-uint min/max of the CENTRES
__global__ void __launch_bounds__(256) lbvh_boxes_kernel(const float4* __restrict__ tris, uint32_t n, float* __restrict__ box,
                                                         unsigned int* scene) {
    const uint32_t k = blockIdx.x * 256u + threadIdx.x;
    if (k >= n) return;
    float b[6];
    leaf_box(tris, k, 1u, b);
    for (int d = 0; d < 6; ++d) box[6ull * k + d] = b[d];
    for (int d = 0; d < 3; ++d) {
        const float c = mul(add(b[2 * d], b[2 * d + 1]), 0.5f);
        atomicMin(&scene[2 * d], lbvh_ord(c));
        atomicMax(&scene[2 * d + 1], lbvh_ord(c));
    }
}

__device__ __forceinline__ uint32_t lbvh_spread(uint32_t v) {          // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void __launch_bounds__(256) lbvh_codes_kernel(const float* __restrict__ box, uint32_t n, const unsigned int* __restrict__ scene,
                                                         uint32_t* __restrict__ code, uint32_t* __restrict__ idx) {
    const uint32_t k = blockIdx.x * 256u + threadIdx.x;
    if (k >= n) return;
    uint32_t q[3];
    for (int d = 0; d < 3; ++d) {
        const float lo = lbvh_unord(scene[2 * d]), hi = lbvh_unord(scene[2 * d + 1]);
        const float c = 0.5f * (box[6ull * k + 2 * d] + box[6ull * k + 2 * d + 1]);
        const float ext = hi - lo;
        float t = ext > 0.0f ? (c - lo) / ext : 0.0f;
        t = fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f);
        q[d] = (uint32_t)t;
    }
    code[k] = (lbvh_spread(q[0]) << 2) | (lbvh_spread(q[1]) << 1) | lbvh_spread(q[2]);
    idx[k] = k;
}

// common-prefix length of the keys (code << 32 | position) at sorted positions i and j; -1 outside [0, n)
__device__ __forceinline__ int lbvh_delta(const uint32_t* __restrict__ code, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = ((unsigned long long)code[i] << 32) | (unsigned)i, b = ((unsigned long long)code[j] << 32) | (unsigned)j;
    return __clzll((long long)(a ^ b));
}

// Karras 2012, section 4: inner node i covers the keys [min(i,j), max(i,j)] and splits them after position gamma.
// child[2i], child[2i+1]: inner node index, or leaf position | LBVH_LEAF.  par_*: parent inner node << 1 | side.
__global__ void __launch_bounds__(256) lbvh_tree_kernel(const uint32_t* __restrict__ code, int n, uint32_t* __restrict__ child,
                                                        uint32_t* __restrict__ par_inner, uint32_t* __restrict__ par_leaf) {
    const int i = (int)(blockIdx.x * 256u + threadIdx.x);
    if (i >= n - 1) return;
    const int d = lbvh_delta(code, n, i, i + 1) - lbvh_delta(code, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(code, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(code, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(code, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(code, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
        if (lbvh_delta(code, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + (d < 0 ? d : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    const uint32_t L = lo == gamma ? ((uint32_t)gamma | LBVH_LEAF) : (uint32_t)gamma;
    const uint32_t R = hi == gamma + 1 ? ((uint32_t)(gamma + 1) | LBVH_LEAF) : (uint32_t)(gamma + 1);
    child[2 * i] = L; child[2 * i + 1] = R;
    if (L & LBVH_LEAF) par_leaf[L & ~LBVH_LEAF] = 2u * i; else par_inner[L] = 2u * i;
    if (R & LBVH_LEAF) par_leaf[R & ~LBVH_LEAF] = 2u * i + 1u; else par_inner[R] = 2u * i + 1u;
}

struct LbvhNode { float bounds[6]; uint32_t primitive_count; uint32_t first_child_or_primitive; };   // bvh.hpp:25-30

// children of inner node i live at nodes[1 + 2 i] and nodes[2 + 2 i]; the root (inner node 0) at nodes[0]
__global__ void __launch_bounds__(256) lbvh_emit_kernel(const uint32_t* __restrict__ child, const uint32_t* __restrict__ idx,
                                                        const float* __restrict__ box, int n, LbvhNode* __restrict__ nodes) {
    const int i = (int)(blockIdx.x * 256u + threadIdx.x);
    if (i >= n - 1) return;
    if (i == 0) { nodes[0].primitive_count = 0u; nodes[0].first_child_or_primitive = 1u; }
    for (int s = 0; s < 2; ++s) {
        const uint32_t c = child[2 * i + s];
        LbvhNode& nd = nodes[1 + 2 * i + s];
        if (c & LBVH_LEAF) {
            const uint32_t k = c & ~LBVH_LEAF;
            const float* b = box + 6ull * idx[k];
            for (int q = 0; q < 6; ++q) nd.bounds[q] = b[q];
            nd.primitive_count = 1u; nd.first_child_or_primitive = k;
        } else {
            nd.primitive_count = 0u; nd.first_child_or_primitive = 1u + 2u * c;
        }
    }
}

__global__ void __launch_bounds__(256) lbvh_fit_kernel(const uint32_t* __restrict__ par_inner, const uint32_t* __restrict__ par_leaf,
                                                       uint32_t* arrived, int n, LbvhNode* nodes) {
    const int k = (int)(blockIdx.x * 256u + threadIdx.x);
    if (k >= n) return;
    uint32_t p = par_leaf[k] >> 1;
    for (;;) {
        if (atomicAdd(&arrived[p], 1u) == 0u) return;           // the sibling subtree is not finished yet
        __threadfence();
        const float* l = nodes[1 + 2 * p].bounds; const float* r = nodes[2 + 2 * p].bounds;
        float u[6];
        for (int d = 0; d < 3; ++d) {                            // left.extend(right), bounding_box.hpp:23-27
            u[2 * d] = min_std(__ldcg(l + 2 * d), __ldcg(r + 2 * d));
            u[2 * d + 1] = max_std(__ldcg(l + 2 * d + 1), __ldcg(r + 2 * d + 1));
        }
        const uint32_t up = p == 0u ? 0u : par_inner[p];
        float* dst = p == 0u ? nodes[0].bounds : nodes[1 + 2 * (up >> 1) + (up & 1u)].bounds;
        for (int q = 0; q < 6; ++q) __stcg(dst + q, u[q]);
        if (p == 0u) return;
        __threadfence();
        p = up >> 1;
    }
}
