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
#include <cub/device/device_scan.cuh>

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
    const bool live = k < n;
    float b[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (live) {
        leaf_box(tris, k, 1u, b);
        for (int d = 0; d < 6; ++d) box[6ull * k + d] = b[d];
    }
    // six words for the whole scene: one atomic per warp, not per triangle
    for (int d = 0; d < 3; ++d) {
        const unsigned int c = lbvh_ord(mul(add(b[2 * d], b[2 * d + 1]), 0.5f));
        const unsigned int lo = __reduce_min_sync(0xFFFFFFFFu, live ? c : 0xFFFFFFFFu), hi = __reduce_max_sync(0xFFFFFFFFu, live ? c : 0u);
        if ((threadIdx.x & 31u) == 0u && lo <= hi) { atomicMin(&scene[2 * d], lo); atomicMax(&scene[2 * d + 1], hi); }
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

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018; the reference sketches the same idea on the host in
// lib/bvh/locally_ordered_clustering_builder.hpp:186-234) ---------------------------------------------------------------------
// The Morton-sorted leaves are the initial clusters.  Every round each cluster looks `radius` places to either side for the
// partner whose union with it has the smallest surface area; mutual choices merge into a new inner node, the array is compacted
// (order kept) and the round repeats until one cluster is left.  The tree is as good as a binned-SAH build (dragon 4K: see
// profiles/r2/sweeps.md) where the radix tree of the Morton codes alone is ~19 % slower to render.  Same intermediate form as
// the radix tree (child[], par_inner[], par_leaf[], inner node 0 = root), so emit / fit / depth kernels are shared: inner nodes
// are numbered DOWNWARDS from n - 2 as they are created, the last merge makes node 0.
struct PlocState { uint32_t m, next_id, merges, pad; };      // clusters left, next inner id (exclusive upper end), merges of this round

__device__ __forceinline__ float ploc_area(const float* a, const float* b) {
    const float dx = fmaxf(a[1], b[1]) - fminf(a[0], b[0]), dy = fmaxf(a[3], b[3]) - fminf(a[2], b[2]), dz = fmaxf(a[5], b[5]) - fminf(a[4], b[4]);
    return dx * dy + dy * dz + dz * dx;
}

// nn[i] = the cluster within `radius` positions of i whose union with i has the smallest half-area; ties go to the nearer
// position and then to i ^ 1, so that runs of identical boxes (coincident triangles) pair up (0,1)(2,3)... instead of chaining
__global__ void __launch_bounds__(256) ploc_nn_kernel(const float* __restrict__ cbox, const PlocState* st, int radius, uint32_t* __restrict__ nn) {
    const int m = (int)st->m, i = (int)(blockIdx.x * 256u + threadIdx.x);
    if (i >= m) return;
    float me[6];
    for (int q = 0; q < 6; ++q) me[q] = __ldg(cbox + 6ull * i + q);
    float best = 3.4e38f; int bj = i;
    const int lo = max(0, i - radius), hi = min(m - 1, i + radius);
    for (int j = lo; j <= hi; ++j) {
        if (j == i) continue;
        float o[6];
        for (int q = 0; q < 6; ++q) o[q] = __ldg(cbox + 6ull * j + q);
        const float a = ploc_area(me, o);
        const int dj = abs(j - i), db = abs(bj - i);
        if (a < best || (a == best && (dj < db || (dj == db && j == (i ^ 1))))) { best = a; bj = j; }
    }
    nn[i] = (uint32_t)bj;
}

// flag[i] = valid | merge << 32: a mutual pair (i, nn[i]) survives at its lower position and merges there
__global__ void __launch_bounds__(256) ploc_flag_kernel(const uint32_t* __restrict__ nn, const PlocState* st, unsigned long long* __restrict__ flag) {
    const uint32_t m = st->m, i = blockIdx.x * 256u + threadIdx.x;
    if (i >= m) return;
    const uint32_t j = nn[i];
    const bool mutual = j != i && nn[j] == i;
    flag[i] = (mutual && i > j ? 0ull : 1ull) | (mutual && i < j ? 1ull << 32 : 0ull);
}

__global__ void ploc_totals_kernel(const unsigned long long* __restrict__ flag, const unsigned long long* __restrict__ scan, PlocState* st) {
    const uint32_t m = st->m;
    const unsigned long long t = scan[m - 1] + flag[m - 1];
    st->merges = (uint32_t)(t >> 32);
    st->pad = (uint32_t)t;                 // clusters of the next round (applied by ploc_advance_kernel after the merge pass)
}

__global__ void __launch_bounds__(256) ploc_merge_kernel(const uint32_t* __restrict__ cid, const float* __restrict__ cbox, const uint32_t* __restrict__ nn,
                                                         const unsigned long long* __restrict__ flag, const unsigned long long* __restrict__ scan,
                                                         const PlocState* st, uint32_t* __restrict__ cid_out, float* __restrict__ cbox_out,
                                                         uint32_t* __restrict__ child, uint32_t* __restrict__ par_inner, uint32_t* __restrict__ par_leaf) {
    const uint32_t m = st->m, i = blockIdx.x * 256u + threadIdx.x;
    if (i >= m) return;
    const unsigned long long fl = flag[i];
    if (!(fl & 1ull)) return;                                   // merged into its partner's position
    const uint32_t pos = (uint32_t)scan[i];
    float b[6];
    for (int q = 0; q < 6; ++q) b[q] = cbox[6ull * i + q];
    uint32_t id = cid[i];
    if (fl >> 32) {
        const uint32_t j = nn[i], node = st->next_id - st->merges + (uint32_t)(scan[i] >> 32);
        const uint32_t L = id, R = cid[j];
        child[2 * node] = L; child[2 * node + 1] = R;
        if (L & LBVH_LEAF) par_leaf[L & ~LBVH_LEAF] = 2u * node; else par_inner[L] = 2u * node;
        if (R & LBVH_LEAF) par_leaf[R & ~LBVH_LEAF] = 2u * node + 1u; else par_inner[R] = 2u * node + 1u;
        for (int d = 0; d < 3; ++d) {
            b[2 * d] = min_std(b[2 * d], cbox[6ull * j + 2 * d]);
            b[2 * d + 1] = max_std(b[2 * d + 1], cbox[6ull * j + 2 * d + 1]);
        }
        id = node;
    }
    cid_out[pos] = id;
    for (int q = 0; q < 6; ++q) cbox_out[6ull * pos + q] = b[q];
}

__global__ void ploc_advance_kernel(PlocState* st) { st->next_id -= st->merges; st->m = st->pad; }

// initial clusters: leaf k (sorted position) with the box of triangle idx[k]
__global__ void __launch_bounds__(256) ploc_init_kernel(const uint32_t* __restrict__ idx, const float* __restrict__ box, uint32_t n,
                                                        uint32_t* __restrict__ cid, float* __restrict__ cbox) {
    const uint32_t k = blockIdx.x * 256u + threadIdx.x;
    if (k >= n) return;
    cid[k] = k | LBVH_LEAF;
    for (int q = 0; q < 6; ++q) cbox[6ull * k + q] = box[6ull * idx[k] + q];
}

// ---- leaf collapsing (SAH) ---------------------------------------------------------------------------------------------------
// Both builders above end with one triangle per leaf; the reference's SweepSah trees keep up to max_leaf_size triangles in a
// leaf where that is cheaper (top_down_builder.hpp:41, sweep_sah_builder.hpp: a range stays a leaf when no split beats
// count x area).  The same rule bottom-up: cost(node) = min(area x count x Ci, area x Ct + cost(left) + cost(right)); a node whose
// first term wins (and whose count fits) becomes a leaf, the topmost such node on a path wins.  The leaves are renumbered in
// depth-first order so that every subtree owns a contiguous range of slots, and the surviving inner nodes are renumbered in
// pre-order.
struct CollapseParams { float ct, ci; uint32_t max_leaf; };

__device__ __forceinline__ const float* bvhc_box(const LbvhNode* nodes, const uint32_t* par_inner, uint32_t p) {
    if (p == 0u) return nodes[0].bounds;
    const uint32_t up = par_inner[p];
    return nodes[1 + 2 * (up >> 1) + (up & 1u)].bounds;
}
__device__ __forceinline__ float bvhc_half_area(const float* b) {
    const float dx = b[1] - b[0], dy = b[3] - b[2], dz = b[5] - b[4];
    return dx * dy + dy * dz + dz * dx;
}

// one thread per leaf climbs while it is the second child to arrive: cnt[], cost[], collapse[] of the inner nodes
__global__ void __launch_bounds__(256) bvhc_cost_kernel(const uint32_t* __restrict__ child, const uint32_t* __restrict__ par_inner,
                                                        const uint32_t* __restrict__ par_leaf, const LbvhNode* __restrict__ nodes, int n, CollapseParams q,
                                                        uint32_t* arrived, uint32_t* cnt, float* cost, uint32_t* collapse, uint32_t* fl) {
    const int k = (int)(blockIdx.x * 256u + threadIdx.x);
    if (k >= n) return;
    uint32_t p = par_leaf[k] >> 1;
    for (;;) {
        if (atomicAdd(&arrived[p], 1u) == 0u) return;
        __threadfence();
        uint32_t c_n = 0, c_fl = 0; float c_cost = 0.0f;
        for (int s = 0; s < 2; ++s) {
            const uint32_t c = child[2 * p + s];
            if (c & LBVH_LEAF) { c_n += 1u; c_fl += 1u; c_cost += bvhc_half_area(nodes[1 + 2 * p + s].bounds) * q.ci; }
            else { c_n += __ldcg(cnt + c); c_fl += __ldcg(fl + c); c_cost += __ldcg(cost + c); }
        }
        const float a = bvhc_half_area(bvhc_box(nodes, par_inner, p));
        const float as_leaf = a * (float)c_n * q.ci, as_tree = a * q.ct + c_cost;
        const bool col = p != 0u && c_n <= q.max_leaf && as_leaf <= as_tree;
        __stcg(cnt + p, c_n); __stcg(cost + p, col ? as_leaf : as_tree); __stcg(collapse + p, col ? 1u : 0u);
        __stcg(fl + p, col ? 1u : c_fl);                        // leaves of the FINAL tree under p
        if (p == 0u) return;
        __threadfence();
        p = par_inner[p] >> 1;
    }
}

// lbvh_fit_kernel and bvhc_cost_kernel in one climb (both go from every leaf to the root behind the same arrival counters): the
// second thread to arrive at inner node p unites the children's boxes, stores the box where p lives, and evaluates p's leaf /
// subtree costs from it.  Same arithmetic as the two kernels, one pass over the tree less.
__global__ void __launch_bounds__(256) bvhc_fit_cost_kernel(const uint32_t* __restrict__ child, const uint32_t* __restrict__ par_inner,
                                                            const uint32_t* __restrict__ par_leaf, LbvhNode* nodes, int n, CollapseParams q,
                                                            uint32_t* arrived, uint32_t* cnt, float* cost, uint32_t* collapse, uint32_t* fl) {
    const int k = (int)(blockIdx.x * 256u + threadIdx.x);
    if (k >= n) return;
    uint32_t p = par_leaf[k] >> 1;
    for (;;) {
        if (atomicAdd(&arrived[p], 1u) == 0u) return;           // the sibling subtree is not finished yet
        __threadfence();
        const float* l = nodes[1 + 2 * p].bounds; const float* r = nodes[2 + 2 * p].bounds;
        float lb[6], rb[6], u[6];
        for (int d = 0; d < 6; ++d) { lb[d] = __ldcg(l + d); rb[d] = __ldcg(r + d); }
        for (int d = 0; d < 3; ++d) {                            // left.extend(right), bounding_box.hpp:23-27
            u[2 * d] = min_std(lb[2 * d], rb[2 * d]);
            u[2 * d + 1] = max_std(lb[2 * d + 1], rb[2 * d + 1]);
        }
        const uint32_t up = p == 0u ? 0u : par_inner[p];
        float* dst = p == 0u ? nodes[0].bounds : nodes[1 + 2 * (up >> 1) + (up & 1u)].bounds;
        for (int d = 0; d < 6; ++d) __stcg(dst + d, u[d]);
        uint32_t c_n = 0, c_fl = 0; float c_cost = 0.0f;
        for (int s = 0; s < 2; ++s) {
            const uint32_t c = child[2 * p + s];
            if (c & LBVH_LEAF) { c_n += 1u; c_fl += 1u; c_cost += bvhc_half_area(s ? rb : lb) * q.ci; }
            else { c_n += __ldcg(cnt + c); c_fl += __ldcg(fl + c); c_cost += __ldcg(cost + c); }
        }
        const float a = bvhc_half_area(u);
        const float as_leaf = a * (float)c_n * q.ci, as_tree = a * q.ct + c_cost;
        const bool col = p != 0u && c_n <= q.max_leaf && as_leaf <= as_tree;
        __stcg(cnt + p, c_n); __stcg(cost + p, col ? as_leaf : as_tree); __stcg(collapse + p, col ? 1u : 0u);
        __stcg(fl + p, col ? 1u : c_fl);                        // leaves of the FINAL tree under p
        if (p == 0u) return;
        __threadfence();
        p = up >> 1;
    }
}

// thread t < n: leaf t; thread n + i: inner node i.  Depth-first rank of the node's first leaf (sum, over the ancestors it
// hangs under on the right, of the left sibling's leaf count) and whether an ancestor collapses (then the node disappears).
// A surviving inner node also gets its number in the final tree: its PRE-ORDER index among the inner nodes = ancestors +
// sum over the subtrees hanging to the left of its path of (final leaves - 1) -- the order in which ceres_scene_create numbers
// the pairs of a host-built tree, so that a descent walks forward in memory.
__global__ void __launch_bounds__(256) bvhc_rank_kernel(const uint32_t* __restrict__ child, const uint32_t* __restrict__ par_inner,
                                                        const uint32_t* __restrict__ par_leaf, const uint32_t* __restrict__ cnt,
                                                        const uint32_t* __restrict__ collapse, const uint32_t* __restrict__ fl,
                                                        const uint32_t* __restrict__ idx, int n,
                                                        uint32_t* __restrict__ idx_out, uint32_t* __restrict__ rank_inner, uint32_t* __restrict__ keep,
                                                        uint32_t* __restrict__ newid) {
    const int t = (int)(blockIdx.x * 256u + threadIdx.x);
    if (t >= 2 * n - 1) return;
    const bool leaf = t < n;
    const uint32_t me = leaf ? (uint32_t)t : (uint32_t)(t - n);
    if (!leaf && me == 0u) { rank_inner[0] = 0u; keep[0] = 1u; newid[0] = 0u; return; }
    uint32_t w = leaf ? par_leaf[me] : par_inner[me];
    uint32_t rank = 0, pre = 0; bool dead = false;
    for (;;) {
        const uint32_t p = w >> 1;
        pre += 1u;                                               // the ancestor itself
        if (w & 1u) {
            const uint32_t l = child[2 * p];
            rank += (l & LBVH_LEAF) ? 1u : cnt[l];
            pre += (l & LBVH_LEAF) ? 0u : fl[l] - 1u;            // inner nodes of the final tree in the left sibling's subtree
        }
        if (collapse[p]) dead = true;
        if (p == 0u) break;
        w = par_inner[p];
    }
    if (leaf) idx_out[rank] = idx[me];
    else { rank_inner[me] = rank; keep[me] = (!dead && !collapse[me]) ? 1u : 0u; newid[me] = pre; }
}

// surviving inner node i (compact number newid[i]) writes its two children into the compact reference layout
__global__ void __launch_bounds__(256) bvhc_emit_kernel(const uint32_t* __restrict__ child, const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ collapse,
                                                        const uint32_t* __restrict__ rank_inner, const uint32_t* __restrict__ keep,
                                                        const uint32_t* __restrict__ newid, const LbvhNode* __restrict__ nodes, int n,
                                                        LbvhNode* __restrict__ out) {
    const int i = (int)(blockIdx.x * 256u + threadIdx.x);
    if (i >= n - 1 || !keep[i]) return;
    const uint32_t me = newid[i];
    if (i == 0) {
        for (int q = 0; q < 6; ++q) out[0].bounds[q] = nodes[0].bounds[q];
        out[0].primitive_count = 0u; out[0].first_child_or_primitive = 1u;
    }
    const uint32_t base = rank_inner[i];
    for (int s = 0; s < 2; ++s) {
        const uint32_t c = child[2 * i + s];
        LbvhNode nd = nodes[1 + 2 * i + s];                              // bounds of the child
        const uint32_t first_rank = s == 0 ? base : base + ((child[2 * i] & LBVH_LEAF) ? 1u : cnt[child[2 * i]]);
        if (c & LBVH_LEAF) { nd.primitive_count = 1u; nd.first_child_or_primitive = first_rank; }
        else if (collapse[c]) { nd.primitive_count = cnt[c]; nd.first_child_or_primitive = first_rank; }
        else { nd.primitive_count = 0u; nd.first_child_or_primitive = 1u + 2u * newid[c]; }
        out[1 + 2 * me + s] = nd;
    }
}

// ---- top-down binned SAH over the clusters (algorithm 2) ---------------------------------------------------------------------
// What the reference's own builders do on the host -- BinnedSahBuilder (static.cpp:106-107, binned_sah_builder.hpp) -- as a
// level-synchronous device pass: every open node ("task") bins the centroids of its members over its centroid bounds, takes
// the split with the smallest area(left) x count(left) + area(right) x count(right) over the three axes, and its members move
// to the two child tasks.  No member ever moves in memory: node_of[i] names the task of cluster i, bins and child bounds are
// accumulated with atomics on order-preserving uints.  profiles/tree_lab.py (a CPU restatement judged by the reference
// traverser's node-step counter on the dragon frame) is what picked this over the bottom-up clustering: 1.03 x the steps of the reference's SweepSah
// tree against 1.22 x, with 16 % fewer triangle tests -- and what sized the bins: 16 per axis, 4 for tasks of at most 16
// members (+0.1 % steps, a quarter of the memory and of the atomics of the deep levels).
// Same intermediate form as the builders above (child[], par_inner[], par_leaf[], inner node 0 = root): emit / fit / collapse
// are shared.  Inner ids are handed out from 1 upwards by an atomic counter; m clusters make exactly m - 1 inner nodes.
#define SAH_BINS 16                         // tasks of more than SAH_FEW members
#define SAH_BINS_FEW 4
#define SAH_FEW 16u
#define SAH_BIN_WORDS 7u                    // ~min x,y,z (as atomicMax of the complement), max x,y,z, count: all-zero = empty
#define SAH_TASK_WORDS (3u * SAH_BINS * SAH_BIN_WORDS)
#define SAH_TASK_WORDS_FEW (3u * SAH_BINS_FEW * SAH_BIN_WORDS)
#define SAH_NONE 0xFFFFFFFFu
#define SAH_FEW_SLOT 0x80000000u            // SahTask::slot: index into the level's list of small tasks (their own, narrower bins)
#define SAH_FALLBACK 0x80000000u

struct SahTask {
    uint32_t cb[6];                         // centroid bounds of the members: ~ord(min), ord(max) per axis (zero = empty)
    uint32_t count, imin, imax;             // members; smallest / largest member index (imin kept as the complement: zero = empty)
    uint32_t inner;                         // inner node id of this task (count >= 2), set by the link pass
    uint32_t parent;                        // parent inner id << 1 | side
    uint32_t slot;                          // index into this level's bins (| SAH_FEW_SLOT for a small task); SAH_NONE = closed
    uint32_t split;                         // axis << 8 | last bin of the left side; SAH_FALLBACK | first member index of the right side
    uint32_t child0;                        // task id of the left child (right = + 1)
    uint32_t pad_[2];
};
static_assert(sizeof(SahTask) == 64, "one task per 64 bytes");

struct SahState { uint32_t next_task, next_inner, n_open, n_open_few, error, pad_[3]; };

__device__ __forceinline__ float sah_centroid(const float* __restrict__ b, int d) { return mul(add(b[2 * d], b[2 * d + 1]), 0.5f); }
__device__ __forceinline__ int sah_bin(float c, float lo, float ext, int nb) {
    const int b = (int)(mul((float)nb, __fdiv_rn(sub(c, lo), ext)));
    return b < 0 ? 0 : b > nb - 1 ? nb - 1 : b;
}

// Warp-aggregated atomicMax / atomicAdd: the lanes of `peers` (a __match_any_sync group: same destination) reduce among
// themselves and their first lane issues the one atomic.  The clusters are in Morton order, so at the upper levels a whole warp
// shares one task and a few bins -- without this the same few words take an atomic from every one of the 10 M members
// (850 ms instead of 64 ms for the 10.5 M-triangle scene).
#ifndef SAH_AGGREGATE
#define SAH_AGGREGATE 1
#endif
__device__ __forceinline__ void sah_max(uint32_t* dst, uint32_t v, uint32_t peers, bool leader) {
#if SAH_AGGREGATE
    v = __reduce_max_sync(peers, v);
    if (leader) atomicMax(dst, v);
#else
    atomicMax(dst, v);
#endif
}
__device__ __forceinline__ void sah_add(uint32_t* dst, uint32_t v, uint32_t peers, bool leader) {
#if SAH_AGGREGATE
    v = __reduce_add_sync(peers, v);
    if (leader) atomicAdd(dst, v);
#else
    atomicAdd(dst, v);
#endif
}

// root task: every cluster is a member; `scene` = the ordered-uint bounds of the centres (lbvh_boxes_kernel), which are the
// root's centroid bounds when the clusters are the triangles themselves
__global__ void __launch_bounds__(256) sah_root_kernel(const unsigned int* __restrict__ scene, uint32_t m, SahTask* task, uint32_t* __restrict__ node_of) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= m) return;
    node_of[i] = 0u;
    if (i == 0) {
        SahTask t = {};
        for (int d = 0; d < 3; ++d) { t.cb[2 * d] = ~scene[2 * d]; t.cb[2 * d + 1] = scene[2 * d + 1]; }
        t.count = m; t.imin = ~0u; t.imax = m - 1u; t.inner = 0u; t.parent = 0u; t.slot = m > SAH_FEW ? 0u : SAH_FEW_SLOT;
        task[0] = t;
    }
}

// member i adds its box to the bins of its task; big / few = the level's two bin arrays in global memory, or (big only) a
// CTA's private copy
__device__ __forceinline__ void sah_bin_member(uint32_t i, const float* __restrict__ cbox, const SahTask* __restrict__ task,
                                               const uint32_t* __restrict__ node_of, uint32_t* big, uint32_t* few) {
    const SahTask& t = task[node_of[i]];
    const uint32_t slot = t.slot;
    if (slot == SAH_NONE) return;
    const bool is_few = (slot & SAH_FEW_SLOT) != 0u;
    const int nb = is_few ? SAH_BINS_FEW : SAH_BINS;
    uint32_t* base = is_few ? few + (size_t)(slot & ~SAH_FEW_SLOT) * SAH_TASK_WORDS_FEW : big + (size_t)slot * SAH_TASK_WORDS;
    float b[6];
    for (int q = 0; q < 6; ++q) b[q] = __ldg(cbox + 6ull * i + q);
    uint32_t ob[6];
    for (int d = 0; d < 3; ++d) { ob[2 * d] = ~lbvh_ord(b[2 * d]); ob[2 * d + 1] = lbvh_ord(b[2 * d + 1]); }
    for (int d = 0; d < 3; ++d) {
        const float lo = lbvh_unord(~t.cb[2 * d]), hi = lbvh_unord(t.cb[2 * d + 1]);
        const float ext = sub(hi, lo);
        if (!(ext > 0.0f)) continue;
        const uint32_t bin = (uint32_t)sah_bin(sah_centroid(b, d), lo, ext, nb);
        uint32_t* w = base + ((size_t)d * nb + bin) * SAH_BIN_WORDS;
#if SAH_AGGREGATE
        const uint32_t peers = __match_any_sync(__activemask(), ((unsigned long long)slot << 8) | bin);     // lanes that got here with the same bin
#else
        const uint32_t peers = 0u;
#endif
        const bool leader = (threadIdx.x & 31u) == (uint32_t)(__ffs(peers) - 1);
        for (int q = 0; q < 6; ++q) sah_max(w + q, ob[q], peers, leader);
        sah_add(w + 6, 1u, peers, leader);
    }
}

__global__ void __launch_bounds__(256) sah_bin_kernel(const float* __restrict__ cbox, uint32_t m, const SahTask* __restrict__ task,
                                                      const uint32_t* __restrict__ node_of, uint32_t* bins, uint32_t* bins_few) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < m) sah_bin_member(i, cbox, task, node_of, bins, bins_few);
}

// The first levels have a handful of tasks with millions of members each: every CTA of a persistent grid bins into its own
// copy in shared memory and adds the copy to the global bins once (n_open <= SAH_SMALL_LEVEL big tasks, no small ones: at
// most 42 KB).  Levels 1-6 of the 10.5 M-triangle scene: 25.4 ms -> 6.1 ms.
#define SAH_SMALL_LEVEL 32u
__global__ void __launch_bounds__(256) sah_bin_small_kernel(const float* __restrict__ cbox, uint32_t m, const SahTask* __restrict__ task,
                                                            const uint32_t* __restrict__ node_of, uint32_t* bins, uint32_t n_open) {
    extern __shared__ uint32_t sah_sh[];
    const uint32_t words = n_open * SAH_TASK_WORDS;
    for (uint32_t k = threadIdx.x; k < words; k += 256u) sah_sh[k] = 0u;
    __syncthreads();
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < m; i += gridDim.x * 256u) sah_bin_member(i, cbox, task, node_of, sah_sh, nullptr);
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < words; k += 256u) {
        const uint32_t v = sah_sh[k];
        if (!v) continue;
        if (k % SAH_BIN_WORDS == 6u) atomicAdd(bins + k, v); else atomicMax(bins + k, v);
    }
}

// one thread per open task: the best split over 3 axes x (NB - 1) planes; allocates the two child tasks
template <int NB>
__global__ void __launch_bounds__(128) sah_split_kernel(const uint32_t* __restrict__ open, uint32_t n_open, SahTask* task, const uint32_t* __restrict__ bins,
                                                        SahState* st) {
    const uint32_t s = blockIdx.x * 128u + threadIdx.x;
    if (s >= n_open) return;
    const uint32_t tid = open[s];
    SahTask& t = task[tid];
    float best = 3.4e38f; uint32_t best_split = SAH_NONE;
    for (int d = 0; d < 3; ++d) {
        const float lo = lbvh_unord(~t.cb[2 * d]), hi = lbvh_unord(t.cb[2 * d + 1]);
        if (!(sub(hi, lo) > 0.0f)) continue;
        const uint32_t* w = bins + (size_t)s * (3u * NB * SAH_BIN_WORDS) + (size_t)d * NB * SAH_BIN_WORDS;
        float left_cost[NB - 1]; uint32_t left_n[NB - 1];
        uint32_t acc[6] = {0, 0, 0, 0, 0, 0}, n = 0;
        for (int k = 0; k < NB - 1; ++k) {
            for (int q = 0; q < 6; ++q) acc[q] = max(acc[q], w[k * SAH_BIN_WORDS + q]);
            n += w[k * SAH_BIN_WORDS + 6];
            float a = 0.0f;
            if (n) {
                const float dx = lbvh_unord(acc[1]) - lbvh_unord(~acc[0]), dy = lbvh_unord(acc[3]) - lbvh_unord(~acc[2]), dz = lbvh_unord(acc[5]) - lbvh_unord(~acc[4]);
                a = dx * dy + dy * dz + dz * dx;
            }
            left_cost[k] = a * (float)n; left_n[k] = n;
        }
        for (int q = 0; q < 6; ++q) acc[q] = 0;
        n = 0;
        for (int k = NB - 2; k >= 0; --k) {
            for (int q = 0; q < 6; ++q) acc[q] = max(acc[q], w[(k + 1) * SAH_BIN_WORDS + q]);
            n += w[(k + 1) * SAH_BIN_WORDS + 6];
            if (!n || !left_n[k]) continue;
            const float dx = lbvh_unord(acc[1]) - lbvh_unord(~acc[0]), dy = lbvh_unord(acc[3]) - lbvh_unord(~acc[2]), dz = lbvh_unord(acc[5]) - lbvh_unord(~acc[4]);
            const float c = left_cost[k] + (dx * dy + dy * dz + dz * dx) * (float)n;
            if (c < best) { best = c; best_split = ((uint32_t)d << 8) | (uint32_t)k; }
        }
    }
    if (best_split == SAH_NONE) {
        // every centroid on one point (or non-finite costs): cut the member list in the middle of its index range
        const uint32_t imin = ~t.imin, imax = t.imax;
        best_split = SAH_FALLBACK | (imin + (imax - imin + 1u) / 2u);
    }
    t.split = best_split;
    const uint32_t c0 = atomicAdd(&st->next_task, 2u);
    t.child0 = c0;
    for (uint32_t k = 0; k < 2u; ++k) {
        SahTask c = {};
        c.parent = (t.inner << 1) | k; c.slot = SAH_NONE; c.inner = SAH_NONE;
        task[c0 + k] = c;
    }
}

// The first 9 words of a SahTask -- cb[6], count, imin, imax -- are what the members of a new task accumulate: max for all
// but the count.
#define SAH_ACC_WORDS 9u
static_assert(offsetof(SahTask, count) == 24 && offsetof(SahTask, imin) == 28 && offsetof(SahTask, imax) == 32, "accumulated words first");

// member i moves to the child task its parent's split sends it to; SMALL: the children's words are accumulated in shared memory
template <bool SMALL>
__device__ __forceinline__ void sah_assign_member(uint32_t i, const float* __restrict__ cbox, SahTask* task, uint32_t* __restrict__ node_of, uint32_t* sh) {
    const SahTask& t = task[node_of[i]];
    if (t.slot == SAH_NONE) return;
    const float* b = cbox + 6ull * i;
    uint32_t side;
    if (t.split & SAH_FALLBACK) side = i >= (t.split & ~SAH_FALLBACK) ? 1u : 0u;
    else {
        const int d = (int)(t.split >> 8), k = (int)(t.split & 0xFFu);
        const float lo = lbvh_unord(~t.cb[2 * d]), hi = lbvh_unord(t.cb[2 * d + 1]);
        side = sah_bin(sah_centroid(b, d), lo, sub(hi, lo), (t.slot & SAH_FEW_SLOT) ? SAH_BINS_FEW : SAH_BINS) > k ? 1u : 0u;
    }
    const uint32_t ct = t.child0 + side;
    node_of[i] = ct;
    uint32_t* acc = SMALL ? sh + (2u * t.slot + side) * SAH_ACC_WORDS : reinterpret_cast<uint32_t*>(&task[ct]);
#if SAH_AGGREGATE
    const uint32_t peers = __match_any_sync(__activemask(), ct);
#else
    const uint32_t peers = 0u;
#endif
    const bool leader = (threadIdx.x & 31u) == (uint32_t)(__ffs(peers) - 1);
    for (int d = 0; d < 3; ++d) {
        const uint32_t o = lbvh_ord(sah_centroid(b, d));
        sah_max(acc + 2 * d, ~o, peers, leader); sah_max(acc + 2 * d + 1, o, peers, leader);
    }
    sah_add(acc + 6, 1u, peers, leader);
    sah_max(acc + 7, ~i, peers, leader); sah_max(acc + 8, i, peers, leader);
}

__global__ void __launch_bounds__(256) sah_assign_kernel(const float* __restrict__ cbox, uint32_t m, SahTask* task, uint32_t* __restrict__ node_of) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < m) sah_assign_member<false>(i, cbox, task, node_of, nullptr);
}

__global__ void __launch_bounds__(256) sah_assign_small_kernel(const float* __restrict__ cbox, uint32_t m, SahTask* task, uint32_t* __restrict__ node_of,
                                                               const uint32_t* __restrict__ open, uint32_t n_open) {
    extern __shared__ uint32_t sah_sh[];
    const uint32_t words = 2u * n_open * SAH_ACC_WORDS;
    for (uint32_t k = threadIdx.x; k < words; k += 256u) sah_sh[k] = 0u;
    __syncthreads();
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < m; i += gridDim.x * 256u) sah_assign_member<true>(i, cbox, task, node_of, sah_sh);
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < words; k += 256u) {
        const uint32_t v = sah_sh[k];
        if (!v) continue;
        const uint32_t c = k / SAH_ACC_WORDS, f = k - c * SAH_ACC_WORDS;
        uint32_t* dst = reinterpret_cast<uint32_t*>(&task[task[open[c >> 1]].child0 + (c & 1u)]) + f;
        if (f == 6u) atomicAdd(dst, v); else atomicMax(dst, v);
    }
}

__device__ __forceinline__ void sah_set_parent(uint32_t ptr, uint32_t where, uint32_t* __restrict__ par_inner, uint32_t* __restrict__ par_leaf) {
    if (ptr & LBVH_LEAF) par_leaf[ptr & ~LBVH_LEAF] = where; else par_inner[ptr] = where;
}

// one thread per task created at this level: hook the task into the tree and, with 3 members or more, open it for the next level
__global__ void __launch_bounds__(256) sah_link_kernel(uint32_t first, uint32_t count, SahTask* task, const uint32_t* __restrict__ cid,
                                                       uint32_t* __restrict__ child, uint32_t* __restrict__ par_inner, uint32_t* __restrict__ par_leaf,
                                                       uint32_t* __restrict__ open_next, uint32_t* __restrict__ open_next_few, SahState* st) {
    const uint32_t j = blockIdx.x * 256u + threadIdx.x;
    if (j >= count) return;
    SahTask& t = task[first + j];
    const uint32_t where = t.parent, n = t.count;
    if (n == 0u) { st->error = 1u; return; }
    if (n == 1u) {
        const uint32_t ptr = cid[t.imax];
        child[where] = ptr;
        sah_set_parent(ptr, where, par_inner, par_leaf);
        return;
    }
    const uint32_t id = atomicAdd(&st->next_inner, 1u);
    t.inner = id;
    child[where] = id;
    par_inner[id] = where;
    if (n == 2u) {
        const uint32_t a = cid[~t.imin], b = cid[t.imax];
        child[2 * id] = a; child[2 * id + 1] = b;
        sah_set_parent(a, 2 * id, par_inner, par_leaf); sah_set_parent(b, 2 * id + 1, par_inner, par_leaf);
        return;
    }
    if (n > SAH_FEW) {
        const uint32_t slot = atomicAdd(&st->n_open, 1u);
        open_next[slot] = first + j;
        t.slot = slot;
    } else {
        const uint32_t slot = atomicAdd(&st->n_open_few, 1u);
        open_next_few[slot] = first + j;
        t.slot = slot | SAH_FEW_SLOT;
    }
}

// ---- CTA-level accumulation for the bin and assign passes ----------------------------------------------------------------------
// ncu of the warp-aggregated passes on the 10.5 M-triangle scene: instruction-bound (a redux whose lanes carry different member
// masks runs once per group: 900 warp instructions per warp at a mid level, 2 400 at a deep one, 9 of 32 lanes active) and then
// still 7 global atomics per group.  Here a CTA takes SAH_CTA_MEMBERS consecutive members, gives every destination record it
// meets (a task's bins; a child task's 9 accumulated words) a region of a shared-memory pool through a small hash table,
// accumulates with shared-memory atomics and adds each non-zero word to global memory once.  The members are in Morton order,
// so a CTA meets few records: one to three tasks at the mid levels, ~60 small ones at the deepest.  Records that do not fit
// the pool go straight to global memory.  Results are sums / maxima: identical to the other schemes bit for bit.
#define SAH_CTA_MEMBERS 512u
#define SAH_HASH 1024u
#define SAH_POOL_WORDS 6144u
#define SAH_EMPTY 0xFFFFFFFFu
#define SAH_OVERFLOW 0xFFFFFFFFu

struct SahCta {
    uint32_t keys[SAH_HASH];
    uint32_t offs[SAH_HASH];
    uint32_t list[SAH_CTA_MEMBERS];       // hash entries in use
    uint32_t pool[SAH_POOL_WORDS];
    uint32_t bump, n_list;
};

__device__ __forceinline__ void sah_cta_begin(SahCta& c) {
    for (uint32_t k = threadIdx.x; k < SAH_HASH; k += 256u) c.keys[k] = SAH_EMPTY;
    for (uint32_t k = threadIdx.x; k < SAH_POOL_WORDS; k += 256u) c.pool[k] = 0u;
    if (threadIdx.x == 0) { c.bump = 0u; c.n_list = 0u; }
    __syncthreads();
}
__device__ __forceinline__ uint32_t sah_cta_insert(SahCta& c, uint32_t key) {       // returns the hash entry of `key`
    uint32_t h = (key * 2654435761u) >> 22;                                        // 10 bits
    for (;;) {
        const uint32_t old = atomicCAS(&c.keys[h], SAH_EMPTY, key);
        if (old == SAH_EMPTY || old == key) return h;
        h = (h + 1u) & (SAH_HASH - 1u);
    }
}
// after every member has inserted its key: each used entry gets `words(key)` words of the pool (or SAH_OVERFLOW)
template <class Words>
__device__ __forceinline__ void sah_cta_allocate(SahCta& c, Words words) {
    __syncthreads();
    for (uint32_t e = threadIdx.x; e < SAH_HASH; e += 256u) {
        const uint32_t key = c.keys[e];
        if (key == SAH_EMPTY) continue;
        const uint32_t n = words(key), off = atomicAdd(&c.bump, n);
        if (off + n <= SAH_POOL_WORDS) { c.offs[e] = off; c.list[atomicAdd(&c.n_list, 1u)] = e; }
        else c.offs[e] = SAH_OVERFLOW;
    }
    __syncthreads();
}
// after the accumulation: a warp per record adds the non-zero words to dst(key); word w of a record is a count when is_add(w)
template <class Words, class Dst, class IsAdd>
__device__ __forceinline__ void sah_cta_flush(SahCta& c, Words words, Dst dst, IsAdd is_add) {
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t q = threadIdx.x >> 5; q < c.n_list; q += 8u) {
        const uint32_t e = c.list[q], key = c.keys[e], off = c.offs[e], n = words(key);
        uint32_t* g = dst(key);
        for (uint32_t w = lane; w < n; w += 32u) {
            const uint32_t v = c.pool[off + w];
            if (!v) continue;
            if (is_add(w)) atomicAdd(g + w, v); else atomicMax(g + w, v);
        }
    }
}

__global__ void __launch_bounds__(256) sah_bin_cta_kernel(const float* __restrict__ cbox, uint32_t m, const SahTask* __restrict__ task,
                                                          const uint32_t* __restrict__ node_of, uint32_t* bins, uint32_t* bins_few) {
    __shared__ SahCta c;
    sah_cta_begin(c);
    constexpr uint32_t PER = SAH_CTA_MEMBERS / 256u;
    uint32_t tsk[PER], entry[PER];
    for (uint32_t u = 0; u < PER; ++u) {
        const uint32_t i = blockIdx.x * SAH_CTA_MEMBERS + u * 256u + threadIdx.x;
        tsk[u] = SAH_NONE; entry[u] = 0u;
        if (i >= m) continue;
        const uint32_t t = node_of[i], slot = task[t].slot;
        if (slot == SAH_NONE) continue;
        tsk[u] = t; entry[u] = sah_cta_insert(c, slot);
    }
    auto words = [](uint32_t slot) { return (slot & SAH_FEW_SLOT) ? SAH_TASK_WORDS_FEW : SAH_TASK_WORDS; };
    auto dst = [&](uint32_t slot) { return (slot & SAH_FEW_SLOT) ? bins_few + (size_t)(slot & ~SAH_FEW_SLOT) * SAH_TASK_WORDS_FEW : bins + (size_t)slot * SAH_TASK_WORDS; };
    sah_cta_allocate(c, words);
    for (uint32_t u = 0; u < PER; ++u) {
        if (tsk[u] == SAH_NONE) continue;
        const uint32_t i = blockIdx.x * SAH_CTA_MEMBERS + u * 256u + threadIdx.x;
        const SahTask& t = task[tsk[u]];
        const uint32_t slot = t.slot, off = c.offs[entry[u]];
        const int nb = (slot & SAH_FEW_SLOT) ? SAH_BINS_FEW : SAH_BINS;
        uint32_t* base = off != SAH_OVERFLOW ? c.pool + off : dst(slot);
        float b[6];
        for (int q = 0; q < 6; ++q) b[q] = __ldg(cbox + 6ull * i + q);
        for (int d = 0; d < 3; ++d) {
            const float lo = lbvh_unord(~t.cb[2 * d]), hi = lbvh_unord(t.cb[2 * d + 1]);
            const float ext = sub(hi, lo);
            if (!(ext > 0.0f)) continue;
            uint32_t* w = base + ((size_t)d * nb + (uint32_t)sah_bin(sah_centroid(b, d), lo, ext, nb)) * SAH_BIN_WORDS;
            for (int q = 0; q < 3; ++q) { atomicMax(w + 2 * q, ~lbvh_ord(b[2 * q])); atomicMax(w + 2 * q + 1, lbvh_ord(b[2 * q + 1])); }
            atomicAdd(w + 6, 1u);
        }
    }
    sah_cta_flush(c, words, dst, [](uint32_t w) { return w % SAH_BIN_WORDS == 6u; });
}

__global__ void __launch_bounds__(256) sah_assign_cta_kernel(const float* __restrict__ cbox, uint32_t m, SahTask* task, uint32_t* __restrict__ node_of) {
    __shared__ SahCta c;
    sah_cta_begin(c);
    constexpr uint32_t PER = SAH_CTA_MEMBERS / 256u;
    uint32_t child[PER], entry[PER];
    for (uint32_t u = 0; u < PER; ++u) {
        const uint32_t i = blockIdx.x * SAH_CTA_MEMBERS + u * 256u + threadIdx.x;
        child[u] = SAH_NONE; entry[u] = 0u;
        if (i >= m) continue;
        const SahTask& t = task[node_of[i]];
        if (t.slot == SAH_NONE) continue;
        const float* b = cbox + 6ull * i;
        uint32_t side;
        if (t.split & SAH_FALLBACK) side = i >= (t.split & ~SAH_FALLBACK) ? 1u : 0u;
        else {
            const int d = (int)(t.split >> 8), k = (int)(t.split & 0xFFu);
            const float lo = lbvh_unord(~t.cb[2 * d]), hi = lbvh_unord(t.cb[2 * d + 1]);
            side = sah_bin(sah_centroid(b, d), lo, sub(hi, lo), (t.slot & SAH_FEW_SLOT) ? SAH_BINS_FEW : SAH_BINS) > k ? 1u : 0u;
        }
        child[u] = t.child0 + side;
        node_of[i] = child[u];
        entry[u] = sah_cta_insert(c, child[u]);
    }
    auto words = [](uint32_t) { return SAH_ACC_WORDS; };
    auto dst = [&](uint32_t ct) { return reinterpret_cast<uint32_t*>(&task[ct]); };
    sah_cta_allocate(c, words);
    for (uint32_t u = 0; u < PER; ++u) {
        if (child[u] == SAH_NONE) continue;
        const uint32_t i = blockIdx.x * SAH_CTA_MEMBERS + u * 256u + threadIdx.x;
        const uint32_t off = c.offs[entry[u]];
        uint32_t* acc = off != SAH_OVERFLOW ? c.pool + off : dst(child[u]);
        const float* b = cbox + 6ull * i;
        for (int d = 0; d < 3; ++d) {
            const uint32_t o = lbvh_ord(sah_centroid(b, d));
            atomicMax(acc + 2 * d, ~o); atomicMax(acc + 2 * d + 1, o);
        }
        atomicAdd(acc + 6, 1u);
        atomicMax(acc + 7, ~i); atomicMax(acc + 8, i);
    }
    sah_cta_flush(c, words, dst, [](uint32_t w) { return w == 6u; });
}
