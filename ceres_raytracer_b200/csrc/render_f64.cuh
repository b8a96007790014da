// render_f64.cuh -- the render path for Scalar = double (`anim -d`, include/anim.cpp:154-155),
// included at the end of ceres_b200.cu.
//
// The reference's render<double>() (render.hpp:86-156) is the same loop as render<float>() with every Scalar widened:
// 64-byte nodes with uint64 indices (bvh.hpp:20-28), 96-byte triangles, 72-byte normal triples, std::fma in the slab
// test (utilities.hpp:46-49), DBL_EPSILON in safe_inverse (vector.hpp:69-74), tmax = DBL_MAX (ray.hpp:17-20) -- and the
// shading still in float where render.hpp says float (:46-84).  This is a completeness path, not the headline: one
// kernel, one thread per pixel (a warp owns an 8x4 tile), primary traversal, shadow traversal (any-hit: only bool(hit)
// is consumed, render.hpp:137-138) and shading in sequence.  Every operation through the *_rn double intrinsics, in the
// reference's source order, so that hit ids are bit-exact against the strict double build of the reference.
//
// Device layout: pair p = the two children of an inner node, 128 bytes: child c at 64 c = double bounds[6], uint32
// primitive_count, uint32 first (leaf: first slot; inner: pair index), 8 bytes padding.  Triangles (12 doubles) and
// normals (9 doubles) are permuted into leaf-slot order, slot_id maps a slot back to the primitive index.

namespace f64 {

__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double rcp(double a) { return __ddiv_rn(1.0, a); }
__device__ __forceinline__ double dot3(double ax, double ay, double az, double bx, double by, double bz) {   // vector.hpp:134-141
    return add(add(mul(ax, bx), mul(ay, by)), mul(az, bz));
}
struct D3 { double x, y, z; };
__device__ __forceinline__ D3 normalize(D3 a) {                                                              // vector.hpp:143-154
    const double inv = rcp(__dsqrt_rn(dot3(a.x, a.y, a.z, a.x, a.y, a.z)));
    return {mul(a.x, inv), mul(a.y, inv), mul(a.z, inv)};
}

struct Child { double b[6]; uint32_t count, first; uint32_t pad[2]; };
static_assert(sizeof(Child) == 64, "128-byte pairs");

struct Scene64 {
    const Child* pairs;           // 2 per pair
    const double* tris;           // 12 per slot
    const double* norms;          // 9 per slot
    const uint32_t* slot_id;
    uint32_t root_leaf_count;
    uint32_t n_pairs, n_slots;    // bounds (checked build)
    unsigned long long* check;    // checked build: CHECK_* bits; null otherwise
};

struct Frame64 {
    double eye[3], dir[3], iu[3], iv[3], sun[3];
    uint32_t width, height, flags;
    uint32_t x0, y0, x1, y1;        // columns [x0, x1) x LOCAL rows [y0, y1) are rendered by this launch (a row band of the screen rectangle)
    uint32_t blk_rows, sh_index, sh_count;   // row-block shard (ceres_shard): local row -> frame row; count <= 1: local row = frame row
    float* rgb; uint32_t* ids;      // the colours are float-valued (render.hpp:56-84): 3 floats per pixel on the device
    unsigned long long* counters;   // rays beyond W*H (= primary hits), hits
    int* rect;                      // min i, min j, max i, max j over the pixels whose primary ray hit
};

struct Ray64 { D3 o, d; double tmax; int oct[3]; double inv[3], so[3]; };

__device__ __forceinline__ void ray_init(Ray64& r) {            // node_intersectors.hpp:20-26,89-94
    const double d[3] = {r.d.x, r.d.y, r.d.z}, o[3] = {r.o.x, r.o.y, r.o.z};
    for (int k = 0; k < 3; ++k) {
        r.oct[k] = signbit(d[k]) ? 1 : 0;
        r.inv[k] = rcp(fabs(d[k]) < DBL_EPSILON ? copysign(DBL_EPSILON, d[k]) : d[k]);
        r.so[k] = mul(-o[k], r.inv[k]);
    }
}

__device__ __forceinline__ void slab(const Ray64& r, const Child& c, double& entry, double& exit_) {   // node_intersectors.hpp:34-47,96-100
    const double e0 = __fma_rn(c.b[0 + r.oct[0]], r.inv[0], r.so[0]);
    const double e1 = __fma_rn(c.b[2 + r.oct[1]], r.inv[1], r.so[1]);
    const double e2 = __fma_rn(c.b[4 + r.oct[2]], r.inv[2], r.so[2]);
    const double x0 = __fma_rn(c.b[1 - r.oct[0]], r.inv[0], r.so[0]);
    const double x1 = __fma_rn(c.b[3 - r.oct[1]], r.inv[1], r.so[1]);
    const double x2 = __fma_rn(c.b[5 - r.oct[2]], r.inv[2], r.so[2]);
    auto rmax = [](double x, double y) { return x > y ? x : y; };          // utilities.hpp:57-67
    auto rmin = [](double x, double y) { return x < y ? x : y; };
    entry = rmax(e0, rmax(e1, rmax(e2, 0.0)));                              // tmin = 0 for both ray kinds
    exit_ = rmin(x0, rmin(x1, rmin(x2, r.tmax)));
}

struct Hit64 { uint32_t slot; double u, v; bool hit; };

// triangle.hpp:95-115 for the slots [first, first + count), ascending (single_ray_traverser.hpp:41-63)
template <bool ANY>
__device__ __forceinline__ void leaf(const Scene64& sc, Ray64& r, uint32_t first, uint32_t count, Hit64& best) {
    for (uint32_t s0 = first; s0 < first + count; ++s0) {
        uint32_t s = s0;
        CERES_CLAMP(s, sc.n_slots, CHECK_SLOT_INDEX, sc);
        const double* t = sc.tris + 12ull * s;
        const double cx = sub(t[0], r.o.x), cy = sub(t[1], r.o.y), cz = sub(t[2], r.o.z);
        const double rx = sub(mul(r.d.y, cz), mul(r.d.z, cy)), ry = sub(mul(r.d.z, cx), mul(r.d.x, cz)), rz = sub(mul(r.d.x, cy), mul(r.d.y, cx));
        const double inv_det = rcp(dot3(t[9], t[10], t[11], r.d.x, r.d.y, r.d.z));
        const double u = mul(dot3(rx, ry, rz, t[6], t[7], t[8]), inv_det);
        const double v = mul(dot3(rx, ry, rz, t[3], t[4], t[5]), inv_det);
        const double w = sub(sub(1.0, u), v);
        if (u >= 0 && v >= 0 && w >= 0) {
            const double tt = mul(dot3(t[9], t[10], t[11], cx, cy, cz), inv_det);
            if (tt >= 0.0 && tt <= r.tmax) {
                best.hit = true; best.slot = s; best.u = u; best.v = v;
                r.tmax = tt;
                if (ANY) return;
            }
        }
    }
}

// single_ray_traverser.hpp:65-126: ordered, eager leaves, far child pushed; ties keep the left child in front
template <bool ANY>
__device__ Hit64 traverse(const Scene64& sc, Ray64& r) {
    Hit64 best; best.hit = false; best.slot = 0xFFFFFFFFu; best.u = best.v = 0.0;
    if (sc.root_leaf_count) { leaf<ANY>(sc, r, 0u, sc.root_leaf_count, best); return best; }
    ray_init(r);
    uint32_t stack[64];
    int sp = 0;
    uint32_t pair = 0;
    const uint32_t NONE = 0xFFFFFFFFu;
    for (;;) {
        CERES_CLAMP(pair, sc.n_pairs, CHECK_PAIR_INDEX, sc);
        const Child& l = sc.pairs[2ull * pair];
        const Child& rt = sc.pairs[2ull * pair + 1];
        double le, lx, re, rx;
        slab(r, l, le, lx);
        slab(r, rt, re, rx);                      // before the left leaf shrinks tmax (:86-87)
        uint32_t nl = NONE, nr = NONE;
        if (le <= lx) { if (l.count) leaf<ANY>(sc, r, l.first, l.count, best); else nl = l.first; }
        if (ANY && best.hit) return best;
        if (re <= rx) { if (rt.count) leaf<ANY>(sc, r, rt.first, rt.count, best); else nr = rt.first; }
        if (ANY && best.hit) return best;
        if (nl != NONE) {
            if (nr != NONE) {
                if (le > re) { const uint32_t t = nl; nl = nr; nr = t; }
                CERES_CK(sp < 64, CHECK_STACK_DEPTH, sc);
                stack[sp++] = nr;
            }
            pair = nl;
        } else if (nr != NONE) {
            pair = nr;
        } else {
            if (sp == 0) break;
            pair = stack[--sp];
        }
    }
    return best;
}

__device__ __forceinline__ float clampf01(float x) { return x < 0.0f ? 0.0f : (1.0f < x ? 1.0f : x); }

// render.hpp:56-84 with Scalar = double: float diffuse / specular / colour accumulators, double weights
__device__ __forceinline__ void smooth_shading(D3 sun_line, const double* nm, D3 view, double u, double v, float* c) {
    const double w[3] = {u, v, sub(sub(1.0, u), v)};
    const D3 mv = {mul(view.x, -1.0), mul(view.y, -1.0), mul(view.z, -1.0)};
    c[0] = c[1] = c[2] = 0.0f;
    for (int k = 0; k < 3; ++k) {
        const double nx = nm[3 * k], ny = nm[3 * k + 1], nz = nm[3 * k + 2];
        const float lam = (float)fabs(dot3(sun_line.x, sun_line.y, sun_line.z, nx, ny, nz));
        const float diffuse = __fmul_rn(0.5f, lam);
        const D3 h = normalize({add(sun_line.x, mv.x), add(sun_line.y, mv.y), add(sun_line.z, mv.z)});
        const double b = dot3(nx, ny, nz, h.x, h.y, h.z);
        const double b2 = mul(b, b), b3 = mul(b2, b), b6 = mul(b3, b3), b12 = mul(b6, b6);      // pow(b, 24)
        const float specular = __fmul_rn(0.8f, (float)mul(b12, b12));
        const float base = __fadd_rn(0.2f, diffuse);
        const float col[3] = {clampf01(__fadd_rn(__fmul_rn(base, 0.5f), specular)), clampf01(__fadd_rn(__fmul_rn(base, 0.0f), specular)),
                              clampf01(__fadd_rn(__fmul_rn(base, 0.8f), specular))};
        for (int q = 0; q < 3; ++q) c[q] = (float)add((double)c[q], mul(w[k], (double)col[q]));
    }
}

// One thread per pixel; a warp covers an 8 x 4 tile, a 128-thread CTA a 16 x 8 block.
__global__ void __launch_bounds__(128) render_kernel(const Scene64 sc, const Frame64 f) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t i = f.x0 + blockIdx.x * 16u + (warp & 1u) * 8u + (lane & 7u);
    const uint32_t lr = f.y0 + blockIdx.y * 8u + (warp >> 1) * 4u + (lane >> 3);
    uint32_t j = lr;                                                   // frame row of this local row
    if (f.sh_count > 1u) { const uint32_t blk = lr / f.blk_rows; j = (blk * f.sh_count + f.sh_index) * f.blk_rows + (lr - blk * f.blk_rows); }
    unsigned int primary_hit = 0, hits = 0;
    if (i < f.x1 && lr < f.y1) {
        const size_t pix = (size_t)f.width * lr + i;                  // device buffers are packed by local row
        const double u = sub(__ddiv_rn(mul(2.0, add((double)i, 0.5)), (double)f.width), 1.0);      // render.hpp:109-110
        const double v = sub(__ddiv_rn(mul(2.0, add((double)j, 0.5)), (double)f.height), 1.0);
        const D3 view = normalize({add(add(mul(f.iu[0], u), mul(f.iv[0], v)), f.dir[0]), add(add(mul(f.iu[1], u), mul(f.iv[1], v)), f.dir[1]),
                                   add(add(mul(f.iu[2], u), mul(f.iv[2], v)), f.dir[2])});
        Ray64 ray; ray.o = {f.eye[0], f.eye[1], f.eye[2]}; ray.d = view; ray.tmax = DBL_MAX;
        const Hit64 hit = traverse<false>(sc, ray);
        float rgb[3] = {0.0f, 0.0f, 0.0f};
        if (hit.hit) {
            primary_hit = 1; hits = 1;
            const double* t = sc.tris + 12ull * hit.slot;
            const D3 normal = normalize({t[9], t[10], t[11]});
            const double hu = hit.u, hv = hit.v, hw = sub(sub(1.0, hu), hv);
            double P[3];
            for (int k = 0; k < 3; ++k) {                                     // u p0 + v p1() + (1-u-v) p2(), render.hpp:129
                const double p1 = sub(t[k], t[3 + k]), p2 = add(t[k], t[6 + k]);
                P[k] = add(add(mul(t[k], hu), mul(p1, hv)), mul(p2, hw));
            }
            P[0] = add(P[0], mul(normal.x, -0.00001)); P[1] = add(P[1], mul(normal.y, -0.00001)); P[2] = add(P[2], mul(normal.z, -0.00001));
            const D3 sun_line = normalize({sub(f.sun[0], P[0]), sub(f.sun[1], P[1]), sub(f.sun[2], P[2])});
            bool occluded = false;
            if (!(f.flags & CERES_NO_SHADOWS)) {
                Ray64 sray; sray.o = {P[0], P[1], P[2]}; sray.d = sun_line; sray.tmax = DBL_MAX;
                occluded = traverse<true>(sc, sray).hit;
            }
            if (!occluded) smooth_shading(sun_line, sc.norms + 9ull * hit.slot, view, hu, hv, rgb);
            else hits = 2;
        }
        f.rgb[3 * pix] = rgb[0]; f.rgb[3 * pix + 1] = rgb[1]; f.rgb[3 * pix + 2] = rgb[2];
        if (f.ids) f.ids[pix] = hit.hit ? sc.slot_id[hit.slot] : 0xFFFFFFFFu;
    }
    const unsigned int shadow_rays = (f.flags & CERES_NO_SHADOWS) ? 0u : primary_hit;
    const unsigned int r = __reduce_add_sync(0xFFFFFFFFu, shadow_rays), h = __reduce_add_sync(0xFFFFFFFFu, hits);
    if (h) {        // warp-uniform: some lane hit -> widen the frame's hit rectangle
        const int x0 = __reduce_min_sync(0xFFFFFFFFu, primary_hit ? (int)i : 0x7FFFFFFF), y0 = __reduce_min_sync(0xFFFFFFFFu, primary_hit ? (int)j : 0x7FFFFFFF);
        const int x1 = __reduce_max_sync(0xFFFFFFFFu, primary_hit ? (int)i : -1), y1 = __reduce_max_sync(0xFFFFFFFFu, primary_hit ? (int)j : -1);
        if (lane == 0) {
            atomicAdd(&f.counters[0], (unsigned long long)r); atomicAdd(&f.counters[1], (unsigned long long)h);
            atomicMin(&f.rect[0], x0); atomicMin(&f.rect[1], y0); atomicMax(&f.rect[2], x1); atomicMax(&f.rect[3], y1);
        }
    }
}

}  // namespace f64

// ---- host side ---------------------------------------------------------------------------------------------------
struct ceres_scene_f64 {
    int device = 0;
    f64::Scene64 dev{};
    void* d_pairs = nullptr; double* d_tris = nullptr; double* d_norms = nullptr; uint32_t* d_slot_id = nullptr;
    float* d_rgb = nullptr; size_t pixels_cap = 0;
    uint32_t* d_ids = nullptr; size_t ids_cap = 0;
    unsigned long long* d_counters = nullptr;           // 2 counters followed by the 4 ints of the hit rectangle
    float* h_rgb = nullptr; uint32_t* h_ids = nullptr; size_t stage_cap = 0;      // page-locked staging of the screen rectangle
    size_t tri_count = 0, node_count = 0;
    double bounds[6] = {0, 0, 0, 0, 0, 0};                  // root box (nodes[0].bounds): the conservative screen rectangle comes from it
    cudaStream_t band_stream[4] = {}; cudaEvent_t band_done[4] = {};
};

namespace {
struct RefNode64 { double bounds[6]; uint64_t primitive_count; uint64_t first_child_or_primitive; };      // bvh.hpp:20-28
static_assert(sizeof(RefNode64) == 64, "Bvh<double>::Node");
}

extern "C" {

int ceres_scene_f64_create(const void* nodes_v, size_t node_count, const size_t* prim_idx, const void* tris_v, const void* norms_v,
                           size_t tri_count, int device, ceres_scene_f64** out) {
    if (!out) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: out is null");
    *out = nullptr;
    if (!nodes_v || !prim_idx || !tris_v || !node_count || !tri_count) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: null or empty scene arrays");
    if (tri_count >= 0x7FFFFFFFull || node_count >= 0xFFFFFFFFull) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: scene too large for 32-bit slots");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return fail(CERES_ERR_NO_DEVICE, "no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: bad device index");
    const RefNode64* nodes = static_cast<const RefNode64*>(nodes_v);
    const double* tris = static_cast<const double*>(tris_v);
    const double* norms = static_cast<const double*>(norms_v);
    for (size_t k = 0; k < tri_count; ++k)
        if (prim_idx[k] >= tri_count) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: primitive index out of range");

    // pairs in depth-first order, left child first
    std::vector<f64::Child> pairs;
    uint32_t root_leaf_count = 0, depth_max = 1;
    const RefNode64& root = nodes[0];
    if (root.primitive_count != 0) {
        if (root.first_child_or_primitive != 0 || root.primitive_count != tri_count)
            return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: leaf root must cover all primitives from slot 0");
        root_leaf_count = (uint32_t)root.primitive_count;
        pairs.resize(2);
    } else {
        struct Item { uint64_t first_child; uint32_t pair, depth; };
        std::vector<Item> work;
        pairs.reserve(node_count + 1);
        auto new_pair = [&]() { const uint32_t id = (uint32_t)(pairs.size() / 2); pairs.resize(pairs.size() + 2); return id; };
        size_t visited = 1;
        work.push_back({root.first_child_or_primitive, new_pair(), 1u});
        while (!work.empty()) {
            const Item it = work.back(); work.pop_back();
            if (it.first_child + 1 >= node_count) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: child index out of range");
            visited += 2;
            if (visited > node_count) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: BVH is not a tree (cycle or shared nodes)");
            depth_max = std::max(depth_max, it.depth);
            uint32_t child_pair[2] = {0, 0};
            for (int c = 0; c < 2; ++c) {
                const RefNode64& n = nodes[it.first_child + c];
                f64::Child ch{};
                for (int b = 0; b < 6; ++b) ch.b[b] = n.bounds[b];
                if (n.primitive_count != 0) {
                    if (n.first_child_or_primitive + n.primitive_count > tri_count) return fail(CERES_ERR_INVALID, "ceres_scene_f64_create: leaf range out of bounds");
                    ch.count = (uint32_t)n.primitive_count; ch.first = (uint32_t)n.first_child_or_primitive;
                } else {
                    ch.count = 0; ch.first = child_pair[c] = new_pair();
                }
                pairs[2 * (size_t)it.pair + c] = ch;
            }
            for (int c = 1; c >= 0; --c) {
                const RefNode64& n = nodes[it.first_child + c];
                if (n.primitive_count == 0) work.push_back({n.first_child_or_primitive, child_pair[c], it.depth + 1});
            }
        }
    }
    if (depth_max > 64) return fail(CERES_ERR_UNSUPPORTED, "ceres_scene_f64_create: BVH deeper than the reference's 64-entry stack");
    std::vector<double> tri_slots(tri_count * 12), norm_slots(tri_count * 9, 0.0);
    std::vector<uint32_t> slot_id(tri_count);
    for (size_t s = 0; s < tri_count; ++s) {
        slot_id[s] = (uint32_t)prim_idx[s];
        std::memcpy(&tri_slots[s * 12], tris + prim_idx[s] * 12, 96);
        if (norms) std::memcpy(&norm_slots[s * 9], norms + prim_idx[s] * 9, 72);
    }

    DeviceGuard guard(device);
    if (!guard.ok) return fail(CERES_ERR_CUDA, "cudaSetDevice failed");
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(CERES_ERR_NO_DEVICE, std::string("device is not Blackwell-class (sm_100a required): ") + prop.name);
    ceres_scene_f64* s = new (std::nothrow) ceres_scene_f64;
    if (!s) return fail(CERES_ERR_INVALID, "out of host memory");
    s->device = device; s->tri_count = tri_count; s->node_count = node_count;
    for (int b = 0; b < 6; ++b) s->bounds[b] = root.bounds[b];
#define TRY_D(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail(CERES_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); ceres_scene_f64_destroy(s); return CERES_ERR_CUDA; } } while (0)
    TRY_D(cudaMalloc(&s->d_pairs, pairs.size() * sizeof(f64::Child)));
    TRY_D(cudaMalloc((void**)&s->d_tris, tri_slots.size() * 8));
    TRY_D(cudaMalloc((void**)&s->d_norms, norm_slots.size() * 8));
    TRY_D(cudaMalloc((void**)&s->d_slot_id, slot_id.size() * 4));
    TRY_D(cudaMalloc((void**)&s->d_counters, 5 * sizeof(unsigned long long)));
    TRY_D(cudaMemset(s->d_counters, 0, 5 * sizeof(unsigned long long)));
    TRY_D(cudaMemcpy(s->d_pairs, pairs.data(), pairs.size() * sizeof(f64::Child), cudaMemcpyHostToDevice));
    TRY_D(cudaMemcpy(s->d_tris, tri_slots.data(), tri_slots.size() * 8, cudaMemcpyHostToDevice));
    TRY_D(cudaMemcpy(s->d_norms, norm_slots.data(), norm_slots.size() * 8, cudaMemcpyHostToDevice));
    TRY_D(cudaMemcpy(s->d_slot_id, slot_id.data(), slot_id.size() * 4, cudaMemcpyHostToDevice));
#undef TRY_D
    s->dev.pairs = (const f64::Child*)s->d_pairs; s->dev.tris = s->d_tris; s->dev.norms = s->d_norms; s->dev.slot_id = s->d_slot_id;
    s->dev.root_leaf_count = root_leaf_count;
    s->dev.n_pairs = (uint32_t)(pairs.size() / 2); s->dev.n_slots = (uint32_t)tri_count;
    cudaDeviceSynchronize();                     // uploads (legacy stream) complete before the non-blocking band streams read them
#ifdef CERES_CHECKED
    s->dev.check = s->d_counters + 4;
#endif
    *out = s;
    return CERES_OK;
}

void ceres_scene_f64_destroy(ceres_scene_f64* s) {
    if (!s) return;
    DeviceGuard guard(s->device);
    cudaFree(s->d_pairs); cudaFree(s->d_tris); cudaFree(s->d_norms); cudaFree(s->d_slot_id);
    cudaFree(s->d_rgb); cudaFree(s->d_ids); cudaFree(s->d_counters);
    for (int k = 0; k < 4; ++k) { if (s->band_stream[k]) cudaStreamDestroy(s->band_stream[k]); if (s->band_done[k]) cudaEventDestroy(s->band_done[k]); }
    if (s->h_rgb) cudaFreeHost(s->h_rgb);
    if (s->h_ids) cudaFreeHost(s->h_ids);
    delete s;
}

// One frame, or one row-block shard of it (shard != null, count > 1: only the shard's rows of the full-size frame are
// written -- zeros included -- and the counters are the shard's: rays = own rows * width + own primary hits).
static int render_f64_impl(ceres_scene_f64* s, const ceres_camera_f64* cam, const double sun[3], size_t width, size_t height, unsigned flags,
                           const ceres_shard* shard, double* pixels, uint32_t* ids, uint64_t counters[2], double* kernel_ms) {
    if (!s || !cam || !sun || !pixels) return fail(CERES_ERR_INVALID, "ceres_render_f64: null argument");
    if (!width || !height || (uint64_t)width * height >= (1ull << 31)) return fail(CERES_ERR_INVALID, "ceres_render_f64: bad image size");
    if ((height + 7) / 8 > 65535u) return fail(CERES_ERR_INVALID, "ceres_render_f64: image too tall");
    const bool sharded = shard && shard->count > 1;
    if (sharded && (shard->block_rows == 0 || shard->index >= shard->count)) return fail(CERES_ERR_INVALID, "ceres_render_f64: bad shard");
    // the rows this call owns, as runs [first, second) of frame rows in ascending order (block b -> shard b % count)
    std::vector<std::pair<size_t, size_t>> own;
    size_t own_rows = 0;
    if (!sharded) own.push_back({0, height});
    else for (size_t b = shard->index; b * shard->block_rows < height; b += shard->count)
        own.push_back({b * shard->block_rows, std::min<size_t>(height, (b + 1) * shard->block_rows)});
    for (auto& r : own) own_rows += r.second - r.first;
    DeviceGuard guard(s->device);
    const size_t npix = width * height;
    if (s->pixels_cap < npix) {
        cudaFree(s->d_rgb); s->d_rgb = nullptr; s->pixels_cap = 0;
        CUDA_TRY(cudaMalloc((void**)&s->d_rgb, npix * 12));
        s->pixels_cap = npix;
    }
    if (ids && s->ids_cap < npix) {
        cudaFree(s->d_ids); s->d_ids = nullptr; s->ids_cap = 0;
        CUDA_TRY(cudaMalloc((void**)&s->d_ids, npix * 4));
        s->ids_cap = npix;
    }
    f64::Frame64 f{};
    {   // render.hpp:91-97 in binary64, source order (host code of this file is compiled with -ffp-contract=off)
        auto dot = [](const double* a, const double* b) { double t = a[0] * b[0]; t += a[1] * b[1]; t += a[2] * b[2]; return t; };
        auto normalize = [&](double* a) { const double inv = 1.0 / std::sqrt(dot(a, a)); a[0] = a[0] * inv; a[1] = a[1] * inv; a[2] = a[2] * inv; };
        auto cross = [](const double* a, const double* b, double* c) {
            c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
        };
        double dir[3] = {cam->dir[0], cam->dir[1], cam->dir[2]}, iu[3], iv[3];
        normalize(dir);
        cross(dir, cam->up, iu); normalize(iu);
        cross(iu, dir, iv); normalize(iv);
        const double iw = std::tan(cam->fov * (3.14159265 * (1.0 / 180.0) * 0.5));
        const double ratio = (double)height / (double)width;
        for (int k = 0; k < 3; ++k) {
            f.eye[k] = cam->eye[k]; f.dir[k] = dir[k]; f.iu[k] = iu[k] * iw; f.iv[k] = iv[k] * iw * ratio; f.sun[k] = sun[k];
        }
    }
    f.width = (uint32_t)width; f.height = (uint32_t)height; f.flags = flags;
    f.rgb = s->d_rgb; f.ids = ids ? s->d_ids : nullptr; f.counters = s->d_counters; f.rect = reinterpret_cast<int*>(s->d_counters + 2);
    // Conservative screen rectangle of the scene (the float path's rule, in double: the root box inflated by 1e-4 of the scene
    // scale, corners projected, one pixel of margin; a corner at or behind the eye plane disables it): pixels outside it cannot
    // hit anything, so they are never traced -- their zeros are written by the host pool while the GPU works on the rest.
    size_t x0 = 0, x1 = width, y0 = 0, y1 = height;
    {
        double scale = 0.0;
        for (int k = 0; k < 6; ++k) scale = std::max(scale, std::fabs(s->bounds[k]));
        for (int k = 0; k < 3; ++k) scale = std::max(scale, std::fabs(f.eye[k]));
        auto dot = [](const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
        const double iu2 = dot(f.iu, f.iu), iv2 = dot(f.iv, f.iv), d2 = dot(f.dir, f.dir);
        bool ok = std::isfinite(scale) && scale > 0.0 && iu2 > 0.0 && iv2 > 0.0 && d2 > 0.0 && std::isfinite(iu2 + iv2 + d2) && !s->dev.root_leaf_count;
        double umin = INFINITY, umax = -INFINITY, vmin = INFINITY, vmax = -INFINITY;
        const double m = 1e-4 * scale;
        for (int c = 0; c < 8 && ok; ++c) {
            double p[3];
            for (int k = 0; k < 3; ++k) p[k] = ((c >> k) & 1 ? s->bounds[2 * k + 1] + m : s->bounds[2 * k] - m) - f.eye[k];
            const double z = dot(p, f.dir) / d2;
            if (!(z > 1e-6 * scale)) { ok = false; break; }
            const double u = dot(p, f.iu) / iu2 / z, v = dot(p, f.iv) / iv2 / z;
            umin = std::min(umin, u); umax = std::max(umax, u); vmin = std::min(vmin, v); vmax = std::max(vmax, v);
        }
        if (ok && std::isfinite(umin + umax + vmin + vmax)) {
            auto lo_px = [](double u, double n) { double x = std::floor((u + 1.0) * n / 2.0 - 0.5) - 1.0; return x < 0.0 ? 0.0 : (x > n ? n : x); };
            auto hi_px = [](double u, double n) { double x = std::ceil((u + 1.0) * n / 2.0 - 0.5) + 2.0; return x < 0.0 ? 0.0 : (x > n ? n : x); };
            x0 = (size_t)lo_px(umin, (double)width); x1 = std::max(x0, (size_t)hi_px(umax, (double)width));
            y0 = (size_t)lo_px(vmin, (double)height); y1 = std::max(y0, (size_t)hi_px(vmax, (double)height));
        }
    }
    // the rectangle's rows in LOCAL rows of this shard (frame_row is increasing in the local row)
    auto frame_row = [&](size_t lr) -> size_t {
        if (!sharded) return lr;
        const size_t blk = lr / shard->block_rows;
        return (blk * shard->count + shard->index) * shard->block_rows + (lr - blk * shard->block_rows);
    };
    size_t l0 = 0;
    while (l0 < own_rows && frame_row(l0) < y0) ++l0;
    size_t l1 = l0;
    while (l1 < own_rows && frame_row(l1) < y1) ++l1;
    const size_t rw = x1 - x0, rh = l1 - l0;
    const bool any = rw > 0 && rh > 0;
    f.blk_rows = sharded ? shard->block_rows : 0u; f.sh_index = sharded ? shard->index : 0u; f.sh_count = sharded ? shard->count : 1u;
    // zeros (ids: CERES_MISS) outside the rectangle: host pool, from now on
    ceres_host::Job zeros;
    {
        auto outside = [&](unsigned char* base, size_t px_bytes, unsigned char byte) {
            for (auto& r : own) {
                if (!any) { zeros.add(base + width * r.first * px_bytes, width * (r.second - r.first) * px_bytes, byte); continue; }
                const size_t a = std::min(r.second, std::max(r.first, y0)), b = std::max(a, std::min(r.second, y1));   // rows [a, b) cross the rectangle
                if (a > r.first) zeros.add(base + width * r.first * px_bytes, width * (a - r.first) * px_bytes, byte);
                if (r.second > b) zeros.add(base + width * b * px_bytes, width * (r.second - b) * px_bytes, byte);
                for (size_t j = a; j < b; ++j) {
                    unsigned char* row = base + width * j * px_bytes;
                    if (x0) zeros.segs.push_back({row, x0 * px_bytes, byte});
                    if (x1 < width) zeros.segs.push_back({row + x1 * px_bytes, (width - x1) * px_bytes, byte});
                }
            }
        };
        outside(reinterpret_cast<unsigned char*>(pixels), 24, 0x00);
        if (ids) outside(reinterpret_cast<unsigned char*>(ids), 4, 0xFF);
    }
    ceres_host::Pool& pool = ceres_host::Pool::instance();
    pool.submit(&zeros);
    struct Waiter { ceres_host::Pool& p; ceres_host::Job* j; ~Waiter() { p.wait(j); } } waiter{pool, &zeros};
    struct Out { unsigned long long c[2]; int rect[4]; } o = {{0, 0}, {0x7FFFFFFF, 0x7FFFFFFF, -1, -1}};
    cudaError_t err = cudaSuccess;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    std::vector<std::unique_ptr<ceres_host::Job>> bodies;
    struct BodyWaiter { ceres_host::Pool& p; std::vector<std::unique_ptr<ceres_host::Job>>& v; ~BodyWaiter() { for (auto& j : v) p.wait(j.get()); } } body_waiter{pool, bodies};
    if (any) {
        if (rw * rh > s->stage_cap) {
            if (s->h_rgb) cudaFreeHost(s->h_rgb);
            if (s->h_ids) cudaFreeHost(s->h_ids);
            s->h_rgb = nullptr; s->h_ids = nullptr; s->stage_cap = 0;
            CUDA_TRY(cudaMallocHost((void**)&s->h_rgb, rw * rh * 12));
            CUDA_TRY(cudaMallocHost((void**)&s->h_ids, rw * rh * 4));
            s->stage_cap = rw * rh;
        }
        for (int k = 0; k < 4; ++k) {
            if (!s->band_stream[k]) CUDA_TRY(cudaStreamCreateWithFlags(&s->band_stream[k], cudaStreamNonBlocking));
            if (!s->band_done[k]) CUDA_TRY(cudaEventCreateWithFlags(&s->band_done[k], cudaEventDisableTiming));
        }
        CUDA_TRY(cudaMemcpy(s->d_counters, &o, sizeof o, cudaMemcpyHostToDevice));
        CUDA_TRY(cudaStreamSynchronize(0));      // a small pageable upload may still be in flight; the band streams are non-blocking
        if (kernel_ms) { CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1)); CUDA_TRY(cudaEventRecord(e0, 0)); CUDA_TRY(cudaEventSynchronize(e0)); }
        // Row bands of the rectangle, each on its own stream: kernel -> DMA of the band's float pixels into page-locked staging;
        // the host widens a band into the caller's double frame as soon as its copy has landed, while later bands render.
        // Row bands of the rectangle IN LOCAL ROWS (a shard's rows are packed on the device), each on its own stream:
        // kernel -> DMA of the band's float pixels into page-locked staging; the host widens a band into the caller's double
        // frame -- at its frame rows -- as soon as its copy has landed, while later bands render.
        const size_t nb = std::max<size_t>(1, std::min<size_t>(4, (rw * rh) >> 18));
        const size_t band_rows = ((rh + nb - 1) / nb + 7) / 8 * 8;
        size_t launched = 0;
        for (size_t k = 0, r0 = l0; r0 < l1; ++k, r0 += band_rows) {
            const size_t r1 = std::min(l1, r0 + band_rows);
            f.x0 = (uint32_t)x0; f.x1 = (uint32_t)x1; f.y0 = (uint32_t)r0; f.y1 = (uint32_t)r1;
            cudaStream_t st = s->band_stream[k];
            f64::render_kernel<<<dim3((unsigned)((rw + 15) / 16), (unsigned)((r1 - r0 + 7) / 8)), 128, 0, st>>>(s->dev, f);
            g_launches++;
            err = cudaGetLastError();
            if (err == cudaSuccess) err = cudaMemcpy2DAsync(s->h_rgb + 3 * rw * (r0 - l0), rw * 12, s->d_rgb + 3 * (width * r0 + x0), width * 12, rw * 12, r1 - r0, cudaMemcpyDeviceToHost, st);
            if (err == cudaSuccess && ids) err = cudaMemcpy2DAsync(s->h_ids + rw * (r0 - l0), rw * 4, s->d_ids + (width * r0 + x0), width * 4, rw * 4, r1 - r0, cudaMemcpyDeviceToHost, st);
            if (err == cudaSuccess) err = cudaEventRecord(s->band_done[k], st);
            if (err != cudaSuccess) break;
            launched = k + 1;
        }
        for (size_t k = 0, r0 = l0; k < launched; ++k, r0 += band_rows) {
            const size_t r1 = std::min(l1, r0 + band_rows);
            const cudaError_t e = cudaEventSynchronize(s->band_done[k]);
            if (e != cudaSuccess) { if (err == cudaSuccess) err = e; continue; }
            auto body = std::make_unique<ceres_host::Job>();
            for (size_t lr = r0; lr < r1; ++lr) {
                const size_t j = frame_row(lr);
                body->add_widen(pixels + 3 * (width * j + x0), s->h_rgb + 3 * rw * (lr - l0), 3 * rw);
                if (ids) body->add_copy(ids + width * j + x0, s->h_ids + rw * (lr - l0), rw * 4);
            }
            pool.submit(body.get());
            bodies.push_back(std::move(body));
        }
        if (kernel_ms && err == cudaSuccess) {
            // device time of the frame: first launch to last band done (the bands' streams have all been synchronised above)
            cudaEventRecord(e1, 0); cudaEventSynchronize(e1);
            float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1); *kernel_ms = ms;
        }
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        if (err == cudaSuccess) err = cudaMemcpy(&o, s->d_counters, sizeof o, cudaMemcpyDeviceToHost);
#ifdef CERES_CHECKED
        if (err == cudaSuccess) {
            unsigned long long w = 0;
            err = cudaMemcpy(&w, s->d_counters + 4, sizeof w, cudaMemcpyDeviceToHost);
            if (err == cudaSuccess && w) {
                cudaMemset(s->d_counters + 4, 0, sizeof w);
                return fail(CERES_ERR_CUDA, "ceres_render_f64: checked build: an index left its bound (bits " + std::to_string(w) + ")");
            }
        }
#endif
    }
    pool.wait(&zeros);
    for (auto& j : bodies) pool.wait(j.get());
    if (err != cudaSuccess) return fail(CERES_ERR_CUDA, std::string("ceres_render_f64: ") + cudaGetErrorString(err));
    const unsigned long long c[2] = {o.c[0], o.c[1]};
    if (counters) { counters[0] = (uint64_t)own_rows * width + c[0]; counters[1] = c[1]; }      // render.hpp:115,138 / :119,148
    return CERES_OK;
}

int ceres_render_f64(ceres_scene_f64* s, const ceres_camera_f64* cam, const double sun[3], size_t width, size_t height, unsigned flags,
                     double* pixels, uint32_t* ids, uint64_t counters[2], double* kernel_ms) {
    return render_f64_impl(s, cam, sun, width, height, flags, nullptr, pixels, ids, counters, kernel_ms);
}

int ceres_render_f64_shard(ceres_scene_f64* s, const ceres_camera_f64* cam, const double sun[3], size_t width, size_t height, unsigned flags,
                           const ceres_shard* shard, double* pixels, uint32_t* ids, uint64_t counters[2], double* kernel_ms) {
    return render_f64_impl(s, cam, sun, width, height, flags, shard, pixels, ids, counters, kernel_ms);
}

}  // extern "C"
