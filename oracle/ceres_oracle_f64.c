/* oracle/ceres_oracle_f64.c -- CPU RESTATEMENT of the render loop for Scalar = double.  TEST INFRASTRUCTURE ONLY.
 *
 * The double-precision instantiation of the reference (`anim -d`, include/anim.cpp:154-155) restated in plain C
 * (gnu11, -ffp-contract=off), next to oracle/ceres_oracle.c (Scalar = float).  Every function cites the reference
 * file:line it follows.  Parity status: PINNED -- tests/test_oracle_f64.py checks it bit for bit (hit ids, RGB,
 * counters) against the reference's own headers compiled unchanged with Scalar = double
 * (oracle/_ref/libceres_ref64_strict.so, oracle/ref_driver_f64.cpp) and against tests/golden/f64_bunny.npz made from
 * that library (tests/golden/make_golden_f64.py).
 *
 * What changes with Scalar = double (and what does not):
 *   nodes      64 bytes: double bounds[6], uint64 primitive_count, uint64 first_child_or_primitive (bvh.hpp:20-28)
 *   triangles  96 bytes {p0, e1, e2, n}; tri_norms 72 bytes
 *   slab test  std::fma (utilities.hpp:46-49, FP_FAST_FMA is defined under -mfma); safe_inverse threshold DBL_EPSILON
 *   ray        tmax = DBL_MAX (ray.hpp:17-20)
 *   shading    STILL float where render.hpp says float: lambertian / blinn_phong_spec return float, amb / diffuse /
 *              specular are float, the colour accumulators are float and take `double weight * float colour` rounded
 *              to float at every `+=` (render.hpp:46-84); the pixel is that float widened (render.hpp:144-146).
 *
 * Only tests/ and __graft_entry__.smoke() may load this file's library; the product path never does.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define CO_MISS 0xFFFFFFFFu
#define CO_STACK 64 /* lib/bvh/single_ray_traverser.hpp:14 */

typedef struct { double v[3]; } vec3;
typedef struct { double bounds[6]; uint64_t primitive_count; uint64_t first_child_or_primitive; } node64;   /* bvh.hpp:25-30 */
typedef struct { vec3 p0, e1, e2, n; } tri64;                                                             /* triangle.hpp:27 */
typedef struct { vec3 o, d; double tmin, tmax; } ray64;                                                   /* ray.hpp:10-15 */
typedef struct { const node64* nodes; const uint64_t* prim_idx; const tri64* tris; const double* norms; } scene64;

/* ---- lib/bvh/vector.hpp:92-167 ------------------------------------------------------------------------------ */
static inline vec3 v3(double x, double y, double z) { vec3 r = {{x, y, z}}; return r; }
static inline vec3 vadd(vec3 a, vec3 b) { return v3(a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2]); }
static inline vec3 vsub(vec3 a, vec3 b) { return v3(a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2]); }
static inline vec3 vscale(vec3 a, double s) { return v3(a.v[0] * s, a.v[1] * s, a.v[2] * s); }
static inline double vdot(vec3 a, vec3 b) { /* :134-141 */
    double sum = a.v[0] * b.v[0];
    sum += a.v[1] * b.v[1];
    sum += a.v[2] * b.v[2];
    return sum;
}
static inline vec3 vcross(vec3 a, vec3 b) { /* :159-167 */
    return v3(a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2], a.v[0] * b.v[1] - a.v[1] * b.v[0]);
}
static inline vec3 vnormalize(vec3 a) { /* :143-154 */
    double inv = 1.0 / sqrt(vdot(a, a));
    return vscale(a, inv);
}
static inline vec3 tri_p1(const tri64* t) { return vsub(t->p0, t->e1); }   /* triangle.hpp:36 */
static inline vec3 tri_p2(const tri64* t) { return vadd(t->p0, t->e2); }   /* triangle.hpp:37 */

/* ---- lib/bvh/triangle.hpp:95-115 ----------------------------------------------------------------------------- */
static inline int tri_intersect(const tri64* tr, const ray64* ray, double* t_out, double* u_out, double* v_out) {
    vec3 c = vsub(tr->p0, ray->o);
    vec3 r = vcross(ray->d, c);
    double inv_det = 1.0 / vdot(tr->n, ray->d);
    double u = vdot(r, tr->e2) * inv_det;
    double v = vdot(r, tr->e1) * inv_det;
    double w = 1.0 - u - v;
    if (u >= 0 && v >= 0 && w >= 0) {
        double t = vdot(tr->n, c) * inv_det;
        if (t >= ray->tmin && t <= ray->tmax) { *t_out = t; *u_out = u; *v_out = v; return 1; }
    }
    return 0;
}

/* ---- lib/bvh/node_intersectors.hpp:20-26,34-47,89-100; vector.hpp:69-74 ------------------------------------- */
typedef struct { int oct[3]; double inv[3]; double so[3]; } nodeint64;
static inline void nodeint_init(nodeint64* ni, const ray64* ray) {
    const double threshold = DBL_EPSILON;
    for (int k = 0; k < 3; ++k) {
        double d = ray->d.v[k];
        ni->oct[k] = signbit(d) ? 1 : 0;
        ni->inv[k] = 1.0 / (fabs(d) < threshold ? copysign(threshold, d) : d);
        ni->so[k] = -ray->o.v[k] * ni->inv[k];
    }
}
static inline double robust_min(double x, double y) { return x < y ? x : y; }   /* utilities.hpp:57-67 */
static inline double robust_max(double x, double y) { return x > y ? x : y; }
static inline void slab(const nodeint64* ni, const double* b, double tmin, double tmax, double* entry, double* exit_) {
    double e0 = fma(b[0 + ni->oct[0]], ni->inv[0], ni->so[0]);
    double e1 = fma(b[2 + ni->oct[1]], ni->inv[1], ni->so[1]);
    double e2 = fma(b[4 + ni->oct[2]], ni->inv[2], ni->so[2]);
    double x0 = fma(b[0 + 1 - ni->oct[0]], ni->inv[0], ni->so[0]);
    double x1 = fma(b[2 + 1 - ni->oct[1]], ni->inv[1], ni->so[1]);
    double x2 = fma(b[4 + 1 - ni->oct[2]], ni->inv[2], ni->so[2]);
    *entry = robust_max(e0, robust_max(e1, robust_max(e2, tmin)));
    *exit_ = robust_min(x0, robust_min(x1, robust_min(x2, tmax)));
}

typedef struct { uint32_t id; double t, u, v; int hit; } hit64;

/* single_ray_traverser.hpp:41-63, primitive_intersectors.hpp:17-20,48-53 */
static inline void intersect_leaf(const scene64* s, const node64* node, ray64* ray, hit64* best) {
    size_t begin = (size_t)node->first_child_or_primitive, end = begin + (size_t)node->primitive_count;
    for (size_t i = begin; i < end; ++i) {
        size_t id = (size_t)s->prim_idx[i];
        double t, u, v;
        if (tri_intersect(&s->tris[id], ray, &t, &u, &v)) {
            best->hit = 1; best->id = (uint32_t)id; best->t = t; best->u = u; best->v = v;
            ray->tmax = t;
        }
    }
}

/* single_ray_traverser.hpp:65-126 */
static hit64 traverse(const scene64* s, ray64 ray) {
    hit64 best; best.hit = 0; best.id = CO_MISS; best.t = best.u = best.v = 0;
    const node64* nodes = s->nodes;
    if (nodes[0].primitive_count != 0) { intersect_leaf(s, &nodes[0], &ray, &best); return best; }
    nodeint64 ni; nodeint_init(&ni, &ray);
    uint64_t stack[CO_STACK]; size_t sp = 0;
    const node64* left = &nodes[nodes[0].first_child_or_primitive];
    for (;;) {
        const node64* right = left + 1;
        double le, lx, re, rx;
        slab(&ni, left->bounds, ray.tmin, ray.tmax, &le, &lx);
        slab(&ni, right->bounds, ray.tmin, ray.tmax, &re, &rx);
        if (le <= lx) {
            if (left->primitive_count != 0) { intersect_leaf(s, left, &ray, &best); left = NULL; }
        } else left = NULL;
        if (re <= rx) {
            if (right->primitive_count != 0) { intersect_leaf(s, right, &ray, &best); right = NULL; }
        } else right = NULL;
        if (left) {
            if (right) {
                if (le > re) { const node64* tmp = left; left = right; right = tmp; }
                stack[sp++] = right->first_child_or_primitive;
            }
            left = &nodes[left->first_child_or_primitive];
        } else if (right) {
            left = &nodes[right->first_child_or_primitive];
        } else {
            if (sp == 0) break;
            left = &nodes[stack[--sp]];
        }
    }
    return best;
}

/* ---- include/render.hpp:46-84: float results and accumulators also when Scalar = double --------------------- */
static inline float lambertian(vec3 s, vec3 n) { return (float)fabs(s.v[0] * n.v[0] + s.v[1] * n.v[1] + s.v[2] * n.v[2]); }
static inline float blinn_phong_spec(vec3 sun_line, vec3 n, vec3 view) {
    return (float)pow(vdot(n, vnormalize(vadd(sun_line, view))), 24.0);
}
static inline float clamp01(float x) { return x < 0.f ? 0.f : (1.f < x ? 1.f : x); }
static void smooth_shading(vec3 sun_line, const double* norms, vec3 view, double u, double v, float* c) {
    const float amb = 0.2f;
    double w[3]; w[0] = u; w[1] = v; w[2] = 1 - u - v;
    vec3 mview = vscale(view, -1.0);
    c[0] = c[1] = c[2] = 0.f;
    for (int k = 0; k < 3; ++k) {
        vec3 n = v3(norms[3 * k], norms[3 * k + 1], norms[3 * k + 2]);
        float diffuse = 0.5f * lambertian(sun_line, n);
        float specular = 0.8f * blinn_phong_spec(sun_line, n, mview);
        c[0] = (float)((double)c[0] + w[k] * (double)clamp01((amb + diffuse) * 0.5f + specular));
        c[1] = (float)((double)c[1] + w[k] * (double)clamp01((amb + diffuse) * 0.0f + specular));
        c[2] = (float)((double)c[2] + w[k] * (double)clamp01((amb + diffuse) * 0.8f + specular));
    }
}

/* ---- include/render.hpp:91-97.  cam = eye[3], dir[3], up[3], fov; out = dir[3], image_u[3], image_v[3] -------- */
void co64_frame_setup(const double* cam, size_t width, size_t height, double* out) {
    vec3 dir = vnormalize(v3(cam[3], cam[4], cam[5]));
    vec3 up = v3(cam[6], cam[7], cam[8]);
    vec3 iu = vnormalize(vcross(dir, up));
    vec3 iv = vnormalize(vcross(iu, dir));
    double iw = tan(cam[9] * (3.14159265 * (1.0 / 180.0) * 0.5));
    double ratio = (double)height / (double)width;
    iu = vscale(iu, iw);
    iv = vscale(vscale(iv, iw), ratio);
    for (int k = 0; k < 3; ++k) { out[k] = dir.v[k]; out[3 + k] = iu.v[k]; out[6 + k] = iv.v[k]; }
}

/* ---- include/render.hpp:99-155.  flags bit0: no shadow ray.  Outputs nullable; counters[2] = {rays, hits}. ---- */
void co64_render(const void* nodes, const uint64_t* prim_idx, const void* tris, const double* norms,
                 const double* cam, const double* sun, size_t width, size_t height, unsigned flags,
                 double* pixels, uint32_t* ids, uint8_t* state, uint64_t* counters) {
    scene64 s = {(const node64*)nodes, prim_idx, (const tri64*)tris, norms};
    double fs[9];
    co64_frame_setup(cam, width, height, fs);
    vec3 dir = v3(fs[0], fs[1], fs[2]), image_u = v3(fs[3], fs[4], fs[5]), image_v = v3(fs[6], fs[7], fs[8]);
    vec3 eye = v3(cam[0], cam[1], cam[2]);
    vec3 sun_position = v3(sun[0], sun[1], sun[2]);
    const int shadows = !(flags & 1u);
    uint64_t n_rays = 0, n_hits = 0;
    #pragma omp parallel for schedule(dynamic, 4) reduction(+: n_rays, n_hits)
    for (size_t j = 0; j < height; ++j) {
        for (size_t i = 0; i < width; ++i) {
            size_t pix = width * j + i, index = 3 * pix;                         /* :107 */
            double u = 2 * ((double)i + 0.5) / (double)width - 1.0;              /* :109 */
            double v = 2 * ((double)j + 0.5) / (double)height - 1.0;             /* :110 */
            vec3 view = vnormalize(vadd(vadd(vscale(image_u, u), vscale(image_v, v)), dir));
            ray64 ray = {eye, view, 0.0, DBL_MAX};                               /* :113 */
            hit64 hit = traverse(&s, ray);
            n_rays++;
            float rgb[3] = {0, 0, 0};
            uint8_t stt = 0;
            if (hit.hit) {
                n_hits++;                                                        /* :119 */
                const tri64* tr = &s.tris[hit.id];
                vec3 normal = vnormalize(tr->n);
                double hu = hit.u, hv = hit.v;
                vec3 P = vadd(vadd(vscale(tr->p0, hu), vscale(tri_p1(tr), hv)), vscale(tri_p2(tr), 1 - hu - hv));
                double scale = -0.00001;                                         /* :132 */
                P = vadd(P, vscale(normal, scale));
                vec3 sun_line = vnormalize(vsub(sun_position, P));
                int occluded = 0;
                if (shadows) {
                    ray64 sray = {P, sun_line, 0.0, DBL_MAX};
                    occluded = traverse(&s, sray).hit;                           /* :137 */
                    n_rays++;
                }
                if (!occluded) { smooth_shading(sun_line, s.norms + 9 * (size_t)hit.id, view, hu, hv, rgb); stt = 1; }
                else { n_hits++; stt = 2; }
            }
            if (pixels) { pixels[index] = rgb[0]; pixels[index + 1] = rgb[1]; pixels[index + 2] = rgb[2]; }
            if (ids) ids[pix] = hit.hit ? hit.id : CO_MISS;
            if (state) state[pix] = stt;
        }
    }
    if (counters) { counters[0] = n_rays; counters[1] = n_hits; }
}
