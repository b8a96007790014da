/* oracle/ceres_oracle.c -- CPU RESTATEMENT of the ceres-raytracer render loop.  TEST INFRASTRUCTURE ONLY.
 *
 * Plain C (gnu11, -ffp-contract=off) restatement of the reference algorithm, written from the behaviour of the
 * reference headers; every function cites the reference file:line it follows (paths relative to
 * /root/reference).  Parity status: PINNED -- tests/test_oracle.py checks this restatement bit-for-bit (hit ids,
 * t/u/v, RGB, counters, S/T statistics) against the reference's own code compiled unchanged
 * (oracle/_ref/libceres_ref_strict.so, see oracle/ref_driver.cpp) and against the committed fixtures in
 * tests/golden/ that were generated from that library (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; the product path (ceres_raytracer_b200/) never does and fails loudly without its CUDA library.
 *
 * Arithmetic: IEEE-754 binary32 in source order, the only fused operation being the explicit fmaf of the slab
 * test (lib/bvh/utilities.hpp:37-40, active because -mfma defines FP_FAST_FMAF), which is the arithmetic of the
 * reference built with -ffp-contract=off ("strict oracle").
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <float.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define CO_MISS 0xFFFFFFFFu
#define CO_STACK 64 /* lib/bvh/single_ray_traverser.hpp:14 */

typedef struct { float v[3]; } vec3;

/* lib/bvh/bvh.hpp:25-30 */
typedef struct { float bounds[6]; uint32_t primitive_count; uint32_t first_child_or_primitive; } co_node;
/* lib/bvh/triangle.hpp:27 */
typedef struct { vec3 p0, e1, e2, n; } co_tri;
/* lib/bvh/ray.hpp:10-15 */
typedef struct { vec3 o, d; float tmin, tmax; } co_ray;

typedef struct {
    const co_node* nodes;
    const uint64_t* prim_idx; /* size_t in the reference, lib/bvh/bvh.hpp:94 */
    const co_tri* tris;
    const float* norms;       /* 9 floats per triangle, lib/obj_norms.hpp:113-115 */
    size_t node_count, tri_count;
} co_scene;

/* ---- lib/bvh/vector.hpp ---------------------------------------------------------------------------------- */
static inline vec3 v3(float x, float y, float z) { vec3 r = {{x, y, z}}; return r; }
static inline vec3 vadd(vec3 a, vec3 b) { return v3(a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2]); }
static inline vec3 vsub(vec3 a, vec3 b) { return v3(a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2]); }
static inline vec3 vscale(vec3 a, float s) { return v3(a.v[0] * s, a.v[1] * s, a.v[2] * s); }
/* vector.hpp:134-141: sum = a0*b0; sum += a1*b1; sum += a2*b2 */
static inline float vdot(vec3 a, vec3 b) {
    float sum = a.v[0] * b.v[0];
    sum += a.v[1] * b.v[1];
    sum += a.v[2] * b.v[2];
    return sum;
}
/* vector.hpp:159-167: c[i] = a[j]*b[k] - a[k]*b[j], j=(i+1)%3, k=(i+2)%3 */
static inline vec3 vcross(vec3 a, vec3 b) {
    return v3(a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2], a.v[0] * b.v[1] - a.v[1] * b.v[0]);
}
/* vector.hpp:143-154: inv = 1/sqrt(dot(v,v)); v*inv */
static inline vec3 vnormalize(vec3 a) {
    float inv = 1.0f / sqrtf(vdot(a, a));
    return vscale(a, inv);
}

/* ---- lib/bvh/triangle.hpp:30-37 --------------------------------------------------------------------------- */
static inline co_tri tri_make(vec3 p0, vec3 p1, vec3 p2) {
    co_tri t;
    t.p0 = p0; t.e1 = vsub(p0, p1); t.e2 = vsub(p2, p0);
    t.n = vcross(t.e1, t.e2); /* LeftHandedNormal = true */
    return t;
}
static inline vec3 tri_p1(const co_tri* t) { return vsub(t->p0, t->e1); }
static inline vec3 tri_p2(const co_tri* t) { return vadd(t->p0, t->e2); }

/* ---- lib/bvh/triangle.hpp:95-115 Moeller-Trumbore with precomputed n --------------------------------------- */
static inline int tri_intersect(const co_tri* tr, const co_ray* ray, float* t_out, float* u_out, float* v_out) {
    vec3 c = vsub(tr->p0, ray->o);
    vec3 r = vcross(ray->d, c);
    float inv_det = 1.0f / vdot(tr->n, ray->d);
    float u = vdot(r, tr->e2) * inv_det;
    float v = vdot(r, tr->e1) * inv_det;
    float w = 1.0f - u - v;
    if (u >= 0 && v >= 0 && w >= 0) { /* false on NaN, triangle.hpp:104-106 */
        float t = vdot(tr->n, c) * inv_det;
        if (t >= ray->tmin && t <= ray->tmax) { *t_out = t; *u_out = u; *v_out = v; return 1; }
    }
    return 0;
}

/* ---- lib/bvh/node_intersectors.hpp:20-26,89-94 + vector.hpp:69-74 ------------------------------------------ */
typedef struct { int oct[3]; float inv[3]; float so[3]; } co_nodeint;
static inline void nodeint_init(co_nodeint* ni, const co_ray* ray) {
    const float threshold = FLT_EPSILON;
    for (int k = 0; k < 3; ++k) {
        float d = ray->d.v[k];
        ni->oct[k] = signbit(d) ? 1 : 0;
        ni->inv[k] = 1.0f / (fabsf(d) < threshold ? copysignf(threshold, d) : d);
        ni->so[k] = -ray->o.v[k] * ni->inv[k];
    }
}
/* utilities.hpp:57-67: accumulator on the right */
static inline float robust_min(float x, float y) { return x < y ? x : y; }
static inline float robust_max(float x, float y) { return x > y ? x : y; }
/* node_intersectors.hpp:34-47,96-100 */
static inline void slab(const co_nodeint* ni, const float* bounds, float tmin, float tmax, float* entry, float* exit_) {
    float e0 = fmaf(bounds[0 + ni->oct[0]], ni->inv[0], ni->so[0]);
    float e1 = fmaf(bounds[2 + ni->oct[1]], ni->inv[1], ni->so[1]);
    float e2 = fmaf(bounds[4 + ni->oct[2]], ni->inv[2], ni->so[2]);
    float x0 = fmaf(bounds[0 + 1 - ni->oct[0]], ni->inv[0], ni->so[0]);
    float x1 = fmaf(bounds[2 + 1 - ni->oct[1]], ni->inv[1], ni->so[1]);
    float x2 = fmaf(bounds[4 + 1 - ni->oct[2]], ni->inv[2], ni->so[2]);
    *entry = robust_max(e0, robust_max(e1, robust_max(e2, tmin)));
    *exit_ = robust_min(x0, robust_min(x1, robust_min(x2, tmax)));
}

typedef struct { uint32_t id; float t, u, v; int hit; } co_hit;
typedef struct { uint64_t steps, tests; } co_stats;

/* single_ray_traverser.hpp:41-63 with primitive_intersectors.hpp:17-20,48-53 (Permuted=false, closest) */
static inline void intersect_leaf(const co_scene* s, const co_node* node, co_ray* ray, co_hit* best, co_stats* st) {
    size_t begin = node->first_child_or_primitive;
    size_t end = begin + node->primitive_count;
    st->tests += end - begin;
    for (size_t i = begin; i < end; ++i) {
        size_t id = (size_t)s->prim_idx[i];
        float t, u, v;
        if (tri_intersect(&s->tris[id], ray, &t, &u, &v)) {
            best->hit = 1; best->id = (uint32_t)id; best->t = t; best->u = u; best->v = v;
            ray->tmax = t;
        }
    }
}

/* single_ray_traverser.hpp:65-126: ordered, eager-leaf, stack-based closest-hit traversal */
static co_hit traverse(const co_scene* s, co_ray ray, co_stats* st) {
    co_hit best; best.hit = 0; best.id = CO_MISS; best.t = best.u = best.v = 0;
    const co_node* nodes = s->nodes;
    if (nodes[0].primitive_count != 0) { intersect_leaf(s, &nodes[0], &ray, &best, st); return best; }
    co_nodeint ni; nodeint_init(&ni, &ray);
    uint32_t stack[CO_STACK]; size_t sp = 0;
    const co_node* left = &nodes[nodes[0].first_child_or_primitive];
    for (;;) {
        st->steps++;
        const co_node* right = left + 1;
        float le, lx, re, rx;
        slab(&ni, left->bounds, ray.tmin, ray.tmax, &le, &lx);
        slab(&ni, right->bounds, ray.tmin, ray.tmax, &re, &rx); /* before left's leaf shrinks tmax, :86-87 */
        if (le <= lx) {
            if (left->primitive_count != 0) { intersect_leaf(s, left, &ray, &best, st); left = NULL; }
        } else left = NULL;
        if (re <= rx) {
            if (right->primitive_count != 0) { intersect_leaf(s, right, &ray, &best, st); right = NULL; }
        } else right = NULL;
        if (left) {
            if (right) {
                if (le > re) { const co_node* tmp = left; left = right; right = tmp; } /* ties keep left, :111-112 */
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

/* ---- include/render.hpp:46-84 shading ---------------------------------------------------------------------- */
static inline float lambertian(vec3 sun_line, vec3 n) { /* :46-49 */
    return fabsf(sun_line.v[0] * n.v[0] + sun_line.v[1] * n.v[1] + sun_line.v[2] * n.v[2]);
}
static inline float blinn_phong_spec(vec3 sun_line, vec3 n, vec3 view) { /* :51-54: std::pow(float,int) is a double pow */
    return (float)pow((double)vdot(n, vnormalize(vadd(sun_line, view))), 24.0);
}
static inline float clamp01(float x) { return x < 0.f ? 0.f : (1.f < x ? 1.f : x); } /* std::clamp */
static void smooth_shading(vec3 sun_line, const float* norms, vec3 view, float u, float v, float* c) { /* :56-84 */
    const float amb = 0.2f;
    float w[3]; w[0] = u; w[1] = v; w[2] = 1 - u - v;
    vec3 mview = vscale(view, -1.0f);
    c[0] = c[1] = c[2] = 0.f;
    for (int k = 0; k < 3; ++k) {
        vec3 n = v3(norms[3 * k], norms[3 * k + 1], norms[3 * k + 2]);
        float diffuse = 0.5f * lambertian(sun_line, n);
        float specular = 0.8f * blinn_phong_spec(sun_line, n, mview);
        c[0] += w[k] * clamp01((amb + diffuse) * 0.5f + specular);
        c[1] += w[k] * clamp01((amb + diffuse) * 0.0f + specular);
        c[2] += w[k] * clamp01((amb + diffuse) * 0.8f + specular);
    }
}

/* ---- include/render.hpp:91-97 frame setup ------------------------------------------------------------------ */
/* cam = eye[3], dir[3], up[3], fov.  out = dir[3], image_u[3], image_v[3] */
void co_frame_setup(const float* cam, size_t width, size_t height, float* out) {
    vec3 dir = vnormalize(v3(cam[3], cam[4], cam[5]));
    vec3 up = v3(cam[6], cam[7], cam[8]);
    vec3 iu = vnormalize(vcross(dir, up));
    vec3 iv = vnormalize(vcross(iu, dir));
    float iw = tanf(cam[9] * (float)(3.14159265 * (1.0 / 180.0) * 0.5));
    float ratio = (float)height / (float)width;
    iu = vscale(iu, iw);
    iv = vscale(vscale(iv, iw), ratio);
    for (int k = 0; k < 3; ++k) { out[k] = dir.v[k]; out[3 + k] = iu.v[k]; out[6 + k] = iv.v[k]; }
}

/* ---- include/render.hpp:99-155 the per-pixel loop ----------------------------------------------------------- */
/* flags bit0: no shadow ray (shade every primary hit).  All outputs nullable.
 * stats u64[8] = {S_miss,T_miss,S_hit,T_hit,S_shadow,T_shadow,rays,hits}; rays/hits are render.hpp's counters. */
void co_render(const void* nodes, size_t node_count, const uint64_t* prim_idx, const void* tris, const float* norms,
               size_t tri_count, const float* cam, const float* sun, size_t width, size_t height, unsigned flags,
               float* pixels, uint32_t* ids, uint8_t* state, float* tuv, uint64_t* stats) {
    co_scene s = {(const co_node*)nodes, prim_idx, (const co_tri*)tris, norms, node_count, tri_count};
    float fs[9];
    co_frame_setup(cam, width, height, fs);
    vec3 dir = v3(fs[0], fs[1], fs[2]), image_u = v3(fs[3], fs[4], fs[5]), image_v = v3(fs[6], fs[7], fs[8]);
    vec3 eye = v3(cam[0], cam[1], cam[2]);
    vec3 sun_position = v3(sun[0], sun[1], sun[2]);
    const int shadows = !(flags & 1u);
    uint64_t s_miss = 0, t_miss = 0, s_hit = 0, t_hit = 0, s_sh = 0, t_sh = 0, n_rays = 0, n_hits = 0;
    #pragma omp parallel for schedule(dynamic, 4) reduction(+: s_miss, t_miss, s_hit, t_hit, s_sh, t_sh, n_rays, n_hits)
    for (size_t j = 0; j < height; ++j) {
        for (size_t i = 0; i < width; ++i) {
            size_t pix = width * j + i, index = 3 * pix;                     /* :107 */
            float u = 2 * ((float)i + 0.5f) / (float)width - 1.0f;           /* :109 */
            float v = 2 * ((float)j + 0.5f) / (float)height - 1.0f;          /* :110 */
            vec3 view = vnormalize(vadd(vadd(vscale(image_u, u), vscale(image_v, v)), dir)); /* :111 */
            co_ray ray = {eye, view, 0.0f, FLT_MAX};                          /* :113 */
            co_stats st = {0, 0};
            co_hit hit = traverse(&s, ray, &st);                             /* :114 */
            n_rays++;
            float rgb[3] = {0, 0, 0};
            uint8_t stt = 0;
            if (!hit.hit) { s_miss += st.steps; t_miss += st.tests; }
            else {
                s_hit += st.steps; t_hit += st.tests;
                n_hits++;                                                    /* :119 */
                const co_tri* tr = &s.tris[hit.id];
                vec3 normal = vnormalize(tr->n);                             /* :122 */
                float hu = hit.u, hv = hit.v;
                vec3 P = vadd(vadd(vscale(tr->p0, hu), vscale(tri_p1(tr), hv)), vscale(tri_p2(tr), 1 - hu - hv)); /* :129 (Q1) */
                float scale = -0.00001f;                                     /* :132 */
                P = vadd(P, vscale(normal, scale));                          /* :133 */
                vec3 sun_line = vnormalize(vsub(sun_position, P));           /* :135 */
                int occluded = 0;
                if (shadows) {
                    co_ray sray = {P, sun_line, 0.0f, FLT_MAX};               /* :136 */
                    co_stats st2 = {0, 0};
                    co_hit h2 = traverse(&s, sray, &st2);                    /* :137 closest-hit, only bool used */
                    n_rays++;
                    s_sh += st2.steps; t_sh += st2.tests;
                    occluded = h2.hit;
                }
                if (!occluded) { smooth_shading(sun_line, s.norms + 9 * (size_t)hit.id, view, hu, hv, rgb); stt = 1; }
                else { n_hits++; stt = 2; }                                  /* :148 */
            }
            if (pixels) { pixels[index] = rgb[0]; pixels[index + 1] = rgb[1]; pixels[index + 2] = rgb[2]; }
            if (ids) ids[pix] = hit.hit ? hit.id : CO_MISS;
            if (state) state[pix] = stt;
            if (tuv) { tuv[index] = hit.hit ? hit.t : 0; tuv[index + 1] = hit.hit ? hit.u : 0; tuv[index + 2] = hit.hit ? hit.v : 0; }
        }
    }
    if (stats) {
        stats[0] = s_miss; stats[1] = t_miss; stats[2] = s_hit; stats[3] = t_hit;
        stats[4] = s_sh; stats[5] = t_sh; stats[6] = n_rays; stats[7] = n_hits;
    }
}

/* ---- known-answer helpers ------------------------------------------------------------------------------------ */
/* rays: 8 floats each (o[3], d[3], tmin, tmax) */
void co_trace_rays(const void* nodes, size_t node_count, const uint64_t* prim_idx, const void* tris, size_t tri_count,
                   const float* rays, size_t n, uint32_t* out_id, float* out_tuv) {
    co_scene s = {(const co_node*)nodes, prim_idx, (const co_tri*)tris, NULL, node_count, tri_count};
    #pragma omp parallel for schedule(dynamic, 256)
    for (size_t k = 0; k < n; ++k) {
        const float* r = rays + 8 * k;
        co_ray ray = {v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), r[6], r[7]};
        co_stats st = {0, 0};
        co_hit h = traverse(&s, ray, &st);
        out_id[k] = h.hit ? h.id : CO_MISS;
        out_tuv[3 * k] = h.hit ? h.t : 0; out_tuv[3 * k + 1] = h.hit ? h.u : 0; out_tuv[3 * k + 2] = h.hit ? h.v : 0;
    }
}

void co_slab(const float* boxes, const float* rays, size_t n, float* out_entry_exit) {
    for (size_t k = 0; k < n; ++k) {
        const float* r = rays + 8 * k;
        co_ray ray = {v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), r[6], r[7]};
        co_nodeint ni; nodeint_init(&ni, &ray);
        slab(&ni, boxes + 6 * k, ray.tmin, ray.tmax, &out_entry_exit[2 * k], &out_entry_exit[2 * k + 1]);
    }
}

/* tris: 9 floats (p0,p1,p2) */
void co_tri_intersect(const float* tris, const float* rays, size_t n, uint8_t* out_hit, float* out_tuv) {
    for (size_t k = 0; k < n; ++k) {
        const float* p = tris + 9 * k;
        const float* r = rays + 8 * k;
        co_tri t = tri_make(v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8]));
        co_ray ray = {v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), r[6], r[7]};
        float tt = 0, u = 0, v = 0;
        int h = tri_intersect(&t, &ray, &tt, &u, &v);
        out_hit[k] = (uint8_t)h;
        out_tuv[3 * k] = h ? tt : 0; out_tuv[3 * k + 1] = h ? u : 0; out_tuv[3 * k + 2] = h ? v : 0;
    }
}

/* in: sun_line[3], norms[9], view[3], u, v */
void co_smooth_shading(const float* in, size_t n, float* out_rgb) {
    for (size_t k = 0; k < n; ++k) {
        const float* p = in + 17 * k;
        smooth_shading(v3(p[0], p[1], p[2]), p + 3, v3(p[12], p[13], p[14]), p[15], p[16], out_rgb + 3 * k);
    }
}

/* verts 9 floats per triangle -> 12 floats {p0,e1,e2,n} (triangle.hpp:30-34) */
void co_make_triangles(const float* verts, size_t n, float* out) {
    for (size_t k = 0; k < n; ++k) {
        const float* p = verts + 9 * k;
        co_tri t = tri_make(v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8]));
        memcpy(out + 12 * k, &t, sizeof t);
    }
}

/* include/render.hpp:24-44 rotate_triangles<Axis>: rebuilds each Triangle from rotated p0,p1(),p2() */
void co_rotate_triangles(int axis, float degrees, float* tris12, size_t n) {
    const float pi = 3.14159265359f;
    float c = cosf(degrees * pi / 180.0f);
    float s = sinf(degrees * pi / 180.0f);
    for (size_t k = 0; k < n; ++k) {
        co_tri* t = (co_tri*)(tris12 + 12 * k);
        vec3 p[3]; p[0] = t->p0; p[1] = tri_p1(t); p[2] = tri_p2(t);
        for (int q = 0; q < 3; ++q) {
            vec3 a = p[q];
            if (axis == 0) p[q] = v3(a.v[0], a.v[1] * c - a.v[2] * s, a.v[1] * s + a.v[2] * c);
            else if (axis == 1) p[q] = v3(a.v[0] * c + a.v[2] * s, a.v[1], -a.v[0] * s + a.v[2] * c);
            else p[q] = v3(a.v[0] * c - a.v[1] * s, a.v[0] * s + a.v[1] * c, a.v[2]);
        }
        *t = tri_make(p[0], p[1], p[2]);
    }
}

/* include/transform.hpp:106-129: p' = a*p + v, each row summed left to right; the Triangle is rebuilt from the
 * transformed p0, p1(), p2() (triangle.hpp:30-37).  a is row-major 3x3. */
void co_transform_triangles(const float* a, const float* tv, float* tris12, size_t n) {
    for (size_t k = 0; k < n; ++k) {
        co_tri* t = (co_tri*)(tris12 + 12 * k);
        vec3 p[3]; p[0] = t->p0; p[1] = tri_p1(t); p[2] = tri_p2(t);
        for (int q = 0; q < 3; ++q) {
            vec3 x = p[q];
            p[q] = v3(a[0] * x.v[0] + a[1] * x.v[1] + a[2] * x.v[2] + tv[0],
                      a[3] * x.v[0] + a[4] * x.v[1] + a[5] * x.v[2] + tv[1],
                      a[6] * x.v[0] + a[7] * x.v[1] + a[8] * x.v[2] + tv[2]);
        }
        *t = tri_make(p[0], p[1], p[2]);
    }
}

/* lib/bvh/hierarchy_refitter.hpp:17-32 + bottom_up_algorithm.hpp:47-85: every leaf box becomes the union of its
 * primitives' bounding boxes (triangle.hpp:39-44; min/max as bounding_box.hpp:23-33 -> vector.hpp min/max =
 * std::min/std::max per component), every inner box the union of its two children.  Sequential post-order here;
 * min/max are exact, so the order of the reference's parallel walk does not matter.
 * nodes: 32-byte reference nodes {bounds[6] = minx,maxx,miny,maxy,minz,maxz; count; first}. */
static float fmin_std(float a, float b) { return b < a ? b : a; }   /* std::min(a, b) */
static float fmax_std(float a, float b) { return a < b ? b : a; }   /* std::max(a, b) */
void co_refit(void* nodes_v, size_t node_count, const size_t* prim_idx, const float* tris12) {
    co_node* nodes = (co_node*)nodes_v;
    /* children always follow their parent in neither layout, so walk with an explicit post-order stack */
    uint32_t* stack = (uint32_t*)malloc(sizeof(uint32_t) * 2 * 128);
    size_t cap = 256, sp = 0;
    stack[sp++] = 0;                       /* entries: node index << 1 | visited */
    while (sp) {
        uint32_t e = stack[--sp];
        uint32_t i = e >> 1;
        co_node* nd = &nodes[i];
        if (nd->primitive_count) {
            float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
            for (uint32_t k = 0; k < nd->primitive_count; ++k) {
                const co_tri* t = (const co_tri*)(tris12 + 12 * prim_idx[nd->first_child_or_primitive + k]);
                vec3 p1 = tri_p1(t), p2 = tri_p2(t);
                for (int c = 0; c < 3; ++c) {
                    /* bbox(p0).extend(p1).extend(p2), then accumulated into the leaf box */
                    float mn = fmin_std(fmin_std(t->p0.v[c], p1.v[c]), p2.v[c]);
                    float mx = fmax_std(fmax_std(t->p0.v[c], p1.v[c]), p2.v[c]);
                    lo[c] = fmin_std(lo[c], mn); hi[c] = fmax_std(hi[c], mx);
                }
            }
            for (int c = 0; c < 3; ++c) { nd->bounds[2 * c] = lo[c]; nd->bounds[2 * c + 1] = hi[c]; }
        } else if (e & 1u) {
            const co_node* l = &nodes[nd->first_child_or_primitive]; const co_node* r = l + 1;
            for (int c = 0; c < 3; ++c) {
                nd->bounds[2 * c] = fmin_std(l->bounds[2 * c], r->bounds[2 * c]);
                nd->bounds[2 * c + 1] = fmax_std(l->bounds[2 * c + 1], r->bounds[2 * c + 1]);
            }
        } else {
            if (sp + 3 > cap) { cap *= 2; stack = (uint32_t*)realloc(stack, sizeof(uint32_t) * cap); }
            stack[sp++] = (i << 1) | 1u;
            stack[sp++] = (nd->first_child_or_primitive + 1) << 1;
            stack[sp++] = nd->first_child_or_primitive << 1;
        }
    }
    free(stack);
    (void)node_count;
}

int co_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
