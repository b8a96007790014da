// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A C-ABI wrapper around the UNMODIFIED reference headers, compiled from where they lie
// (-I/root/reference/lib -I/root/reference/include) into oracle/_ref/libceres_ref_{strict,fast}.so by
// oracle/Makefile.  Nothing from /root/reference is copied into this repository; this file only *calls*
// the reference:
//   obj::load_from_file            lib/obj_norms.hpp:120-127
//   rotate_triangles<Axis>         include/render.hpp:24-44
//   compute_bounding_boxes_*       lib/bvh/utilities.hpp:142-171
//   SweepSahBuilder / NodeLayoutOptimizer   (as include/anim.cpp:62-65)
//   BinnedSahBuilder<Bvh,16>       (as include/static.cpp:106-107)
//   render()                       include/render.hpp:86-156      (cref_render: the unmodified function)
//   SingleRayTraverser::traverse   lib/bvh/single_ray_traverser.hpp:142-163 (cref_render_ex / cref_trace_rays)
//   smooth_shading                 include/render.hpp:56-84
//   transform_triangles / Transform::operator()   include/transform.hpp:106-129   (cref_transform_triangles)
//   HierarchyRefitter::refit       lib/bvh/hierarchy_refitter.hpp:17-45            (cref_refit)
//
// cref_render_ex re-runs the loop body of render.hpp:105-153 through the reference's own traverser with the
// Statistics overload, so that per-pixel primitive ids, shadow verdicts and S/T step counts can be recorded
// (the reference's render() does not return them).  tests/test_oracle_ref.py checks that its RGB output is
// identical to the unmodified render() of the same build.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
#include <cstdint>
#include <cstring>
#include <vector>
#include <array>
#include <string>
#include <memory>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

#include <bvh/bvh.hpp>
#include <bvh/sweep_sah_builder.hpp>
#include <bvh/binned_sah_builder.hpp>
#include <bvh/node_layout_optimizer.hpp>
#include <bvh/single_ray_traverser.hpp>
#include <bvh/primitive_intersectors.hpp>
#include <bvh/triangle.hpp>
#include <bvh/hierarchy_refitter.hpp>

#include "render.hpp"
#include "obj_norms.hpp"

using Scalar   = float;
using Vector3  = bvh::Vector3<Scalar>;
using Triangle = bvh::Triangle<Scalar>;
using Bvh      = bvh::Bvh<Scalar>;
using Norms    = std::array<Vector3, 3>;

static_assert(sizeof(Triangle) == 48, "Triangle<float> is 48 bytes");
static_assert(sizeof(Bvh::Node) == 32, "Bvh<float>::Node is 32 bytes");
static_assert(sizeof(Norms) == 36, "tri_norms entry is 36 bytes");

struct RefScene {
    std::vector<Triangle> triangles;
    std::vector<Norms>    norms;
    Bvh                   bvh;
    double                build_seconds = 0;
};

static void build_bvh(RefScene* s, int builder) {
    using namespace std::chrono;
    auto t0 = high_resolution_clock::now();
    auto bc = bvh::compute_bounding_boxes_and_centers(s->triangles.data(), s->triangles.size());
    auto bboxes = bc.first.get();
    auto centers = bc.second.get();
    auto global_bbox = bvh::compute_bounding_boxes_union(bboxes, s->triangles.size());
    if (builder == 1) {
        bvh::BinnedSahBuilder<Bvh, 16> b(s->bvh);
        b.build(global_bbox, bboxes, centers, s->triangles.size());
    } else {
        bvh::SweepSahBuilder<Bvh> b(s->bvh);
        b.build(global_bbox, bboxes, centers, s->triangles.size());
        bvh::NodeLayoutOptimizer<Bvh> opt(s->bvh);
        opt.optimize();
    }
    s->build_seconds = duration_cast<microseconds>(high_resolution_clock::now() - t0).count() / 1e6;
}

extern "C" {

// builder: 0 = SweepSah + NodeLayoutOptimizer (anim.cpp), 1 = BinnedSah<16> (static.cpp)
// rot_axis: -1 none, 0/1/2 = x/y/z (static.cpp:83-88)
void* cref_scene_load_obj(const char* path, int rot_axis, float rot_deg, int builder) {
    auto loaded = obj::load_from_file<Scalar>(path);
    if (loaded.first.empty()) return nullptr;
    auto* s = new RefScene;
    s->triangles = std::move(loaded.first);
    s->norms = std::move(loaded.second);
    if (rot_axis == 0) rotate_triangles<0>(rot_deg, s->triangles.data(), s->triangles.size());
    else if (rot_axis == 1) rotate_triangles<1>(rot_deg, s->triangles.data(), s->triangles.size());
    else if (rot_axis == 2) rotate_triangles<2>(rot_deg, s->triangles.data(), s->triangles.size());
    build_bvh(s, builder);
    return s;
}

// verts: 9 floats per triangle (p0,p1,p2); norms: 9 floats per triangle (may be null -> zeros)
void* cref_scene_from_triangles(const float* verts, const float* norms, size_t n, int builder) {
    if (!n) return nullptr;
    auto* s = new RefScene;
    s->triangles.resize(n);
    s->norms.resize(n);
    #pragma omp parallel for
    for (size_t i = 0; i < n; ++i) {
        const float* v = verts + 9 * i;
        s->triangles[i] = Triangle(Vector3(v[0], v[1], v[2]), Vector3(v[3], v[4], v[5]), Vector3(v[6], v[7], v[8]));
        if (norms) {
            const float* m = norms + 9 * i;
            s->norms[i] = Norms{Vector3(m[0], m[1], m[2]), Vector3(m[3], m[4], m[5]), Vector3(m[6], m[7], m[8])};
        } else {
            s->norms[i] = Norms{Vector3(0, 0, 0), Vector3(0, 0, 0), Vector3(0, 0, 0)};
        }
    }
    build_bvh(s, builder);
    return s;
}

// Adopt an existing flattened scene (POD layouts of the reference) without rebuilding the BVH.
void* cref_scene_from_pod(const void* nodes, size_t node_count, const size_t* prim_idx,
                          const void* tris, const void* norms, size_t tri_count) {
    auto* s = new RefScene;
    s->triangles.resize(tri_count);
    s->norms.resize(tri_count);
    std::memcpy((void*)s->triangles.data(), tris, tri_count * sizeof(Triangle));
    std::memcpy((void*)s->norms.data(), norms, tri_count * sizeof(Norms));
    s->bvh.nodes = std::make_unique<Bvh::Node[]>(node_count);
    std::memcpy((void*)s->bvh.nodes.get(), nodes, node_count * sizeof(Bvh::Node));
    s->bvh.primitive_indices = std::make_unique<size_t[]>(tri_count);
    std::memcpy(s->bvh.primitive_indices.get(), prim_idx, tri_count * sizeof(size_t));
    s->bvh.node_count = node_count;
    return s;
}

// ---- scene preparation ----------------------------------------------------------------------
// tris: n x 12 floats {p0,e1,e2,n}, transformed in place by the reference's transform_triangles with
// Transform{a (row-major 3x3), v}.
void cref_transform_triangles(const float* a, const float* v, float* tris, size_t n) {
    Transform<Scalar> t;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) t.a[r][c] = a[3 * r + c];
    t.v = Vector3(v[0], v[1], v[2]);
    transform_triangles(t, reinterpret_cast<Triangle*>(tris), n);
}

void cref_rotate_triangles(int axis, float degrees, float* tris, size_t n) {
    Triangle* t = reinterpret_cast<Triangle*>(tris);
    if (axis == 0) rotate_triangles<0>(degrees, t, n);
    else if (axis == 1) rotate_triangles<1>(degrees, t, n);
    else rotate_triangles<2>(degrees, t, n);
}

// Refit the scene's BVH to `tris` (n x 12 floats in primitive order; they also replace the scene's triangles) with
// the reference's HierarchyRefitter.  The reference ships no caller of refit(); the leaf update used here is the
// builders' own definition of a leaf box: the union of Triangle::bounding_box() over the leaf's primitives in slot
// order (triangle.hpp:39-44, bounding_box.hpp:23-33).
void cref_refit(void* h, const float* tris) {
    auto* s = static_cast<RefScene*>(h);
    std::memcpy((void*)s->triangles.data(), tris, s->triangles.size() * sizeof(Triangle));
    bvh::HierarchyRefitter<Bvh> refitter(s->bvh);
    refitter.refit([&](Bvh::Node& leaf) {
        auto bbox = bvh::BoundingBox<Scalar>::empty();
        for (size_t k = 0; k < leaf.primitive_count; ++k)
            bbox.extend(s->triangles[s->bvh.primitive_indices[leaf.first_child_or_primitive + k]].bounding_box());
        leaf.bounding_box_proxy() = bbox;
    });
}

void cref_scene_destroy(void* h) { delete static_cast<RefScene*>(h); }
size_t cref_scene_node_count(void* h) { return static_cast<RefScene*>(h)->bvh.node_count; }
size_t cref_scene_tri_count(void* h) { return static_cast<RefScene*>(h)->triangles.size(); }
double cref_scene_build_seconds(void* h) { return static_cast<RefScene*>(h)->build_seconds; }
const void* cref_scene_nodes(void* h) { return static_cast<RefScene*>(h)->bvh.nodes.get(); }
const size_t* cref_scene_prim_indices(void* h) { return static_cast<RefScene*>(h)->bvh.primitive_indices.get(); }
const void* cref_scene_triangles(void* h) { return static_cast<RefScene*>(h)->triangles.data(); }
const void* cref_scene_norms(void* h) { return static_cast<RefScene*>(h)->norms.data(); }
int cref_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

static Camera<Scalar> make_camera(const float* c) {
    return Camera<Scalar>{Vector3(c[0], c[1], c[2]), Vector3(c[3], c[4], c[5]), Vector3(c[6], c[7], c[8]), c[9]};
}

// The unmodified reference render() (render.hpp:86-156). cam = eye[3],dir[3],up[3],fov. Returns seconds.
double cref_render(void* h, const float* cam, const float* sun, float* pixels, size_t W, size_t H,
                   int* rays, int* hits) {
    auto* s = static_cast<RefScene*>(h);
    auto camera = make_camera(cam);
    Vector3 sun_position(sun[0], sun[1], sun[2]);
    using namespace std::chrono;
    auto t0 = high_resolution_clock::now();
    auto r = render(camera, sun_position, s->bvh, s->triangles.data(), s->norms.data(), pixels, W, H);
    auto t1 = high_resolution_clock::now();
    if (rays) *rays = r.first;
    if (hits) *hits = r.second;
    return duration_cast<nanoseconds>(t1 - t0).count() / 1e9;
}

// Instrumented loop: same statements as render.hpp:91-153, run through the reference traverser with
// Statistics. flags bit0 = no shadow ray (shade every primary hit; D3 "primary rays only").
// Optional outputs (nullable): ids u32[W*H] (0xFFFFFFFF miss), state u8[W*H] (0 miss,1 lit,2 occluded),
// tuv float[3*W*H], stats u64[8] = {S_miss,T_miss,S_hit,T_hit,S_shadow,T_shadow,rays,hits}.
double cref_render_ex(void* h, const float* cam, const float* sun, size_t width, size_t height, unsigned flags,
                      float* pixels, uint32_t* ids, uint8_t* state, float* tuv, uint64_t* stats) {
    auto* s = static_cast<RefScene*>(h);
    auto camera = make_camera(cam);
    Vector3 sun_position(sun[0], sun[1], sun[2]);
    const Bvh& bvh = s->bvh;
    const Triangle* triangles = s->triangles.data();
    const Norms* tri_norms = s->norms.data();
    const bool shadows = !(flags & 1u);

    auto dir = bvh::normalize(camera.dir);
    auto image_u = bvh::normalize(bvh::cross(dir, camera.up));
    auto image_v = bvh::normalize(bvh::cross(image_u, dir));
    auto image_w = std::tan(camera.fov * Scalar(3.14159265 * (1.0 / 180.0) * 0.5));
    auto ratio = Scalar(height) / Scalar(width);
    image_u = image_u * image_w;
    image_v = image_v * image_w * ratio;

    bvh::ClosestPrimitiveIntersector<Bvh, Triangle, false> intersector(bvh, triangles);
    bvh::SingleRayTraverser<Bvh> traverser(bvh);
    using Stats = bvh::SingleRayTraverser<Bvh>::Statistics;

    uint64_t s_miss = 0, t_miss = 0, s_hit = 0, t_hit = 0, s_sh = 0, t_sh = 0, n_rays = 0, n_hits = 0;
    using namespace std::chrono;
    auto t0 = high_resolution_clock::now();
    #pragma omp parallel for schedule(dynamic, 4) reduction(+: s_miss, t_miss, s_hit, t_hit, s_sh, t_sh, n_rays, n_hits)
    for (size_t j = 0; j < height; ++j) {
        for (size_t i = 0; i < width; ++i) {
            size_t pix = width * j + i;
            size_t index = 3 * pix;
            auto u = 2 * (i + Scalar(0.5)) / Scalar(width)  - Scalar(1);
            auto v = 2 * (j + Scalar(0.5)) / Scalar(height) - Scalar(1);
            auto view = bvh::normalize(image_u * u + image_v * v + dir);
            bvh::Ray<Scalar> ray(camera.eye, view);
            Stats st;
            auto hit = traverser.traverse(ray, intersector, st);
            n_rays++;
            float rgb[3] = {0, 0, 0};
            uint32_t id = 0xFFFFFFFFu; uint8_t stt = 0; float ht = 0, hu = 0, hv = 0;
            if (!hit) {
                s_miss += st.traversal_steps; t_miss += st.intersections;
            } else {
                s_hit += st.traversal_steps; t_hit += st.intersections;
                n_hits++;
                auto ind = hit->primitive_index;
                id = uint32_t(ind);
                auto tri = triangles[ind];
                auto normal = bvh::normalize(tri.n);
                auto u = hit->intersection.u;
                auto v = hit->intersection.v;
                ht = hit->intersection.t; hu = u; hv = v;
                Vector3 intersect_point = (u*tri.p0 + v*tri.p1() + (1-u-v)*tri.p2());
                Scalar scale = -0.00001;
                intersect_point = intersect_point + scale*normal;
                Vector3 sun_line = bvh::normalize(sun_position - intersect_point);
                bool occluded = false;
                if (shadows) {
                    bvh::Ray<Scalar> sray(intersect_point, sun_line);
                    Stats st2;
                    auto hit2 = traverser.traverse(sray, intersector, st2);
                    n_rays++;
                    s_sh += st2.traversal_steps; t_sh += st2.intersections;
                    occluded = bool(hit2);
                }
                if (!occluded) {
                    auto c = smooth_shading(sun_line, tri_norms[ind], view, u, v);
                    rgb[0] = c[0]; rgb[1] = c[1]; rgb[2] = c[2];
                    stt = 1;
                } else {
                    n_hits++;
                    stt = 2;
                }
            }
            if (pixels) { pixels[index] = rgb[0]; pixels[index + 1] = rgb[1]; pixels[index + 2] = rgb[2]; }
            if (ids) ids[pix] = id;
            if (state) state[pix] = stt;
            if (tuv) { tuv[index] = ht; tuv[index + 1] = hu; tuv[index + 2] = hv; }
        }
    }
    auto t1 = high_resolution_clock::now();
    if (stats) {
        stats[0] = s_miss; stats[1] = t_miss; stats[2] = s_hit; stats[3] = t_hit;
        stats[4] = s_sh; stats[5] = t_sh; stats[6] = n_rays; stats[7] = n_hits;
    }
    return duration_cast<nanoseconds>(t1 - t0).count() / 1e9;
}

// Known-answer helpers ---------------------------------------------------------------------------------
// rays: 8 floats each (origin[3], direction[3], tmin, tmax). out_id u32 (0xFFFFFFFF miss), out_tuv 3 floats.
void cref_trace_rays(void* h, const float* rays, size_t n, uint32_t* out_id, float* out_tuv) {
    auto* s = static_cast<RefScene*>(h);
    bvh::ClosestPrimitiveIntersector<Bvh, Triangle, false> intersector(s->bvh, s->triangles.data());
    bvh::SingleRayTraverser<Bvh> traverser(s->bvh);
    #pragma omp parallel for schedule(dynamic, 256)
    for (size_t k = 0; k < n; ++k) {
        const float* r = rays + 8 * k;
        bvh::Ray<Scalar> ray(Vector3(r[0], r[1], r[2]), Vector3(r[3], r[4], r[5]), r[6], r[7]);
        auto hit = traverser.traverse(ray, intersector);
        out_id[k] = hit ? uint32_t(hit->primitive_index) : 0xFFFFFFFFu;
        out_tuv[3 * k] = hit ? hit->intersection.t : 0;
        out_tuv[3 * k + 1] = hit ? hit->intersection.u : 0;
        out_tuv[3 * k + 2] = hit ? hit->intersection.v : 0;
    }
}

// Slab test (node_intersectors.hpp:34-47,82-103). boxes: 6 floats {minx,maxx,miny,maxy,minz,maxz}.
void cref_slab(const float* boxes, const float* rays, size_t n, float* out_entry_exit) {
    for (size_t k = 0; k < n; ++k) {
        const float* r = rays + 8 * k;
        bvh::Ray<Scalar> ray(Vector3(r[0], r[1], r[2]), Vector3(r[3], r[4], r[5]), r[6], r[7]);
        Bvh::Node node;
        for (int q = 0; q < 6; ++q) node.bounds[q] = boxes[6 * k + q];
        node.primitive_count = 0; node.first_child_or_primitive = 0;
        bvh::FastNodeIntersector<Bvh> ni(ray);
        auto d = ni.intersect(node, ray);
        out_entry_exit[2 * k] = d.first;
        out_entry_exit[2 * k + 1] = d.second;
    }
}

// Triangle test (triangle.hpp:95-115). tris: 9 floats (p0,p1,p2). out: hit flag + t,u,v.
void cref_tri_intersect(const float* tris, const float* rays, size_t n, uint8_t* out_hit, float* out_tuv) {
    for (size_t k = 0; k < n; ++k) {
        const float* v = tris + 9 * k;
        const float* r = rays + 8 * k;
        Triangle tri(Vector3(v[0], v[1], v[2]), Vector3(v[3], v[4], v[5]), Vector3(v[6], v[7], v[8]));
        bvh::Ray<Scalar> ray(Vector3(r[0], r[1], r[2]), Vector3(r[3], r[4], r[5]), r[6], r[7]);
        auto hit = tri.intersect(ray);
        out_hit[k] = hit ? 1 : 0;
        out_tuv[3 * k] = hit ? hit->t : 0;
        out_tuv[3 * k + 1] = hit ? hit->u : 0;
        out_tuv[3 * k + 2] = hit ? hit->v : 0;
    }
}

// smooth_shading (render.hpp:56-84). in: sun_line[3], norms[9], view[3], u, v = 17 floats. out rgb.
void cref_smooth_shading(const float* in, size_t n, float* out_rgb) {
    for (size_t k = 0; k < n; ++k) {
        const float* p = in + 17 * k;
        Norms nm{Vector3(p[3], p[4], p[5]), Vector3(p[6], p[7], p[8]), Vector3(p[9], p[10], p[11])};
        auto c = smooth_shading(Vector3(p[0], p[1], p[2]), nm, Vector3(p[12], p[13], p[14]), p[15], p[16]);
        out_rgb[3 * k] = c[0]; out_rgb[3 * k + 1] = c[1]; out_rgb[3 * k + 2] = c[2];
    }
}

// Turntable camera sequence, the statements of anim.cpp:77-89 (Transform::rotate about `axis`, applied to the
// PREVIOUS camera each frame).  cams_out: n_frames x 13 floats (eye,dir,up,fov,sun).
void cref_turntable(const float* cam, const float* sun, const float* axis, int n_frames, float* cams_out) {
    constexpr Scalar pi = Scalar(3.14159265359);
    auto camera = make_camera(cam);
    Vector3 sun_position(sun[0], sun[1], sun[2]);
    Scalar rotation_step_degrees = 360.0f / n_frames;
    auto t_cam = Transform<Scalar>().rotate(Vector3(axis[0], axis[1], axis[2]), rotation_step_degrees / 180.0f * pi);
    auto t_sun = Transform<Scalar>().rotate(Vector3(axis[0], axis[1], axis[2]), rotation_step_degrees / 180.0f * pi);
    for (int i = 0; i < n_frames; i++) {
        camera.eye = t_cam(camera.eye);
        camera.dir = t_cam(camera.dir);
        sun_position = t_sun(sun_position);
        float* o = cams_out + 13 * i;
        for (int k = 0; k < 3; ++k) { o[k] = camera.eye[k]; o[3 + k] = camera.dir[k]; o[6 + k] = camera.up[k]; o[10 + k] = sun_position[k]; }
        o[9] = camera.fov;
    }
}

} // extern "C"
