// oracle/ref_driver_f64.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The Scalar = double instantiation of the reference (`anim -d`, include/anim.cpp:154-155), compiled from the
// UNMODIFIED headers where they lie into oracle/_ref/libceres_ref64_strict.so by oracle/Makefile.  Same structure as
// oracle/ref_driver.cpp (which see for the list of reference entry points called); only what the double-precision
// parity tests need: scene from an OBJ file / from POD arrays, the unmodified render<double>(), and the instrumented
// loop that records per-pixel primitive ids.
//
// Layouts in double precision: Bvh<double>::Node is 64 bytes (6 doubles + two uint64 indices, lib/bvh/bvh.hpp:20-28),
// Triangle<double> 96 bytes (lib/bvh/triangle.hpp:27), tri_norms entries 72 bytes (lib/obj_norms.hpp:113-115).
#include <cstdint>
#include <cstring>
#include <vector>
#include <array>
#include <memory>
#include <chrono>

#include <bvh/bvh.hpp>
#include <bvh/sweep_sah_builder.hpp>
#include <bvh/node_layout_optimizer.hpp>
#include <bvh/single_ray_traverser.hpp>
#include <bvh/primitive_intersectors.hpp>
#include <bvh/triangle.hpp>

#include "render.hpp"
#include "obj_norms.hpp"

using Scalar   = double;
using Vector3  = bvh::Vector3<Scalar>;
using Triangle = bvh::Triangle<Scalar>;
using Bvh      = bvh::Bvh<Scalar>;
using Norms    = std::array<Vector3, 3>;

static_assert(sizeof(Triangle) == 96, "Triangle<double> is 96 bytes");
static_assert(sizeof(Bvh::Node) == 64, "Bvh<double>::Node is 64 bytes");
static_assert(sizeof(Norms) == 72, "tri_norms entry is 72 bytes");

struct RefScene64 {
    std::vector<Triangle> triangles;
    std::vector<Norms>    norms;
    Bvh                   bvh;
};

static Camera<Scalar> make_camera(const double* c) {
    return Camera<Scalar>{Vector3(c[0], c[1], c[2]), Vector3(c[3], c[4], c[5]), Vector3(c[6], c[7], c[8]), c[9]};
}

extern "C" {

// rot_axis: -1 none, 0/1/2 = x/y/z.  BVH as anim.cpp:46-65 (SweepSah + NodeLayoutOptimizer).
void* cref64_scene_load_obj(const char* path, int rot_axis, double rot_deg) {
    auto loaded = obj::load_from_file<Scalar>(path);
    if (loaded.first.empty()) return nullptr;
    auto* s = new RefScene64;
    s->triangles = std::move(loaded.first);
    s->norms = std::move(loaded.second);
    if (rot_axis == 0) rotate_triangles<0>(Scalar(rot_deg), s->triangles.data(), s->triangles.size());
    else if (rot_axis == 1) rotate_triangles<1>(Scalar(rot_deg), s->triangles.data(), s->triangles.size());
    else if (rot_axis == 2) rotate_triangles<2>(Scalar(rot_deg), s->triangles.data(), s->triangles.size());
    auto bc = bvh::compute_bounding_boxes_and_centers(s->triangles.data(), s->triangles.size());
    auto global_bbox = bvh::compute_bounding_boxes_union(bc.first.get(), s->triangles.size());
    bvh::SweepSahBuilder<Bvh> b(s->bvh);
    b.build(global_bbox, bc.first.get(), bc.second.get(), s->triangles.size());
    bvh::NodeLayoutOptimizer<Bvh> opt(s->bvh);
    opt.optimize();
    return s;
}

void* cref64_scene_from_pod(const void* nodes, size_t node_count, const size_t* prim_idx,
                            const void* tris, const void* norms, size_t tri_count) {
    auto* s = new RefScene64;
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

void cref64_scene_destroy(void* h) { delete static_cast<RefScene64*>(h); }
size_t cref64_scene_node_count(void* h) { return static_cast<RefScene64*>(h)->bvh.node_count; }
size_t cref64_scene_tri_count(void* h) { return static_cast<RefScene64*>(h)->triangles.size(); }
const void* cref64_scene_nodes(void* h) { return static_cast<RefScene64*>(h)->bvh.nodes.get(); }
const size_t* cref64_scene_prim_indices(void* h) { return static_cast<RefScene64*>(h)->bvh.primitive_indices.get(); }
const void* cref64_scene_triangles(void* h) { return static_cast<RefScene64*>(h)->triangles.data(); }
const void* cref64_scene_norms(void* h) { return static_cast<RefScene64*>(h)->norms.data(); }

// The unmodified render<double>() (render.hpp:86-156).  cam = eye[3], dir[3], up[3], fov.  Returns seconds.
double cref64_render(void* h, const double* cam, const double* sun, double* pixels, size_t W, size_t H, int* rays, int* hits) {
    auto* s = static_cast<RefScene64*>(h);
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

// The statements of render.hpp:91-153 run through the reference's traverser, recording what render() does not return:
// ids u32[W*H] (0xFFFFFFFF miss), state u8[W*H] (0 miss, 1 lit, 2 occluded), counters[2] = {rays, hits}.  All nullable.
// flags bit0: no shadow ray.
void cref64_render_ex(void* h, const double* cam, const double* sun, size_t width, size_t height, unsigned flags,
                      double* pixels, uint32_t* ids, uint8_t* state, uint64_t* counters) {
    auto* s = static_cast<RefScene64*>(h);
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
    uint64_t n_rays = 0, n_hits = 0;
    #pragma omp parallel for schedule(dynamic, 4) reduction(+: n_rays, n_hits)
    for (size_t j = 0; j < height; ++j) {
        for (size_t i = 0; i < width; ++i) {
            size_t pix = width * j + i;
            size_t index = 3 * pix;
            auto u = 2 * (i + Scalar(0.5)) / Scalar(width)  - Scalar(1);
            auto v = 2 * (j + Scalar(0.5)) / Scalar(height) - Scalar(1);
            auto view = bvh::normalize(image_u * u + image_v * v + dir);
            bvh::Ray<Scalar> ray(camera.eye, view);
            auto hit = traverser.traverse(ray, intersector);
            n_rays++;
            Scalar rgb[3] = {0, 0, 0};
            uint32_t id = 0xFFFFFFFFu; uint8_t stt = 0;
            if (hit) {
                n_hits++;
                auto ind = hit->primitive_index;
                id = uint32_t(ind);
                auto tri = triangles[ind];
                auto normal = bvh::normalize(tri.n);
                auto u = hit->intersection.u;
                auto v = hit->intersection.v;
                Vector3 intersect_point = (u*tri.p0 + v*tri.p1() + (1-u-v)*tri.p2());
                Scalar scale = -0.00001;
                intersect_point = intersect_point + scale*normal;
                Vector3 sun_line = bvh::normalize(sun_position - intersect_point);
                bool occluded = false;
                if (shadows) {
                    bvh::Ray<Scalar> sray(intersect_point, sun_line);
                    auto hit2 = traverser.traverse(sray, intersector);
                    n_rays++;
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
        }
    }
    if (counters) { counters[0] = n_rays; counters[1] = n_hits; }
}

// Turntable camera sequence in double: the statements of anim.cpp:77-89 with Scalar = double (Transform::rotate about
// `axis`, applied to the PREVIOUS camera each frame).  cams_out: n_frames x 13 doubles (eye, dir, up, fov, sun).
void cref64_turntable(const double* cam, const double* sun, const double* axis, int n_frames, double* cams_out) {
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
        double* o = cams_out + 13 * i;
        for (int k = 0; k < 3; ++k) { o[k] = camera.eye[k]; o[3 + k] = camera.dir[k]; o[6 + k] = camera.up[k]; o[10 + k] = sun_position[k]; }
        o[9] = camera.fov;
    }
}

} // extern "C"
