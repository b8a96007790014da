// include/ceres/render.hpp -- drop-in replacement for the reference's include/render.hpp.
//
// Put this directory before the reference's include/ on the include path (or copy the file over include/render.hpp)
// and link libceres_b200.so: static.cpp / anim.cpp / any caller of
//
//     std::pair<int,int> render(const Camera<Scalar>&, const bvh::Vector3<Scalar>& sun_position,
//                               const bvh::Bvh<Scalar>&, const bvh::Triangle<Scalar>* triangles,
//                               std::array<bvh::Vector3<Scalar>,3>* tri_norms, Scalar* pixels,
//                               size_t width, size_t height);                      (reference render.hpp:86-89)
//
// compile unchanged and render on the B200 through the C ABI of ceres_b200.h.  Same include guard as the reference
// header, same names: Camera (render.hpp:16-22), rotate_triangles<Axis> (render.hpp:24-44), render (render.hpp:86-156).
// The shading helpers lambertian / blinn_phong_spec / smooth_shading (render.hpp:46-84) are device code now and are
// not re-exported.
//
// Written from scratch against the POD layouts of lib/bvh (Node 32 B, Triangle<float> 48 B, tri_norms 36 B); the
// only reference code this header needs is the lib/bvh type definitions of the caller's own tree.
//
// Differences a caller can observe:
//   * STALE GEOMETRY IS POSSIBLE: the GPU copy of (bvh, triangles, tri_norms) is re-used while the same pointers, counts and
//     a SAMPLED fingerprint (root node, 64 windows of 32 bytes and the tail of each array) repeat; the reference reads the live
//     arrays every call.  A caller that edits a few triangles or nodes in place between calls must call
//     ceres::invalidate_scene() -- see the last item.
//   * Scalar = double (anim -d, anim.cpp:154-155) renders through ceres_render_f64: one kernel on one GPU, same
//     results as the reference's double build (hit ids bit for bit), without the float path's host-frame pipeline.
//   * tri_norms may be nullptr (static.cpp:130 passes nullptr, which does not even deduce against the reference
//     signature): normals are then taken as zero, as the reference would read them if it could.
//   * errors (no GPU, CUDA failure, malformed BVH) throw std::runtime_error; the reference has no error path.
//   * the caller's `pixels` buffer is page-locked (ceres_host_register) the first time it is seen and stays so while the
//     same (pointer, size) is passed again, so that the GPU writes the frame into it by DMA -- a plain make_unique /
//     malloc frame (anim.cpp:76, static.cpp:117) then costs the same per frame as a cudaHostAlloc'ed one (the first call pays
//     the registration, ~0.3 ms per MB).  A new (pointer, size) drops the old registration.  HAZARD: a caller that frees the
//     frame and gets the SAME address back from the allocator for a new one between two calls keeps a registration of
//     pages that no longer back the buffer; such callers set CERES_NO_AUTOPIN=1 (pageable path, ~2x slower at 4K) or call
//     ceres::release_frame() before freeing.  Float frames only.
//   * opt-in, CERES_FRAME_RETAINED=1 in the environment: render() stops zero-filling the ~70 % of the frame outside the scene's
//     screen rectangle on every call, on the caller's word that the frame is not written between calls (the reference's
//     anim.cpp only reads it).  Same frames; 0.84 -> 0.6 ms per 4K frame.  Off by default: render() then overwrites every pixel
//     whatever the buffer held, exactly as render.hpp:107,117 do.
//   * the scene (bvh, triangles, tri_norms) is flattened and uploaded on the first call and re-used while the same
//     arrays are passed again (the reference builds them once, outside the frame loop, anim.cpp:46-65).  The cache
//     key is the three pointers + counts + a sampled fingerprint of the contents; call ceres::invalidate_scene()
//     after modifying geometry in place if you want to be certain.
#ifndef __RENDER_H
#define __RENDER_H

#include <array>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

#include <bvh/bvh.hpp>
#include <bvh/triangle.hpp>
#include <bvh/vector.hpp>

#include "ceres_b200.h"
#if __has_include("transform.hpp")
#include "transform.hpp"      // the reference's render.hpp pulls it in for anim.cpp (render.hpp:14)
#endif

template <typename Scalar>
struct Camera {
    bvh::Vector3<Scalar> eye;
    bvh::Vector3<Scalar> dir;
    bvh::Vector3<Scalar> up;
    Scalar fov;
};

// render.hpp:24-44: rotation about a coordinate axis, every triangle rebuilt from its rotated corners so that the
// edges and the normal are recomputed by the Triangle constructor.  Host code (scene preparation, outside the
// timed path).
template <size_t Axis, typename Scalar>
static void rotate_triangles(Scalar degrees, bvh::Triangle<Scalar>* triangles, size_t triangle_count) {
    static_assert(Axis < 3, "axis is 0 (x), 1 (y) or 2 (z)");
    const Scalar angle = degrees * Scalar(3.14159265359) / Scalar(180);
    const Scalar c = std::cos(angle), s = std::sin(angle);
    constexpr size_t A = (Axis + 1) % 3, B = (Axis + 2) % 3;      // the plane that turns
    auto turn = [c, s](const bvh::Vector3<Scalar>& p) {
        bvh::Vector3<Scalar> q = p;
        // x: (y,z) -> (y c - z s, y s + z c);  y: (z,x) -> (z c - x s, z s + x c) written as the reference orders
        // its terms (x c + z s, -x s + z c);  z: (x,y) -> (x c - y s, x s + y c)
        if (Axis == 1) { q[0] = p[0] * c + p[2] * s; q[2] = -p[0] * s + p[2] * c; }
        else           { q[A] = p[A] * c - p[B] * s; q[B] = p[A] * s + p[B] * c; }
        return q;
    };
    #pragma omp parallel for
    for (size_t i = 0; i < triangle_count; ++i) {
        const auto a = turn(triangles[i].p0), b = turn(triangles[i].p1()), d = turn(triangles[i].p2());
        triangles[i] = bvh::Triangle<Scalar>(a, b, d);
    }
}

namespace ceres {

namespace detail {

struct SceneCache {
    ceres_scene* scene = nullptr;
    const void* nodes = nullptr; const void* tris = nullptr; const void* norms = nullptr;
    size_t node_count = 0, tri_count = 0;
    uint64_t fingerprint = 0;
    ~SceneCache() { if (scene) ceres_scene_destroy(scene); }
};
inline SceneCache& cache() { static SceneCache c; return c; }

// FNV-1a over up to 64 evenly spaced 32-byte windows: cheap enough to run every frame, catches in-place edits
inline uint64_t sample(const void* p, size_t bytes, uint64_t h) {
    const unsigned char* b = static_cast<const unsigned char*>(p);
    if (!b || !bytes) return h;
    const size_t windows = 64, span = 32;
    const size_t step = bytes > windows * span ? (bytes - span) / (windows - 1) : span;
    for (size_t off = 0; off + span <= bytes; off += step)
        for (size_t k = 0; k < span; ++k) { h ^= b[off + k]; h *= 1099511628211ull; }
    for (size_t k = bytes > span ? bytes - span : 0; k < bytes; ++k) { h ^= b[k]; h *= 1099511628211ull; }     // always the tail
    return h;
}

[[noreturn]] inline void fail(const char* what) {
    throw std::runtime_error(std::string(what) + ": " + ceres_last_error());
}

}  // namespace detail

namespace detail {
struct FramePin { void* p = nullptr; size_t bytes = 0; bool ours = false; };
inline FramePin& frame_pin() { static FramePin f; return f; }     // never unregistered at exit: the frame may be gone by then
}  // namespace detail

// Forget (and un-page-lock) the frame render() saw last; call before freeing a frame whose address may be reused.
inline void release_frame() {
    auto& f = detail::frame_pin();
    if (f.ours) ceres_host_unregister(f.p);
    f = detail::FramePin{};
}

// Page-lock the caller's frame while the same (pointer, size) keeps coming (see the header comment).
inline void autopin_frame(void* pixels, size_t bytes) {
    static const bool off = std::getenv("CERES_NO_AUTOPIN") != nullptr;
    auto& f = detail::frame_pin();
    if (f.p == pixels && f.bytes == bytes) return;
    release_frame();
    f.p = pixels; f.bytes = bytes;
    if (off || !pixels || !bytes || ceres_host_is_pinned(pixels)) return;
    f.ours = ceres_host_register(pixels, bytes) == CERES_OK;        // on failure: the pageable path still works
}

inline void invalidate_scene() {
    auto& c = detail::cache();
    if (c.scene) ceres_scene_destroy(c.scene);
    c = detail::SceneCache{};
}

// GPU used by render(); default device 0 (CERES_DEVICE in the environment overrides)
inline int device() {
    const char* e = std::getenv("CERES_DEVICE");
    return e ? std::atoi(e) : 0;
}

// the flattened, GPU-resident form of (bvh, triangles, tri_norms); created on first use
inline ceres_scene* scene_for(const bvh::Bvh<float>& bvh, const bvh::Triangle<float>* triangles,
                              const std::array<bvh::Vector3<float>, 3>* tri_norms) {
    static_assert(sizeof(bvh::Bvh<float>::Node) == 32, "lib/bvh/bvh.hpp:25-30");
    static_assert(sizeof(bvh::Triangle<float>) == 48, "lib/bvh/triangle.hpp:27");
    static_assert(sizeof(std::array<bvh::Vector3<float>, 3>) == 36, "lib/obj_norms.hpp:113-115");
    auto& c = detail::cache();
    // number of primitives = sum of leaf counts = what primitive_indices holds; the reference passes no count, so
    // derive it from the tree once per new tree
    const void* nodes = bvh.nodes.get();
    size_t tri_count = c.tri_count;
    const bool same_arrays = c.scene && c.nodes == nodes && c.tris == triangles && c.norms == tri_norms &&
                             c.node_count == bvh.node_count;
    if (!same_arrays) {
        tri_count = 0;
        for (size_t k = 0; k < bvh.node_count; ++k) tri_count += bvh.nodes[k].primitive_count;
    }
    uint64_t fp = 1469598103934665603ull;
    fp = detail::sample(nodes, bvh.node_count * 32, fp);
    fp = detail::sample(triangles, tri_count * 48, fp);
    fp = detail::sample(tri_norms, tri_norms ? tri_count * 36 : 0, fp);
    if (same_arrays && fp == c.fingerprint) return c.scene;
    invalidate_scene();
    ceres_scene* s = nullptr;
    if (ceres_scene_create(nodes, bvh.node_count, bvh.primitive_indices.get(), triangles, tri_norms, tri_count,
                           device(), &s) != CERES_OK)
        detail::fail("ceres_scene_create");
    c.scene = s; c.nodes = nodes; c.tris = triangles; c.norms = tri_norms;
    c.node_count = bvh.node_count; c.tri_count = tri_count; c.fingerprint = fp;
    return s;
}

}  // namespace ceres

namespace ceres {

namespace detail {
struct SceneCacheF64 {
    ceres_scene_f64* scene = nullptr;
    const void* nodes = nullptr; const void* tris = nullptr; const void* norms = nullptr;
    size_t node_count = 0, tri_count = 0;
    uint64_t fingerprint = 0;
    ~SceneCacheF64() { if (scene) ceres_scene_f64_destroy(scene); }
};
inline SceneCacheF64& cache_f64() { static SceneCacheF64 c; return c; }
}  // namespace detail

// the double-precision scene on the GPU; same caching rule as scene_for above
inline ceres_scene_f64* scene_for(const bvh::Bvh<double>& bvh, const bvh::Triangle<double>* triangles,
                                  const std::array<bvh::Vector3<double>, 3>* tri_norms) {
    static_assert(sizeof(bvh::Bvh<double>::Node) == 64, "lib/bvh/bvh.hpp:20-28");
    static_assert(sizeof(bvh::Triangle<double>) == 96, "lib/bvh/triangle.hpp:27");
    static_assert(sizeof(std::array<bvh::Vector3<double>, 3>) == 72, "lib/obj_norms.hpp:113-115");
    auto& c = detail::cache_f64();
    const void* nodes = bvh.nodes.get();
    size_t tri_count = c.tri_count;
    const bool same_arrays = c.scene && c.nodes == nodes && c.tris == triangles && c.norms == tri_norms && c.node_count == bvh.node_count;
    if (!same_arrays) {
        tri_count = 0;
        for (size_t k = 0; k < bvh.node_count; ++k) tri_count += bvh.nodes[k].primitive_count;
    }
    uint64_t fp = 1469598103934665603ull;
    fp = detail::sample(nodes, bvh.node_count * 64, fp);
    fp = detail::sample(triangles, tri_count * 96, fp);
    fp = detail::sample(tri_norms, tri_norms ? tri_count * 72 : 0, fp);
    if (same_arrays && fp == c.fingerprint) return c.scene;
    if (c.scene) ceres_scene_f64_destroy(c.scene);
    c = detail::SceneCacheF64{};
    ceres_scene_f64* s = nullptr;
    if (ceres_scene_f64_create(nodes, bvh.node_count, bvh.primitive_indices.get(), triangles, tri_norms, tri_count, device(), &s) != CERES_OK)
        detail::fail("ceres_scene_f64_create");
    c.scene = s; c.nodes = nodes; c.tris = triangles; c.norms = tri_norms;
    c.node_count = bvh.node_count; c.tri_count = tri_count; c.fingerprint = fp;
    return s;
}

}  // namespace ceres

template <typename Scalar>
std::pair<int, int> render(const Camera<Scalar>& camera, const bvh::Vector3<Scalar>& sun_position,
                           const bvh::Bvh<Scalar>& bvh, const bvh::Triangle<Scalar>* triangles,
                           std::array<bvh::Vector3<Scalar>, 3>* tri_norms, Scalar* pixels,
                           size_t width, size_t height) {
    static_assert(std::is_same<Scalar, float>::value || std::is_same<Scalar, double>::value, "Scalar is float or double");
    uint64_t counters[2] = {0, 0};
    if constexpr (std::is_same<Scalar, float>::value) {
        ceres_scene* scene = ceres::scene_for(bvh, triangles, tri_norms);
        ceres_camera cam;
        for (int k = 0; k < 3; ++k) { cam.eye[k] = camera.eye[k]; cam.dir[k] = camera.dir[k]; cam.up[k] = camera.up[k]; }
        cam.fov = camera.fov;
        const float sun[3] = {sun_position[0], sun_position[1], sun_position[2]};
        ceres::autopin_frame(pixels, 3 * width * height * sizeof(float));
        // CERES_FRAME_RETAINED=1 in the environment: the caller states that nobody writes into `pixels` between two render()
        // calls (true for anim.cpp:76-127, which only reads the frame) -- see CERES_FRAME_RETAINED in ceres_b200.h
        static const unsigned retained = std::getenv("CERES_FRAME_RETAINED") ? CERES_FRAME_RETAINED : 0u;
        if (ceres_render(scene, &cam, sun, width, height, retained, pixels, nullptr, counters) != CERES_OK)
            ceres::detail::fail("ceres_render");
    } else {
        ceres_scene_f64* scene = ceres::scene_for(bvh, triangles, tri_norms);
        ceres_camera_f64 cam;
        for (int k = 0; k < 3; ++k) { cam.eye[k] = camera.eye[k]; cam.dir[k] = camera.dir[k]; cam.up[k] = camera.up[k]; }
        cam.fov = camera.fov;
        const double sun[3] = {sun_position[0], sun_position[1], sun_position[2]};
        if (ceres_render_f64(scene, &cam, sun, width, height, 0u, pixels, nullptr, counters, nullptr) != CERES_OK)
            ceres::detail::fail("ceres_render_f64");
    }
    return std::pair<int, int>(static_cast<int>(counters[0]), static_cast<int>(counters[1]));   // render.hpp:155
}

// static.cpp:130 calls render(..., nullptr, ...): give that call something to bind to
template <typename Scalar>
std::pair<int, int> render(const Camera<Scalar>& camera, const bvh::Vector3<Scalar>& sun_position,
                           const bvh::Bvh<Scalar>& bvh, const bvh::Triangle<Scalar>* triangles, std::nullptr_t,
                           Scalar* pixels, size_t width, size_t height) {
    return render<Scalar>(camera, sun_position, bvh, triangles,
                          static_cast<std::array<bvh::Vector3<Scalar>, 3>*>(nullptr), pixels, width, height);
}

#endif  // __RENDER_H
