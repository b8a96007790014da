// apps/shim_bench.cpp -- times the drop-in render<float>() of include/ceres/render.hpp on a scene dump, into the three kinds
// of frame a caller can hand it:
//   make_unique   std::make_unique<float[]> exactly as anim.cpp:76 / static.cpp:117 (the shim page-locks it on first sight)
//   pageable      the same with CERES_NO_AUTOPIN semantics (ceres::release_frame() before every call is NOT used; the
//                 process is started with CERES_NO_AUTOPIN=1 for this mode)
//   pinned        a ceres_host_alloc'ed frame
// bench.py writes the dump from the same scene arrays it gives the GPU arm and reports the result as `e2e_shim`.
//
//   ./shim_bench scene.bin width height frames  eye[3] dir[3] up[3] fov sun[3]
// scene.bin: u64 node_count, u64 tri_count, nodes (32 B each), primitive indices (u64 each), triangles (48 B), tri_norms (36 B).
#include <array>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include <bvh/bvh.hpp>
#include <bvh/triangle.hpp>
#include <bvh/vector.hpp>

#include "render.hpp"

using Vector3 = bvh::Vector3<float>;
using Bvh = bvh::Bvh<float>;
using Triangle = bvh::Triangle<float>;

static bool read_all(std::FILE* f, void* p, size_t n) { return std::fread(p, 1, n, f) == n; }

int main(int argc, char** argv) {
    if (argc < 18) { std::fprintf(stderr, "usage: shim_bench scene.bin W H frames eye[3] dir[3] up[3] fov sun[3]\n"); return 2; }
    const size_t W = std::strtoull(argv[2], nullptr, 10), H = std::strtoull(argv[3], nullptr, 10);
    const int frames = std::atoi(argv[4]);
    float v[13];
    for (int k = 0; k < 13; ++k) v[k] = std::strtof(argv[5 + k], nullptr);
    std::FILE* f = std::fopen(argv[1], "rb");
    if (!f) { std::perror(argv[1]); return 1; }
    uint64_t node_count = 0, tri_count = 0;
    if (!read_all(f, &node_count, 8) || !read_all(f, &tri_count, 8)) return 1;
    static_assert(sizeof(Bvh::Node) == 32 && sizeof(Triangle) == 48 && sizeof(std::array<Vector3, 3>) == 36 && sizeof(size_t) == 8, "POD layouts");
    Bvh bvh;
    bvh.node_count = node_count;
    bvh.nodes = std::make_unique<Bvh::Node[]>(node_count);
    bvh.primitive_indices = std::make_unique<size_t[]>(tri_count);
    std::vector<Triangle> triangles(tri_count);
    std::vector<std::array<Vector3, 3>> tri_norms(tri_count);
    if (!read_all(f, bvh.nodes.get(), node_count * 32) || !read_all(f, bvh.primitive_indices.get(), tri_count * 8) ||
        !read_all(f, (void*)triangles.data(), tri_count * 48) || !read_all(f, (void*)tri_norms.data(), tri_count * 36)) {
        std::fprintf(stderr, "shim_bench: short read\n"); return 1;
    }
    std::fclose(f);
    Camera<float> camera = {Vector3(v[0], v[1], v[2]), Vector3(v[3], v[4], v[5]), Vector3(v[6], v[7], v[8]), v[9]};
    Vector3 sun(v[10], v[11], v[12]);
    const size_t n = 3 * W * H;
    auto time_into = [&](float* pixels, const char* name) {
        std::pair<int, int> r{0, 0};
        for (int k = 0; k < 3; ++k) r = render(camera, sun, bvh, triangles.data(), tri_norms.data(), pixels, W, H);
        double best = 1e30, sum = 0;
        for (int k = 0; k < frames; ++k) {
            const auto t0 = std::chrono::steady_clock::now();
            r = render(camera, sun, bvh, triangles.data(), tri_norms.data(), pixels, W, H);
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            best = ms < best ? ms : best; sum += ms;
        }
        double check = 0;
        for (size_t k = 0; k < n; k += 97) check += pixels[k];
        std::printf("{\"frame\": \"%s\", \"ms_mean\": %.6f, \"ms_best\": %.6f, \"rays\": %d, \"hits\": %d, \"checksum\": %.9e}\n",
                    name, sum / frames, best, r.first, r.second, check);
    };
    try {
        {
            auto pixels = std::make_unique<float[]>(n);                        // anim.cpp:76, static.cpp:117
            std::memset(pixels.get(), 0xFF, n * sizeof(float));
            time_into(pixels.get(), std::getenv("CERES_NO_AUTOPIN") ? "make_unique, CERES_NO_AUTOPIN=1 (pageable)" : "make_unique (page-locked by the shim on first sight)");
            ceres::release_frame();
        }
        {
            float* pixels = static_cast<float*>(ceres_host_alloc(n * sizeof(float)));
            if (!pixels) { std::fprintf(stderr, "ceres_host_alloc failed\n"); return 1; }
            time_into(pixels, "ceres_host_alloc");
            ceres::release_frame();
            ceres_host_free(pixels);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "shim_bench: %s\n", e.what());
        return 1;
    }
    return 0;
}
