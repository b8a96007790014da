// apps/shim_check.cpp -- calls render<float>() and render<double>() through include/ceres/render.hpp exactly as a caller
// of the reference's render.hpp would (same headers, same call, anim.cpp:46-65,105), twice each (the second call re-uses
// the cached GPU scene), and prints the counters and a checksum of the frame.  tests/test_apps.py compares the lines with
// the reference's own float and double builds.
//
//   ./shim_check file.obj [width height]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <vector>

#include <bvh/bvh.hpp>
#include <bvh/sweep_sah_builder.hpp>
#include <bvh/node_layout_optimizer.hpp>
#include <bvh/triangle.hpp>

#include "render.hpp"
#include "obj_norms.hpp"

template <typename Scalar>
int run(const char* name, const char* path, size_t width, size_t height) {
    using Vector3 = bvh::Vector3<Scalar>;
    using Bvh = bvh::Bvh<Scalar>;
    auto loaded = obj::load_from_file<Scalar>(path);
    auto& triangles = loaded.first;
    auto& tri_norms = loaded.second;
    if (triangles.empty()) { std::cerr << "The given scene is empty or cannot be loaded" << std::endl; return 1; }
    Bvh bvh;
    {
        auto bc = bvh::compute_bounding_boxes_and_centers(triangles.data(), triangles.size());
        auto global_bbox = bvh::compute_bounding_boxes_union(bc.first.get(), triangles.size());
        bvh::SweepSahBuilder<Bvh> builder(bvh);
        builder.build(global_bbox, bc.first.get(), bc.second.get(), triangles.size());
        bvh::NodeLayoutOptimizer<Bvh> layout(bvh);
        layout.optimize();
    }
    Camera<Scalar> camera = {Vector3(-150.0, -20.0, 0.0), Vector3(1, -0.1, 0), Vector3(0, -1, 0), 60};
    Vector3 sun_position(-500.0, 300.0, 10.0);
    auto pixels = std::make_unique<Scalar[]>(3 * width * height);
    for (int pass = 0; pass < 2; ++pass) {
        for (size_t k = 0; k < 3 * width * height; ++k) pixels[k] = Scalar(-1);      // render() must overwrite every pixel
        auto r = render(camera, sun_position, bvh, triangles.data(), tri_norms.data(), pixels.get(), width, height);
        double sum = 0; size_t lit = 0, negative = 0;
        for (size_t k = 0; k < 3 * width * height; ++k) { sum += (double)pixels[k]; lit += pixels[k] > 0; negative += pixels[k] < 0; }
        std::printf("%s pass %d: rays %d hits %d lit %zu negative %zu sum %.9e\n", name, pass, r.first, r.second, lit, negative, sum);
        camera.eye = camera.eye + Vector3(5, 0, 0);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) { std::cerr << "usage: shim_check file.obj [width height]" << std::endl; return 2; }
    const size_t width = argc > 3 ? std::strtoull(argv[2], nullptr, 10) : 200, height = argc > 3 ? std::strtoull(argv[3], nullptr, 10) : 320;
    try {
        if (int rc = run<float>("float", argv[1], width, height)) return rc;
        return run<double>("double", argv[1], width, height);
    } catch (const std::exception& e) {
        std::cerr << "shim_check: " << e.what() << std::endl;
        return 1;
    }
}
