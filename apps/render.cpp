// apps/render.cpp -- the `render` executable of ceres-raytracer (README.md:11, include/static.cpp), rendering on
// B200 GPUs through libceres_b200.so.
//
//   ./render                                             static.cpp's compiled-in scene: ../../data/dragon.obj, rotate x 90,
//                                                        eye (0,-15,2) dir (0,1,0) up (0,0,1) fov 60, sun (-50,-20,0),
//                                                        1920x1080, BinnedSah<16> -> render.ppm   (static.cpp:26,38-47,72-73,106)
//   ./render ../data/bunny.obj --eye 0 .1 -.3 --rotate y -145     the README example (OBJ preview: dir (0,0,1), up (0,1,0))
//
// Written fresh: the upstream static.cpp ignores argv (static.cpp:23-24) and does not compile, so the
// README command line is implemented here.  Host-side scene preparation is the caller's own lib/bvh + obj_norms.hpp
// (the reference tree on the include path, as in a drop-in build); the render() call goes through
// include/ceres/render.hpp -> C ABI -> CUDA.  Output: P6 PPM with static.cpp:135-147 semantics, and the stdout lines
// of static.cpp:96-133.
//
// Additive flags: --dir x y z  --up x y z  --fov deg  --sun x y z  --width W  --height H  -o file.ppm
//   --no-shadow (primary rays only, BASELINE configs 1-2)  --builder binned|sweep|device (device: the tree is built on
//   the GPU, ceres_build_bvh -- top-down binned SAH in milliseconds, into the same Bvh structure)  --gpus N (interleaved 16-row blocks,
//   one host thread per GPU)  --dump-ids file.u32  --dump-rgb file.f32  --frames K (repeat the render call K times)
//   --ppm8 (quantise + flip on the GPU and copy the 8-bit image: static.cpp:135-147 done on the device, 1 GPU)
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include <bvh/bvh.hpp>
#include <bvh/binned_sah_builder.hpp>
#include <bvh/sweep_sah_builder.hpp>
#include <bvh/node_layout_optimizer.hpp>
#include <bvh/triangle.hpp>

#include "render.hpp"        // include/ceres/render.hpp (found first on the include path)
#include "obj_norms.hpp"
#include "ppm.hpp"

using Scalar = float;
using Triangle = bvh::Triangle<Scalar>;
using Vector3 = bvh::Vector3<Scalar>;
using Bvh = bvh::Bvh<Scalar>;

namespace {

struct Options {
    std::string input = "../../data/dragon.obj", output = "render.ppm", dump_ids, dump_rgb, builder = "binned";
    bool ppm8 = false;
    Camera<Scalar> camera{Vector3(0.0, -15.0, 2.0), Vector3(0, 1, 0), Vector3(0, 0, 1), 60};
    Vector3 sun = Vector3(-50.0, -20.0, 0.0);
    int rotation_axis = 0;
    Scalar rotation_degrees = 90;
    size_t width = 1920, height = 1080;
    bool shadows = true;
    int gpus = 1, frames = 1;
};

[[noreturn]] void usage(const char* msg) {
    std::cerr << "render: " << msg << "\n"
              << "usage: render [file.obj] [--eye x y z] [--rotate x|y|z deg] [--dir x y z] [--up x y z] [--fov deg]\n"
              << "              [--sun x y z] [--width W] [--height H] [-o out.ppm] [--no-shadow] [--builder binned|sweep|device]\n"
              << "              [--gpus N] [--dump-ids f.u32] [--dump-rgb f.f32] [--frames K] [--ppm8]\n";
    std::exit(2);
}

Options parse(int argc, char** argv) {
    Options o;
    bool have_input = false, have_dir = false, have_up = false, have_rot = false;
    auto need = [&](int i, int n) { if (i + n >= argc) usage((std::string("missing value after ") + argv[i]).c_str()); };
    auto vec = [&](int& i) { need(i, 3); Vector3 v(std::strtof(argv[i + 1], nullptr), std::strtof(argv[i + 2], nullptr), std::strtof(argv[i + 3], nullptr)); i += 3; return v; };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--eye") o.camera.eye = vec(i);
        else if (a == "--dir") { o.camera.dir = vec(i); have_dir = true; }
        else if (a == "--up") { o.camera.up = vec(i); have_up = true; }
        else if (a == "--sun") o.sun = vec(i);
        else if (a == "--fov") { need(i, 1); o.camera.fov = std::strtof(argv[++i], nullptr); }
        else if (a == "--rotate") {
            need(i, 2);
            const std::string ax = argv[++i];
            if (ax == "x" || ax == "0") o.rotation_axis = 0; else if (ax == "y" || ax == "1") o.rotation_axis = 1;
            else if (ax == "z" || ax == "2") o.rotation_axis = 2; else usage("--rotate axis must be x, y or z");
            o.rotation_degrees = std::strtof(argv[++i], nullptr);
            have_rot = true;
        }
        else if (a == "--width") { need(i, 1); o.width = std::strtoull(argv[++i], nullptr, 10); }
        else if (a == "--height") { need(i, 1); o.height = std::strtoull(argv[++i], nullptr, 10); }
        else if (a == "-o") { need(i, 1); o.output = argv[++i]; }
        else if (a == "--no-shadow") o.shadows = false;
        else if (a == "--builder") { need(i, 1); o.builder = argv[++i]; }
        else if (a == "--gpus") { need(i, 1); o.gpus = std::atoi(argv[++i]); }
        else if (a == "--frames") { need(i, 1); o.frames = std::atoi(argv[++i]); }
        else if (a == "--ppm8") o.ppm8 = true;
        else if (a == "--dump-ids") { need(i, 1); o.dump_ids = argv[++i]; }
        else if (a == "--dump-rgb") { need(i, 1); o.dump_rgb = argv[++i]; }
        else if (!a.empty() && a[0] != '-' && !have_input) { o.input = a; have_input = true; }
        else usage(("unknown argument " + a).c_str());
    }
    if (have_input) {
        // README preview mode (SURVEY.md 8(b) decision D1): OBJ-native y-up frame unless told otherwise
        if (!have_dir) o.camera.dir = Vector3(0, 0, 1);
        if (!have_up) o.camera.up = Vector3(0, 1, 0);
        if (!have_rot) { o.rotation_axis = 0; o.rotation_degrees = 0; }
    }
    if (o.builder != "binned" && o.builder != "sweep" && o.builder != "device") usage("--builder is binned, sweep or device");
    if (o.gpus < 1 || o.frames < 1 || !o.width || !o.height) usage("bad --gpus/--frames/--width/--height");
    return o;
}

double seconds_since(std::chrono::high_resolution_clock::time_point t0) {
    return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::high_resolution_clock::now() - t0).count() / 1e6;
}

}  // namespace

int main(int argc, char** argv) {
    const Options opt = parse(argc, argv);

    auto loaded = obj::load_from_file<Scalar>(opt.input);     // obj_norms.hpp:120-127
    auto& triangles = loaded.first;
    auto& tri_norms = loaded.second;
    if (triangles.size() == 0) {
        std::cerr << "The given scene is empty or cannot be loaded" << std::endl;   // static.cpp:77-80
        return 1;
    }
    if (opt.rotation_degrees != 0) {
        if (opt.rotation_axis == 0) rotate_triangles<0>(opt.rotation_degrees, triangles.data(), triangles.size());
        else if (opt.rotation_axis == 1) rotate_triangles<1>(opt.rotation_degrees, triangles.data(), triangles.size());
        else rotate_triangles<2>(opt.rotation_degrees, triangles.data(), triangles.size());
    }

    Bvh bvh;
    const size_t reference_count = triangles.size();
    std::cout << "Building BVH ( using " << (opt.builder == "binned" ? "BinnedSahBuilder" : opt.builder == "device" ? "the B200 binned-SAH builder" : "SweepSahBuilder")
              << " )..." << std::endl;
    auto t0 = std::chrono::high_resolution_clock::now();
    if (opt.builder == "device") {
        // same data structure, filled by the library: bvh.nodes (32-byte nodes, root 0, siblings adjacent) and
        // bvh.primitive_indices (bvh.hpp:25-30,93-96)
        static_assert(sizeof(Bvh::Node) == 32 && sizeof(Triangle) == 48, "reference POD layouts");
        bvh.nodes = std::make_unique<Bvh::Node[]>(2 * reference_count - 1);
        bvh.primitive_indices = std::make_unique<size_t[]>(reference_count);
        size_t node_count = 0; double build_ms = 0;
        if (ceres_build_bvh(triangles.data(), reference_count, 0, bvh.nodes.get(), &node_count, bvh.primitive_indices.get(), &build_ms) != CERES_OK) {
            std::cerr << "render: " << ceres_last_error() << std::endl;
            return 1;
        }
        bvh.node_count = node_count;
        std::cout << "(device build kernels: " << build_ms << " ms)" << std::endl;
    } else {
        auto bc = bvh::compute_bounding_boxes_and_centers(triangles.data(), triangles.size());
        auto global_bbox = bvh::compute_bounding_boxes_union(bc.first.get(), triangles.size());
        if (opt.builder == "binned") {                                   // static.cpp:106-107
            bvh::BinnedSahBuilder<Bvh, 16> builder(bvh);
            builder.build(global_bbox, bc.first.get(), bc.second.get(), reference_count);
        } else {                                                         // anim.cpp:62-65
            bvh::SweepSahBuilder<Bvh> builder(bvh);
            builder.build(global_bbox, bc.first.get(), bc.second.get(), reference_count);
            bvh::NodeLayoutOptimizer<Bvh> layout(bvh);
            layout.optimize();
        }
    }
    std::cout << seconds_since(t0) << std::endl;
    std::cout << "BVH of " << bvh.node_count << " node(s) and " << reference_count << " reference(s)" << std::endl;

    const size_t npix = opt.width * opt.height;
    // page-locked framebuffer: the device writes it by DMA (any other host pointer also works, slower)
    Scalar* pixels = static_cast<Scalar*>(ceres_host_alloc(3 * npix * sizeof(Scalar)));
    if (!pixels) { std::cerr << "render: " << ceres_last_error() << std::endl; return 1; }
    std::vector<uint32_t> ids(opt.dump_ids.empty() ? 0 : npix);

    const int ndev = ceres_device_count();
    if (ndev < 1) { std::cerr << "render: no CUDA device (this build has no CPU path)" << std::endl; return 1; }
    if (opt.gpus > ndev) { std::cerr << "render: --gpus " << opt.gpus << " but only " << ndev << " visible" << std::endl; return 1; }
    std::cout << "Rendering image (" << opt.width << "x" << opt.height << ") on " << opt.gpus << " B200 GPU(s)..." << std::endl;

    std::pair<int, int> counters(0, 0);
    double render_s = 0;
    std::vector<uint8_t> rgb8;            // --ppm8: the P6 payload comes from the device
    try {
        if (opt.ppm8) {
            if (opt.gpus != 1 || !ids.empty() || !opt.dump_rgb.empty()) throw std::runtime_error("--ppm8 renders on one GPU and produces only the PPM");
            ceres_scene* scene = nullptr;
            if (ceres_scene_create(bvh.nodes.get(), bvh.node_count, bvh.primitive_indices.get(), triangles.data(),
                                   tri_norms.data(), reference_count, 0, &scene) != CERES_OK) throw std::runtime_error(ceres_last_error());
            uint8_t* host8 = static_cast<uint8_t*>(ceres_host_alloc(3 * npix));
            if (!host8) { ceres_scene_destroy(scene); throw std::runtime_error(ceres_last_error()); }
            ceres_camera cam;
            for (int k = 0; k < 3; ++k) { cam.eye[k] = opt.camera.eye[k]; cam.dir[k] = opt.camera.dir[k]; cam.up[k] = opt.camera.up[k]; }
            cam.fov = opt.camera.fov;
            const float sun[3] = {opt.sun[0], opt.sun[1], opt.sun[2]};
            uint64_t cnt[2] = {0, 0};
            int rc = CERES_OK;
            for (int f = -1; f < opt.frames && rc == CERES_OK; ++f) {      // f = -1: warm-up
                auto t1 = std::chrono::high_resolution_clock::now();
                rc = ceres_render_ppm8(scene, &cam, sun, opt.width, opt.height, opt.shadows ? 0u : CERES_NO_SHADOWS, host8, cnt);
                if (f >= 0) render_s += seconds_since(t1) / opt.frames;
            }
            if (rc == CERES_OK) rgb8.assign(host8, host8 + 3 * npix);
            ceres_host_free(host8);
            ceres_scene_destroy(scene);
            if (rc != CERES_OK) throw std::runtime_error(ceres_last_error());
            counters = std::pair<int, int>((int)cnt[0], (int)cnt[1]);
        } else if (opt.gpus == 1 && opt.shadows && ids.empty()) {
            // the reference call, unchanged (static.cpp:130 / anim.cpp:105)
            render(opt.camera, opt.sun, bvh, triangles.data(), tri_norms.data(), pixels, opt.width, opt.height);   // upload + warm-up
            auto t1 = std::chrono::high_resolution_clock::now();
            for (int f = 0; f < opt.frames; ++f)
                counters = render(opt.camera, opt.sun, bvh, triangles.data(), tri_norms.data(), pixels, opt.width, opt.height);
            render_s = seconds_since(t1) / opt.frames;
        } else {
            // flags the reference signature has no room for: straight to the C ABI
            std::vector<ceres_scene*> scenes(opt.gpus, nullptr);
            for (int g = 0; g < opt.gpus; ++g)
                if (ceres_scene_create(bvh.nodes.get(), bvh.node_count, bvh.primitive_indices.get(), triangles.data(),
                                       tri_norms.data(), reference_count, g, &scenes[g]) != CERES_OK)
                    throw std::runtime_error(ceres_last_error());
            ceres_camera cam;
            for (int k = 0; k < 3; ++k) { cam.eye[k] = opt.camera.eye[k]; cam.dir[k] = opt.camera.dir[k]; cam.up[k] = opt.camera.up[k]; }
            cam.fov = opt.camera.fov;
            const float sun[3] = {opt.sun[0], opt.sun[1], opt.sun[2]};
            const unsigned flags = opt.shadows ? 0u : CERES_NO_SHADOWS;
            uint64_t cnt[2] = {0, 0};
            for (int f = -1; f < opt.frames; ++f) {      // f = -1: warm-up
                auto t1 = std::chrono::high_resolution_clock::now();
                if (ceres_render_multi(scenes.data(), opt.gpus, &cam, sun, opt.width, opt.height, flags, pixels,
                                       ids.empty() ? nullptr : ids.data(), cnt) != CERES_OK)
                    throw std::runtime_error(ceres_last_error());
                if (f >= 0) render_s += seconds_since(t1) / opt.frames;
            }
            counters = std::pair<int, int>((int)cnt[0], (int)cnt[1]);
            for (ceres_scene* s : scenes) ceres_scene_destroy(s);
        }
    } catch (const std::exception& e) {
        std::cerr << "render: " << e.what() << std::endl;
        return 1;
    }
    std::cout << render_s << std::endl;
    std::cout << "Rays: " << counters.first << "\tHits: " << counters.second << "\t" << counters.first / render_s / 1e6
              << " Mrays/s" << std::endl;

    if (opt.ppm8 ? !ceres_apps::write_ppm_bytes(opt.output, rgb8.data(), opt.width, opt.height)
                 : !ceres_apps::write_ppm(opt.output, pixels, opt.width, opt.height)) {
        std::cerr << "render: cannot write " << opt.output << std::endl;
        return 1;
    }
    if (!opt.dump_ids.empty()) std::ofstream(opt.dump_ids, std::ios::binary).write(reinterpret_cast<const char*>(ids.data()), ids.size() * 4);
    if (!opt.dump_rgb.empty()) std::ofstream(opt.dump_rgb, std::ios::binary).write(reinterpret_cast<const char*>(pixels), 3 * npix * sizeof(Scalar));
    ceres::invalidate_scene();
    ceres_host_free(pixels);
    return 0;
}
