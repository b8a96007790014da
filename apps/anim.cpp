// apps/anim.cpp -- the `anim` executable of ceres-raytracer (include/anim.cpp): an n-frame turntable in which the
// camera and the sun orbit a static object, rendered on B200 GPUs through libceres_b200.so.
//
//   ./anim [-d] [out_file]      as upstream (anim.cpp:145-152).  Defaults are upstream's constants: 621x1344, 60 frames,
//                               ../../waahh_smoothed.obj, eye (-150,-20,0) dir (1,-0.1,0) up (0,-1,0) fov 60,
//                               sun (-500,300,10), 6 degrees per frame about +y, SweepSah + NodeLayoutOptimizer
//                               (anim.cpp:27-35,62-65,77-89,140-143).
//
// Written fresh.  What differs from upstream, and why:
//   * -d (Scalar = double, anim.cpp:154-155) loads, builds and renders in double through ceres_render_f64 (one kernel per
//     frame, no host-frame pipeline); the colours of render<double>() are float-valued, so the writers take them narrowed.
//   * upstream encodes the frames with ImageMagick 7 Magick++ (anim.cpp:112-127) -- unavailable on this box.  Frames
//     are written as <out_file stem>_NNNN.ppm instead (16-bit P6 with --ppm16, matching the x65535 quantum scaling of
//     anim.cpp:117-119; 8-bit static.cpp-style by default), quantised on the GPU, image row = frame row as in anim.cpp:117-119.
//   * frames are independent once the camera path is known; the path is iterative (each frame transforms the previous
//     camera, anim.cpp:87-89), so it is computed serially on the host first and the frames are then dealt round-robin
//     to the GPUs (--gpus N), one host thread per GPU, scene replicated.
// Additive flags: --input file.obj  --frames n  --width W  --height H  --gpus N  --ppm16  --no-output
//   --scene dragon  (../../data/dragon.obj rotated x 90, static.cpp camera/sun, orbit about +z)
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <future>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include <bvh/bvh.hpp>
#include <bvh/sweep_sah_builder.hpp>
#include <bvh/node_layout_optimizer.hpp>
#include <bvh/triangle.hpp>

#include "render.hpp"        // include/ceres/render.hpp; pulls in the caller tree's transform.hpp
#include "obj_norms.hpp"
#include "ppm.hpp"

namespace {

template <typename Scalar> struct FramePose { Camera<Scalar> camera; bvh::Vector3<Scalar> sun; };

struct Options {
    std::string out_file = "render.mp4", input_file = "../../waahh_smoothed.obj";
    size_t width = 1242 / 2, height = 2688 / 2;
    int n_frames = 60, gpus = 1;
    bool ppm16 = false, no_output = false, dragon = false;
};

double seconds_since(std::chrono::high_resolution_clock::time_point t0) {
    return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::high_resolution_clock::now() - t0).count() / 1e6;
}

std::string frame_name(const std::string& out_file, int i) {
    std::string stem = out_file;
    const size_t dot = stem.find_last_of('.');
    if (dot != std::string::npos && dot > stem.find_last_of('/') + 0) stem = stem.substr(0, dot);
    char buf[32];
    std::snprintf(buf, sizeof buf, "_%04d.ppm", i);
    return stem + buf;
}

}  // namespace

template <typename Scalar> int scene(const Options& opt);

int main(int argc, char** argv) {
    Options o;
    bool use_double = false;
    std::vector<std::string> positional;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::cerr << "anim: missing value after " << a << std::endl; std::exit(2); } return argv[++i]; };
        if (a == "-d") use_double = true;                        // anim.cpp:146-148
        else if (a == "--input") o.input_file = val();
        else if (a == "--frames") o.n_frames = std::atoi(val());
        else if (a == "--width") o.width = std::strtoull(val(), nullptr, 10);
        else if (a == "--height") o.height = std::strtoull(val(), nullptr, 10);
        else if (a == "--gpus") o.gpus = std::atoi(val());
        else if (a == "--ppm16") o.ppm16 = true;
        else if (a == "--no-output") o.no_output = true;
        else if (a == "--scene") { o.dragon = std::string(val()) == "dragon"; }
        else positional.push_back(a);
    }
    if (!positional.empty()) o.out_file = positional.back();     // anim.cpp:145-152
    if (o.n_frames < 1 || o.gpus < 1 || !o.width || !o.height) { std::cerr << "anim: bad --frames/--gpus/--width/--height" << std::endl; return 2; }
    return use_double ? scene<double>(o) : scene<float>(o);      // anim.cpp:154-155
}

template <typename Scalar>
int scene(const Options& opt) {
    using Vector3 = bvh::Vector3<Scalar>;
    using Bvh = bvh::Bvh<Scalar>;
    const std::string out_file = opt.out_file;
    std::string input_file = opt.input_file;
    const size_t width = opt.width, height = opt.height;
    const int n_frames = opt.n_frames, gpus = opt.gpus;
    const bool ppm16 = opt.ppm16, no_output = opt.no_output, dragon = opt.dragon;

    constexpr Scalar pi = Scalar(3.14159265359);
    Camera<Scalar> camera = {Vector3(-150.0, -20.0, 0.0), Vector3(1, -0.1, 0), Vector3(0, -1, 0), 60};
    Vector3 sun_position = Vector3(-500.0, 300.0, 10.0);
    Vector3 axis(0, 1, 0);
    if (dragon) {
        if (input_file == "../../waahh_smoothed.obj") input_file = "../../data/dragon.obj";
        camera = {Vector3(0.0, -15.0, 2.0), Vector3(0, 1, 0), Vector3(0, 0, 1), 60};
        sun_position = Vector3(-50.0, -20.0, 0.0);
        axis = Vector3(0, 0, 1);
    }
    const Scalar rotation_step_degrees = 360.0f / n_frames;

    auto loaded = obj::load_from_file<Scalar>(input_file);
    auto& triangles = loaded.first;
    auto& tri_norms = loaded.second;
    if (triangles.size() == 0) {
        std::cerr << "The given scene is empty or cannot be loaded" << std::endl;
        return 1;
    }
    if (dragon) rotate_triangles<0>(Scalar(90), triangles.data(), triangles.size());

    Bvh bvh;
    const size_t reference_count = triangles.size();
    std::cout << "Building BVH ( using SweepSahBuilder )..." << std::endl;
    auto start = std::chrono::high_resolution_clock::now();
    {
        auto bc = bvh::compute_bounding_boxes_and_centers(triangles.data(), triangles.size());
        auto global_bbox = bvh::compute_bounding_boxes_union(bc.first.get(), triangles.size());
        bvh::SweepSahBuilder<Bvh> builder(bvh);
        builder.build(global_bbox, bc.first.get(), bc.second.get(), reference_count);
        bvh::NodeLayoutOptimizer<Bvh> layout(bvh);
        layout.optimize();
    }
    std::cout << "BVH took " << seconds_since(start) << " s" << std::endl;
    std::cout << "BVH of " << bvh.node_count << " node(s) and " << reference_count << " reference(s)" << std::endl;

    // the camera path, serially (each pose is the previous one rotated, anim.cpp:77-89)
    const auto t_cam = Transform<Scalar>().rotate(axis, rotation_step_degrees / 180.0f * pi);
    const auto t_sun = Transform<Scalar>().rotate(axis, rotation_step_degrees / 180.0f * pi);
    std::vector<FramePose<Scalar>> path(n_frames);
    for (int i = 0; i < n_frames; ++i) {
        camera.eye = t_cam(camera.eye);
        camera.dir = t_cam(camera.dir);
        sun_position = t_sun(sun_position);
        path[i] = FramePose<Scalar>{camera, sun_position};
    }

    const int ndev = ceres_device_count();
    if (ndev < 1) { std::cerr << "anim: no CUDA device (this build has no CPU path)" << std::endl; return 1; }
    if (gpus > ndev) { std::cerr << "anim: --gpus " << gpus << " but only " << ndev << " visible" << std::endl; return 1; }

    std::vector<std::string> log(n_frames);
    std::vector<long> frame_rays(n_frames, 0);
    std::vector<std::string> errors(gpus);
    auto wall = std::chrono::high_resolution_clock::now();
    auto worker_f32 = [&](int g) {
      if constexpr (std::is_same<Scalar, float>::value) {      // the other instantiation never takes this branch
        ceres_scene* scene = nullptr;
        if (ceres_scene_create(bvh.nodes.get(), bvh.node_count, bvh.primitive_indices.get(), triangles.data(),
                               tri_norms.data(), reference_count, g, &scene) != CERES_OK) { errors[g] = ceres_last_error(); return; }
        // The frames leave the GPU as the file's payload: quantised on the device (ceres_render_ppm8 /
        // ceres_render_ppm16), so a quarter / half of the float frame's bytes cross PCIe.  Two page-locked payload
        // buffers per GPU: frame k is written by a writer thread while frame k+1 renders (SURVEY.md 8(f) N1 + N4;
        // upstream encodes synchronously inside the frame loop, anim.cpp:112-127).
        const int bpc = ppm16 ? 2 : 1;
        const size_t payload = 3 * width * height * (size_t)bpc;
        uint8_t* buffers[2] = {static_cast<uint8_t*>(ceres_host_alloc(payload)), static_cast<uint8_t*>(ceres_host_alloc(payload))};
        if (!buffers[0] || !buffers[1]) {
            errors[g] = ceres_last_error(); ceres_host_free(buffers[0]); ceres_host_free(buffers[1]); ceres_scene_destroy(scene); return;
        }
        std::future<std::string> writing[2];
        auto drain = [&](int b) {            // wait for the write that last used buffer b; returns false on a write error
            if (!writing[b].valid()) return true;
            const std::string err = writing[b].get();
            if (!err.empty() && errors[g].empty()) errors[g] = err;
            return err.empty();
        };
        int turn = 0;
        for (int i = g; i < n_frames; i += gpus, turn ^= 1) {
            if (!drain(turn)) break;
            uint8_t* bytes = buffers[turn];
            ceres_camera cam;
            for (int k = 0; k < 3; ++k) { cam.eye[k] = (float)path[i].camera.eye[k]; cam.dir[k] = (float)path[i].camera.dir[k]; cam.up[k] = (float)path[i].camera.up[k]; }
            cam.fov = (float)path[i].camera.fov;
            const float sun[3] = {(float)path[i].sun[0], (float)path[i].sun[1], (float)path[i].sun[2]};
            uint64_t cnt[2] = {0, 0};
            auto t0 = std::chrono::high_resolution_clock::now();
            // image row = frame row, as anim.cpp:117-119 fills its Magick image (static.cpp flips, anim.cpp does not)
            const int rc = ppm16 ? ceres_render_ppm16(scene, &cam, sun, width, height, CERES_PAYLOAD_NO_FLIP, bytes, cnt)
                                 : ceres_render_ppm8(scene, &cam, sun, width, height, CERES_PAYLOAD_NO_FLIP, bytes, cnt);
            if (rc != CERES_OK) { errors[g] = ceres_last_error(); break; }
            const double s = seconds_since(t0);
            std::ostringstream os;
            os << "Rendering image " << i << " (" << width << "x" << height << ") on GPU " << g << "...\n"
               << "Rays: " << (int)cnt[0] << "\tHits: " << (int)cnt[1] << "\n" << "Tracing took " << s << " s\n";
            log[i] = os.str();
            frame_rays[i] = (long)(int)cnt[0];
            if (!no_output) {
                const std::string name = frame_name(out_file, i);
                writing[turn] = std::async(std::launch::async, [=]() -> std::string {
                    return ceres_apps::write_ppm_bytes(name, bytes, width, height, bpc) ? std::string() : "cannot write " + name;
                });
            }
        }
        drain(0); drain(1);
        ceres_host_free(buffers[0]); ceres_host_free(buffers[1]);
        ceres_scene_destroy(scene);
      }
    };
    // Scalar = double: ceres_render_f64 fills a double frame; its colours are float-valued (render.hpp:56-84 accumulates
    // in float), so they are narrowed without loss for the same writers.
    auto worker_f64 = [&](int g) {
      if constexpr (std::is_same<Scalar, double>::value) {
        ceres_scene_f64* scene = nullptr;
        if (ceres_scene_f64_create(bvh.nodes.get(), bvh.node_count, bvh.primitive_indices.get(), triangles.data(),
                                   tri_norms.data(), reference_count, g, &scene) != CERES_OK) { errors[g] = ceres_last_error(); return; }
        std::vector<double> frame(3 * width * height);
        std::vector<float> narrow[2] = {std::vector<float>(3 * width * height), std::vector<float>(3 * width * height)};
        std::future<std::string> writing[2];
        auto drain = [&](int b) {
            if (!writing[b].valid()) return true;
            const std::string err = writing[b].get();
            if (!err.empty() && errors[g].empty()) errors[g] = err;
            return err.empty();
        };
        int turn = 0;
        for (int i = g; i < n_frames; i += gpus, turn ^= 1) {
            if (!drain(turn)) break;
            ceres_camera_f64 cam;
            for (int k = 0; k < 3; ++k) { cam.eye[k] = path[i].camera.eye[k]; cam.dir[k] = path[i].camera.dir[k]; cam.up[k] = path[i].camera.up[k]; }
            cam.fov = path[i].camera.fov;
            const double sun[3] = {(double)path[i].sun[0], (double)path[i].sun[1], (double)path[i].sun[2]};
            uint64_t cnt[2] = {0, 0};
            auto t0 = std::chrono::high_resolution_clock::now();
            if (ceres_render_f64(scene, &cam, sun, width, height, 0u, frame.data(), nullptr, cnt, nullptr) != CERES_OK) { errors[g] = ceres_last_error(); break; }
            const double s = seconds_since(t0);
            std::ostringstream os;
            os << "Rendering image " << i << " (" << width << "x" << height << ") on GPU " << g << "...\n"
               << "Rays: " << (int)cnt[0] << "\tHits: " << (int)cnt[1] << "\n" << "Tracing took " << s << " s\n";
            log[i] = os.str();
            frame_rays[i] = (long)(int)cnt[0];
            if (!no_output) {
                float* pixels = narrow[turn].data();
                for (size_t k = 0; k < frame.size(); ++k) pixels[k] = (float)frame[k];
                const std::string name = frame_name(out_file, i);
                writing[turn] = std::async(std::launch::async, [=]() -> std::string {
                    const bool ok = ppm16 ? ceres_apps::write_ppm16(name, pixels, width, height, false) : ceres_apps::write_ppm(name, pixels, width, height, false);
                    return ok ? std::string() : "cannot write " + name;
                });
            }
        }
        drain(0); drain(1);
        ceres_scene_f64_destroy(scene);
      }
    };
    auto worker = [&](int g) {
        if constexpr (std::is_same<Scalar, float>::value) worker_f32(g); else worker_f64(g);
    };
    std::vector<std::thread> threads;
    for (int g = 1; g < gpus; ++g) threads.emplace_back(worker, g);
    worker(0);
    for (auto& t : threads) t.join();
    const double total_s = seconds_since(wall);
    for (int g = 0; g < gpus; ++g) if (!errors[g].empty()) { std::cerr << "anim: GPU " << g << ": " << errors[g] << std::endl; return 1; }

    long tot_rays = 0;
    for (int i = 0; i < n_frames; ++i) { std::cout << log[i]; tot_rays += frame_rays[i]; }
    std::cout << "Total Rays: " << tot_rays << std::endl;
    std::cout << n_frames << " frames in " << total_s << " s on " << gpus << " GPU(s): " << tot_rays / total_s / 1e6 << " Mrays/s incl. output" << std::endl;
    return 0;
}
