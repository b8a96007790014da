// profiles/sanitize_driver.cpp -- a reduced workload that launches EVERY kernel the product library ships:
// profiles/run_sanitizers.sh builds it and runs it against the product library and against the CHECKED build
// (-DCERES_CHECKED: every data-dependent index compared with its bound, queue protocol invariants verified between frames).
// compute-sanitizer is closed on the GPU pool, so the checked build + the repeat-and-compare loops below are this
// repository's memcheck / racecheck.  Plain C ABI calls + cudart for the device buffers; no Python, no torch.
//
//   sanitize_driver scene.bin scene64.bin      (dumps written by profiles/sanitize_dump.py from tests/golden/*.npz)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ceres_b200.h"

#define CHECK(x) do { int rc_ = (x); if (rc_ != CERES_OK) { std::fprintf(stderr, "FAILED %s: %s\n", #x, ceres_last_error()); return 1; } } while (0)
#define NOERR(scene, what) do { uint64_t w_ = 0; CHECK(ceres_scene_frame_error(scene, &w_)); if (w_) { std::fprintf(stderr, "DEVICE CHECK FAILED after %s: bits 0x%llx\n", what, (unsigned long long)w_); ++g_bad; } } while (0)
static int g_bad = 0;
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "FAILED %s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

static bool slurp(const char* path, std::vector<unsigned char>& out) {
    std::FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END); long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
    out.resize((size_t)n);
    bool ok = std::fread(out.data(), 1, (size_t)n, f) == (size_t)n;
    std::fclose(f);
    return ok;
}

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: sanitize_driver scene.bin scene64.bin\n"); return 2; }
    std::vector<unsigned char> b32, b64;
    if (!slurp(argv[1], b32) || !slurp(argv[2], b64)) { std::fprintf(stderr, "cannot read the scene dumps\n"); return 1; }
    uint64_t nn, nt; std::memcpy(&nn, b32.data(), 8); std::memcpy(&nt, b32.data() + 8, 8);
    const unsigned char* nodes = b32.data() + 16; const size_t* idx = (const size_t*)(nodes + nn * 32);
    const unsigned char* tris = (const unsigned char*)(idx + nt); const unsigned char* norms = tris + nt * 48;
    const size_t W = 320, H = 180;
    ceres_camera cam = {{0.0f, 0.1f, -0.3f}, {0.0f, 0.0f, 1.0f}, {0.0f, 1.0f, 0.0f}, 60.0f};
    const float sun[3] = {-50.0f, -20.0f, 0.0f};
    uint64_t c0[2], c[2];

    ceres_scene *s = nullptr, *s2 = nullptr;
    CHECK(ceres_scene_create(nodes, nn, idx, tris, norms, nt, 0, &s));
    CHECK(ceres_scene_create(nodes, nn, idx, tris, norms, nt, 0, &s2));
    std::vector<float> px(3 * W * H), px2(3 * W * H);
    std::vector<uint32_t> ids(W * H);
    // 1. host frame, with ids (frame kernel in row bands, tight copies)
    CHECK(ceres_render(s, &cam, sun, W, H, 0, px.data(), ids.data(), c0));
    std::printf("host frame: rays %llu hits %llu (%s build)\n", (unsigned long long)c0[0], (unsigned long long)c0[1], ceres_build_is_checked() ? "CHECKED" : "product");
    NOERR(s, "host frame");
    // 1b. the same frame again and again with every band count and sharing on / off: a race in the queue protocol shows up as a
    //     frame that differs from the first one (pixels, ids and counters are compared bit for bit)
    {
        std::vector<uint32_t> ids2(W * H);
        int diffs = 0;
        for (unsigned rep = 0; rep < 40; ++rep) {
            ceres_tuning t{}; t.reserved[0] = 1 + rep % 2; t.reserved[1] = 1 + rep % 8; t.ctas_per_sm = rep % 5 == 4 ? 3 : 0;
            CHECK(ceres_scene_set_tuning(s, &t));
            CHECK(ceres_render(s, &cam, sun, W, H, 0, px2.data(), ids2.data(), c));
            if (c[0] != c0[0] || c[1] != c0[1] || std::memcmp(px.data(), px2.data(), px.size() * 4) || std::memcmp(ids.data(), ids2.data(), ids.size() * 4)) ++diffs;
        }
        CHECK(ceres_scene_set_tuning(s, nullptr));
        std::printf("repeat-and-compare: 40 frames (1-8 bands, sharing on/off, 8 or 3 CTAs per SM), %d differ from the first\n", diffs);
        if (diffs) ++g_bad;
        NOERR(s, "repeat-and-compare");
    }
    // 2. sharing forced on and off, statistics build, timeline build, primary only
    for (unsigned share = 1; share <= 2; ++share) {
        ceres_tuning t{}; t.reserved[0] = share; t.reserved[1] = 1;
        CHECK(ceres_scene_set_tuning(s, &t));
        CHECK(ceres_render(s, &cam, sun, W, H, 0, px2.data(), nullptr, c));
        if (c[0] != c0[0] || c[1] != c0[1] || std::memcmp(px.data(), px2.data(), px.size() * 4)) { std::fprintf(stderr, "share %u: frame differs\n", share); return 1; }
        CHECK(ceres_scene_set_stats(s, 1));
        CHECK(ceres_render(s, &cam, sun, W, H, 0, px2.data(), nullptr, c));
        CHECK(ceres_scene_set_stats(s, 0));
        CHECK(ceres_scene_set_timeline(s, 1));
        CHECK(ceres_render(s, &cam, sun, W, H, 0, px2.data(), nullptr, c));
        CHECK(ceres_scene_set_timeline(s, 0));
    }
    CHECK(ceres_scene_set_tuning(s, nullptr));
    CHECK(ceres_render(s, &cam, sun, W, H, CERES_NO_SHADOWS, px2.data(), nullptr, c));
    NOERR(s, "sharing / statistics / timeline / primary-only frames");
    // 3. shard 3 of 8, scattered into the frame and packed
    {
        ceres_shard sh = {CERES_TILE_BLOCK_ROWS, 3, 8};
        std::vector<float> f(3 * W * H, 0.0f);
        CHECK(ceres_render_shard(s, &cam, sun, W, H, 0, &sh, f.data(), nullptr, c));
        CHECK(ceres_render_shard(s, &cam, sun, W, H, CERES_COMPACT_ROWS, &sh, f.data(), nullptr, c));
        NOERR(s, "shards");
    }
    // 4. P6 payloads
    {
        std::vector<uint8_t> q8(3 * W * H), q16(6 * W * H);
        CHECK(ceres_render_ppm8(s, &cam, sun, W, H, 0, q8.data(), c));
        CHECK(ceres_render_ppm16(s, &cam, sun, W, H, 0, q16.data(), c));
        NOERR(s, "P6 payloads");
    }
    // 5. device frame (phase-0 clear), quantise on the device, trace_rays
    {
        float* d_px = nullptr; uint32_t* d_ids = nullptr; uint64_t* d_cnt = nullptr; uint8_t* d_q = nullptr;
        CU(cudaMalloc((void**)&d_px, 3 * W * H * 4)); CU(cudaMalloc((void**)&d_ids, W * H * 4)); CU(cudaMalloc((void**)&d_cnt, 16)); CU(cudaMalloc((void**)&d_q, 3 * W * H));
        CU(cudaMemset(d_px, 0xFF, 3 * W * H * 4));
        CHECK(ceres_render_device(s, &cam, sun, W, H, 0, nullptr, nullptr, d_px, d_ids, d_cnt));
        CHECK(ceres_quantize_device(s, d_px, W, H, nullptr, d_q));
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(px2.data(), d_px, px2.size() * 4, cudaMemcpyDeviceToHost));
        if (std::memcmp(px.data(), px2.data(), px.size() * 4)) { std::fprintf(stderr, "device frame differs from the host frame\n"); return 1; }
        // the same frame over the library's static scratch, twice (the second call runs over what the first left there)
        for (int rep = 0; rep < 2; ++rep) {
            CU(cudaMemset(d_px, 0xFF, 3 * W * H * 4));
            CHECK(ceres_render_device_static(s, &cam, sun, W, H, 0, nullptr, d_px, d_ids, d_cnt));
            CU(cudaDeviceSynchronize());
            CU(cudaMemcpy(px2.data(), d_px, px2.size() * 4, cudaMemcpyDeviceToHost));
            if (std::memcmp(px.data(), px2.data(), px.size() * 4)) { std::fprintf(stderr, "static-scratch frame differs from the host frame\n"); return 1; }
        }
        if (ceres_render_device_static(s, &cam, sun, 4096, 4096, 0, nullptr, d_px, d_ids, d_cnt) == CERES_OK) { std::fprintf(stderr, "static scratch: an oversized frame was accepted\n"); return 1; }
        ceres_shard sh = {CERES_TILE_BLOCK_ROWS, 1, 3};
        CHECK(ceres_render_device(s, &cam, sun, W, H, CERES_COMPACT_ROWS, &sh, nullptr, d_px, d_ids, d_cnt));
        CU(cudaDeviceSynchronize());
        cudaFree(d_px); cudaFree(d_ids); cudaFree(d_cnt); cudaFree(d_q);
        std::vector<float> rays(8 * 4096), tuv(3 * 4096); std::vector<uint32_t> rid(4096);
        for (size_t k = 0; k < 4096; ++k) {
            float* r = &rays[8 * k];
            r[0] = 0.0f; r[1] = 0.1f; r[2] = -0.3f;
            r[3] = 0.002f * (float)(k % 64) - 0.064f; r[4] = 0.002f * (float)(k / 64) - 0.064f; r[5] = 1.0f; r[6] = 0.0f; r[7] = 3.0e38f;
        }
        CHECK(ceres_trace_rays(s, rays.data(), 4096, rid.data(), tuv.data()));
        NOERR(s, "device frames / trace_rays");
    }
    // 6. peer frame: the owner alone (world 1), then two ranks on this GPU, rank 0 enqueued first on its own stream
    {
        ceres_peer_frame* solo = nullptr;
        CHECK(ceres_peer_frame_create(s, W, H, 1, 1, &solo, nullptr));
        uint64_t* d_cnt = nullptr; CU(cudaMalloc((void**)&d_cnt, 16));
        for (int e = 0; e < 2; ++e) CHECK(ceres_render_peer(s, &cam, sun, 0, nullptr, solo, d_cnt));
        CU(cudaDeviceSynchronize());
        std::printf("peer frame, world 1: error word %d\n", ceres_peer_frame_error(solo));
        ceres_peer_frame_destroy(solo);
        ceres_peer_frame *own = nullptr, *ali = nullptr;
        CHECK(ceres_peer_frame_create(s, W, H, 1, 2, &own, nullptr));
        CHECK(ceres_peer_frame_alias(own, s2, 1, &ali));
        cudaStream_t st0, st1; CU(cudaStreamCreateWithFlags(&st0, cudaStreamNonBlocking)); CU(cudaStreamCreateWithFlags(&st1, cudaStreamNonBlocking));
        ceres_camera cam2 = cam; cam2.eye[0] = 0.05f;
        for (int e = 0; e < 3; ++e) {
            const ceres_camera* v = e == 1 ? &cam2 : &cam;      // a moved rectangle exercises the dirty-rectangle clear
            CHECK(ceres_render_peer(s, v, sun, 0, st0, own, d_cnt));
            CHECK(ceres_render_peer(s2, v, sun, 0, st1, ali, nullptr));
            CU(cudaDeviceSynchronize());
        }
        CU(cudaMemcpy(c, d_cnt, 16, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(px2.data(), ceres_peer_frame_pixels(own), px2.size() * 4, cudaMemcpyDeviceToHost));
        const int err = ceres_peer_frame_error(own);
        std::printf("peer frame, world 2 on one GPU: error word %d (a tool that serialises kernels makes rank 0 time out: 2), rays %llu hits %llu, frame %s\n",
                    err, (unsigned long long)c[0], (unsigned long long)c[1], std::memcmp(px.data(), px2.data(), px.size() * 4) ? "DIFFERS" : "equal");
        if (err || std::memcmp(px.data(), px2.data(), px.size() * 4)) ++g_bad;
        NOERR(s, "peer frames (rank 0)"); NOERR(s2, "peer frames (rank 1)");
        ceres_peer_frame_destroy(ali); ceres_peer_frame_destroy(own);
        cudaStreamDestroy(st0); cudaStreamDestroy(st1); cudaFree(d_cnt);
    }
    // 7. sequence: two frames in flight from two host threads over the same geometry
    {
        std::vector<float> cams(13 * 4), seq(3 * W * H * 4);
        std::vector<uint64_t> cc(8);
        for (int k = 0; k < 4; ++k) {
            float* q = &cams[13 * k];
            q[0] = 0.01f * k; q[1] = 0.1f; q[2] = -0.3f; q[3] = 0; q[4] = 0; q[5] = 1; q[6] = 0; q[7] = 1; q[8] = 0; q[9] = 60.0f; q[10] = sun[0]; q[11] = sun[1]; q[12] = sun[2];
        }
        CHECK(ceres_render_sequence(s, 4, cams.data(), W, H, 0, seq.data(), cc.data()));
        if (std::memcmp(seq.data(), px.data(), px.size() * 4)) { std::fprintf(stderr, "sequence frame 0 differs\n"); return 1; }
        NOERR(s, "sequence");
    }
    // 8. scene preparation: rotate, refit, render, download
    {
        CHECK(ceres_scene_rotate(s2, 1, 10.0f));
        CHECK(ceres_scene_refit(s2));
        CHECK(ceres_render(s2, &cam, sun, W, H, 0, px2.data(), nullptr, c));
        std::vector<unsigned char> nd(nn * 32), td(nt * 48);
        CHECK(ceres_scene_download(s2, nd.data(), nn, td.data(), nt));
        const float a[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, v[3] = {0.01f, 0, 0};
        CHECK(ceres_scene_transform(s2, a, v));
        CHECK(ceres_scene_refit(s2));
        CHECK(ceres_render(s2, &cam, sun, W, H, 0, px2.data(), nullptr, c));
        NOERR(s2, "scene preparation");
    }
    // 9. BVH build on the GPU, host copy and device-resident
    {
        std::vector<unsigned char> nd((2 * nt - 1) * 32); std::vector<size_t> pi(nt); double ms = 0;
        for (int algo = 1; algo <= 3; ++algo) {         // radix tree, top-down binned SAH (the default), locally-ordered clustering
            CHECK(ceres_build_options(algo, 0, 0, 0.0f));
            CHECK(ceres_build_lbvh(tris, nt, 0, nd.data(), pi.data(), &ms));
            ceres_scene* sl = nullptr;
            CHECK(ceres_scene_create_lbvh(tris, norms, nt, 0, &sl, &ms));
            CHECK(ceres_render(sl, &cam, sun, W, H, 0, px2.data(), nullptr, c));
            if (c[0] != c0[0]) { std::fprintf(stderr, "lbvh scene (algorithm %d): rays differ\n", algo); return 1; }
            CHECK(ceres_scene_refit(sl));
            CHECK(ceres_render(sl, &cam, sun, W, H, 0, px2.data(), nullptr, c));
            NOERR(sl, "GPU-built scene");
            ceres_scene_destroy(sl);
        }
        CHECK(ceres_build_options(0, 0, 0, 0.0f));
    }
    // 10. Scalar = double
    {
        uint64_t n2, t2; std::memcpy(&n2, b64.data(), 8); std::memcpy(&t2, b64.data() + 8, 8);
        const unsigned char* nodes64 = b64.data() + 16; const size_t* idx64 = (const size_t*)(nodes64 + n2 * 64);
        const unsigned char* tris64 = (const unsigned char*)(idx64 + t2); const unsigned char* norms64 = tris64 + t2 * 96;
        ceres_scene_f64* d = nullptr;
        CHECK(ceres_scene_f64_create(nodes64, n2, idx64, tris64, norms64, t2, 0, &d));
        ceres_camera_f64 cd = {{0.0, 0.1, -0.3}, {0.0, 0.0, 1.0}, {0.0, 1.0, 0.0}, 60.0};
        const double sd[3] = {-50.0, -20.0, 0.0};
        std::vector<double> pd(3 * W * H); std::vector<uint32_t> idd(W * H);
        CHECK(ceres_render_f64(d, &cd, sd, W, H, 0, pd.data(), idd.data(), c, nullptr));
        std::printf("double frame: rays %llu hits %llu\n", (unsigned long long)c[0], (unsigned long long)c[1]);
        {   // the same frame from 3 row-block shards into a second buffer
            std::vector<double> ps(3 * W * H, -1.0); std::vector<uint32_t> ids2(W * H, 7u);
            uint64_t sum[2] = {0, 0}, cs[2];
            for (uint32_t k = 0; k < 3; ++k) {
                ceres_shard sh = {CERES_TILE_BLOCK_ROWS, k, 3};
                CHECK(ceres_render_f64_shard(d, &cd, sd, W, H, 0, &sh, ps.data(), ids2.data(), cs, nullptr));
                sum[0] += cs[0]; sum[1] += cs[1];
            }
            if (std::memcmp(ps.data(), pd.data(), pd.size() * 8) || std::memcmp(ids2.data(), idd.data(), idd.size() * 4) || sum[0] != c[0] || sum[1] != c[1]) {
                std::fprintf(stderr, "double frame: the shards do not add up to the frame\n"); return 1;
            }
        }
        ceres_scene_f64_destroy(d);
    }
    double gbs = 0;
    CHECK(ceres_probe_read_bandwidth(0, 1 << 20, 2, &gbs));
    ceres_scene_destroy(s2); ceres_scene_destroy(s);
    std::printf("sanitize_driver: all calls returned CERES_OK, %llu kernel launches, %d failed check(s)\n", (unsigned long long)ceres_kernel_launch_count(), g_bad);
    return g_bad ? 3 : 0;
}
