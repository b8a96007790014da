// Probe (development tool): how fast can host threads zero-fill page-locked memory, next to a concurrent D2H DMA?
// build: nvcc -O3 -std=c++17 -Xcompiler -mavx2 -o hostfill_probe hostfill_probe.cu -lpthread
#include <cuda_runtime.h>
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void fill_nt(char* p, size_t n) {
    __m256i z = _mm256_setzero_si256();
    size_t k = 0;
    for (; k + 32 <= n; k += 32) _mm256_stream_si256((__m256i*)(p + k), z);
    _mm_sfence();
}
int main() {
    const size_t N = 100u << 20;
    char *h, *d;
    cudaMallocHost(&h, N); cudaMalloc(&d, N); cudaMemset(d, 1, N);
    memset(h, 1, N);
    cudaStream_t st; cudaStreamCreate(&st);
    printf("hw threads %u\n", std::thread::hardware_concurrency());
    for (int mode = 0; mode < 2; ++mode)
    for (int T : {1, 2, 4, 8, 16}) {
        double best = 1e9, best_dma = 1e9;
        for (int rep = 0; rep < 5; ++rep) {
            double t0 = now();
            std::vector<std::thread> th;
            const size_t half = N / 2, per = half / T / 64 * 64;
            cudaMemcpyAsync(h + half, d, half, cudaMemcpyDeviceToHost, st);   // concurrent DMA into the other half
            for (int t = 0; t < T; ++t) th.emplace_back([=] { if (mode) fill_nt(h + t * per, per); else memset(h + t * per, 0, per); });
            for (auto& x : th) x.join();
            double t1 = now();
            cudaStreamSynchronize(st);
            double t2 = now();
            if (t1 - t0 < best) best = t1 - t0;
            if (t2 - t0 < best_dma) best_dma = t2 - t0;
        }
        printf("%s threads %2d: fill %5.1f MB in %.3f ms = %5.1f GB/s (incl. thread create/join); with 52 MB DMA done at %.3f ms\n",
               mode ? "nt-store" : "memset  ", T, N / 2 / 1e6, best * 1e3, N / 2 / best / 1e9, best_dma * 1e3);
    }
    double t0 = now(); cudaMemcpyAsync(h, d, N, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st);
    printf("plain D2H %zu MB: %.3f ms\n", N >> 20, (now() - t0) * 1e3);
    return 0;
}
