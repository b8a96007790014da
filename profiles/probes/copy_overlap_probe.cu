// Probe (development tool): does a D2H copy overlap a GPU-filling kernel?  linear vs 2D (pitched) copies.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o copy_overlap_probe copy_overlap_probe.cu
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
__global__ void spin(long long cycles, int* sink) {
    long long t0 = clock64();
    int x = 0;
    while (clock64() - t0 < cycles) x += threadIdx.x;
    if (x == 123456789) *sink = x;
}
int main() {
    const size_t W = 3840, H = 2160, rc = 2090, rr = 1182;     // 4K frame, dragon rectangle
    float *d, *h; int* sink;
    cudaMalloc(&d, W * H * 12); cudaMallocHost(&h, W * H * 12); cudaMalloc(&sink, 4);
    cudaMemset(d, 0, W * H * 12);
    cudaStream_t a, b; cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int with_kernel = 0; with_kernel < 2; ++with_kernel)
    for (int mode = 0; mode < 3; ++mode) {
        double best = 1e9, best_all = 1e9;
        for (int rep = 0; rep < 5; ++rep) {
            cudaDeviceSynchronize();
            double t0 = now();
            if (with_kernel) spin<<<sms * 16, 128, 0, a>>>(2000000, sink);     // ~1 ms at 1.9 GHz, all SM slots taken
            if (mode == 0) cudaMemcpyAsync(h, d, rc * rr * 12, cudaMemcpyDeviceToHost, b);                                  // linear, same bytes
            if (mode == 1) cudaMemcpy2DAsync(h, W * 12, d, W * 12, rc * 12, rr, cudaMemcpyDeviceToHost, b);                 // pitched both sides
            if (mode == 2) cudaMemcpy2DAsync(h, W * 12, d, rc * 12, rc * 12, rr, cudaMemcpyDeviceToHost, b);                // compact source
            cudaStreamSynchronize(b);
            double t1 = now();
            cudaDeviceSynchronize();
            double t2 = now();
            if (t1 - t0 < best) best = t1 - t0;
            if (t2 - t0 < best_all) best_all = t2 - t0;
        }
        const char* names[3] = {"linear", "2D pitched src+dst", "2D compact src"};
        printf("%s kernel, %-20s: copy done %.3f ms, all done %.3f ms (%.1f MB)\n", with_kernel ? "with" : "no  ", names[mode], best * 1e3, best_all * 1e3, rc * rr * 12 / 1e6);
    }
    return 0;
}
