// Probe (development tool): drain the frame's screen rectangle to a page-locked HOST frame with SM stores through the
// mapped pointer (no copy engine) -- against the 2-D DMA of copy_overlap_probe.cu.  With and without a kernel that
// occupies half of every SM's thread slots.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o zero_copy_probe zero_copy_probe.cu
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstring>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
__global__ void spin(long long cycles, int* sink) {
    long long t0 = clock64();
    int x = 0;
    while (clock64() - t0 < cycles) x += threadIdx.x;
    if (x == 123456789) *sink = x;
}
// rows of `row_vec` float4 each, pitch `pitch_vec` float4 on both sides
__global__ void __launch_bounds__(256) drain(const float4* __restrict__ src, float4* __restrict__ dst, size_t pitch_vec, size_t row_vec, size_t rows) {
    const size_t total = row_vec * rows;
    for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < total; k += (size_t)gridDim.x * 256u) {
        const size_t r = k / row_vec, c = k - r * row_vec;
        __stcs(dst + r * pitch_vec + c, __ldcs(src + r * pitch_vec + c));
    }
}
int main() {
    const size_t W = 3840, H = 2160, rc = 2092, rr = 1182;     // 4K frame, dragon rectangle (columns rounded to 4)
    float *d, *h; int* sink;
    cudaMalloc(&d, W * H * 12); cudaMallocHost(&h, W * H * 12); cudaMalloc(&sink, 4);
    cudaMemset(d, 0x3f, W * H * 12);
    memset(h, 0, W * H * 12);
    cudaStream_t a, b; cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t pitch_vec = W * 12 / 16, row_vec = rc * 12 / 16;
    const int grids[] = {0, 4, 8, 16, 32, 74, 148, 296, 592};
    for (int with_kernel = 0; with_kernel < 2; ++with_kernel)
    for (int g : grids) {
        double best = 1e9;
        for (int rep = 0; rep < 5; ++rep) {
            cudaDeviceSynchronize();
            double t0 = now();
            if (with_kernel) spin<<<sms * 8, 128, 0, a>>>(2000000, sink);
            if (g == 0) cudaMemcpy2DAsync(h, W * 12, d, W * 12, rc * 12, rr, cudaMemcpyDeviceToHost, b);
            else drain<<<g, 256, 0, b>>>((const float4*)d, (float4*)h, pitch_vec, row_vec, rr);
            cudaStreamSynchronize(b);
            double t1 = now();
            cudaDeviceSynchronize();
            if (t1 - t0 < best) best = t1 - t0;
        }
        const double mb = rc * rr * 12 / 1e6;
        if (g == 0) printf("%s kernel, 2-D DMA              : %.3f ms  %.1f GB/s (%.1f MB)\n", with_kernel ? "with" : "no  ", best * 1e3, mb / best / 1e3, mb);
        else        printf("%s kernel, SM stores, %3d CTAs  : %.3f ms  %.1f GB/s\n", with_kernel ? "with" : "no  ", g, best * 1e3, mb / best / 1e3);
    }
    // correctness of the last drain
    size_t bad = 0;
    for (size_t r = 0; r < rr; ++r) for (size_t c = 0; c < rc * 3; ++c) { unsigned u; memcpy(&u, &h[r * W * 3 + c], 4); bad += u != 0x3f3f3f3fu; }
    printf("mismatching words in the rectangle: %zu\n", bad);
    return 0;
}
