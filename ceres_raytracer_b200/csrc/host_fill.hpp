// host_fill.hpp -- a small pool of host threads that fills regions of the caller's framebuffer with a byte pattern.
//
// Why: render() must write EVERY pixel (render.hpp:107,117,149), but pixels outside the conservative screen rectangle
// of the scene are 0 by construction (ids: CERES_MISS).  Shipping those zeros over PCIe costs 18.5 us per MB
// (measured 54-57 GB/s D2H on the B200 box); a few host cores write them with streaming stores at 110-134 GB/s
// (profiles/probes/hostfill_probe.cu), concurrently with the GPU kernels and with the DMA of the rectangle itself.
// Nothing here computes anything: it is memset for the part of the output the device never needs to touch.
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__AVX2__)
#include <immintrin.h>
#endif

namespace ceres_host {

// src == nullptr: fill n bytes with `byte`; widen == false: copy n bytes from src; widen == true: n floats at src -> n doubles at p
struct Segment { unsigned char* p; size_t n; unsigned char byte; const unsigned char* src = nullptr; bool widen = false; };

// float -> double, exact (the double-precision path's colours are float-valued, render.hpp:56-84): the frame crosses PCIe
// as floats and is widened into the caller's double buffer here
inline void widen_f32(double* dst, const float* src, size_t n) {
    size_t k = 0;
#if defined(__AVX2__)
    for (; k + 4 <= n; k += 4) _mm256_storeu_pd(dst + k, _mm256_cvtps_pd(_mm_loadu_ps(src + k)));
#endif
    for (; k < n; ++k) dst[k] = (double)src[k];
}

inline void fill_bytes(unsigned char* p, size_t n, unsigned char byte) {
#if defined(__AVX2__)
    static const bool cached = std::getenv("CERES_HOST_FILL_CACHED") != nullptr;      // experiment switch: plain memset
    if (n >= 256 && !cached) {
        // streaming stores: the framebuffer is write-only here, do not pull it through the caches
        size_t head = (32 - (reinterpret_cast<uintptr_t>(p) & 31)) & 31;
        std::memset(p, byte, head);
        p += head; n -= head;
        const __m256i v = _mm256_set1_epi8(static_cast<char>(byte));
        size_t k = 0;
        for (; k + 128 <= n; k += 128) {
            _mm256_stream_si256(reinterpret_cast<__m256i*>(p + k), v);
            _mm256_stream_si256(reinterpret_cast<__m256i*>(p + k + 32), v);
            _mm256_stream_si256(reinterpret_cast<__m256i*>(p + k + 64), v);
            _mm256_stream_si256(reinterpret_cast<__m256i*>(p + k + 96), v);
        }
        for (; k + 32 <= n; k += 32) _mm256_stream_si256(reinterpret_cast<__m256i*>(p + k), v);
        _mm_sfence();
        std::memset(p + k, byte, n - k);
        return;
    }
#endif
    std::memset(p, byte, n);
}

// One job = a list of segments; workers (and the submitting thread, in wait()) claim segments through an atomic index.
struct Job {
    std::vector<Segment> segs;
    std::atomic<size_t> next{0};
    std::atomic<size_t> done{0};
    std::atomic<int> holders{0};      // pool workers currently inside this job
    void add(void* p, size_t n, unsigned char byte) {
        // split big ranges so that the pool balances
        unsigned char* b = static_cast<unsigned char*>(p);
        const size_t piece = 1u << 20;
        while (n > piece + piece / 2) { segs.push_back({b, piece, byte}); b += piece; n -= piece; }
        if (n) segs.push_back({b, n, byte});
    }
    void add_copy(void* dst, const void* src, size_t bytes) { if (bytes) segs.push_back({static_cast<unsigned char*>(dst), bytes, 0, static_cast<const unsigned char*>(src), false}); }
    void add_widen(double* dst, const float* src, size_t count) { if (count) segs.push_back({reinterpret_cast<unsigned char*>(dst), count, 0, reinterpret_cast<const unsigned char*>(src), true}); }
    bool work_one() {
        const size_t k = next.fetch_add(1, std::memory_order_relaxed);
        if (k >= segs.size()) return false;
        const Segment& g = segs[k];
        if (!g.src) fill_bytes(g.p, g.n, g.byte);
        else if (g.widen) widen_f32(reinterpret_cast<double*>(g.p), reinterpret_cast<const float*>(g.src), g.n);
        else std::memcpy(g.p, g.src, g.n);
        done.fetch_add(1, std::memory_order_release);
        return true;
    }
    bool finished() const { return done.load(std::memory_order_acquire) >= segs.size(); }
};

class Pool {
public:
    static Pool& instance() { static Pool p; return p; }
    // Starts the job on the pool; the caller keeps ownership and must call wait(job) before destroying it.
    void submit(Job* job) {
        if (job->segs.empty()) return;
        {
            std::lock_guard<std::mutex> lk(m_);
            jobs_.push_back(job);
        }
        cv_.notify_all();
    }
    // The submitting thread helps, then waits until the workers have finished their pieces.
    void wait(Job* job) {
        if (job->segs.empty()) return;
        while (job->work_one()) {}
        while (!job->finished()) std::this_thread::yield();
        {
            std::lock_guard<std::mutex> lk(m_);     // workers pick jobs under this lock: after the erase no new holder appears
            for (auto it = jobs_.begin(); it != jobs_.end(); ++it) if (*it == job) { jobs_.erase(it); break; }
        }
        while (job->holders.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }
    unsigned workers() const { return (unsigned)threads_.size(); }

private:
    Pool() {
        unsigned hw = std::thread::hardware_concurrency();
        unsigned n = hw >= 16 ? 7 : (hw >= 8 ? 3 : (hw >= 4 ? 1 : 0));     // + the submitting thread
        // one process per GPU (torchrun sets LOCAL_WORLD_SIZE): the ranks of a box share its cores -- 8 pools of 8 threads on 32
        // cores fill a shared frame slower than one pool does (profiles/r2/sweeps.md)
        if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) {
            const unsigned ranks = (unsigned)std::max(1, std::atoi(e));
            if (ranks > 1) n = std::min(n, std::max(1u, hw / ranks) - (hw / ranks > 1 ? 1u : 0u));
        }
        if (const char* e = std::getenv("CERES_HOST_FILL_THREADS")) n = (unsigned)std::max(0, std::atoi(e));
        for (unsigned k = 0; k < n; ++k) threads_.emplace_back([this] { run(); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> lk(m_); stop_ = true; }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    void run() {
        for (;;) {
            Job* job = nullptr;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] {
                    if (stop_) return true;
                    for (Job* j : jobs_) if (j->next.load(std::memory_order_relaxed) < j->segs.size()) return true;
                    return false;
                });
                if (stop_) return;
                for (Job* j : jobs_) if (j->next.load(std::memory_order_relaxed) < j->segs.size()) { job = j; break; }
                if (job) job->holders.fetch_add(1, std::memory_order_acq_rel);
            }
            // wait() does not return (and the owner does not destroy the job) while holders != 0
            if (job) { while (job->work_one()) {} job->holders.fetch_sub(1, std::memory_order_acq_rel); }
        }
    }
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Job*> jobs_;
    std::vector<std::thread> threads_;
    bool stop_ = false;
};

}  // namespace ceres_host
