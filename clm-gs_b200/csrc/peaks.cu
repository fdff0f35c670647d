// Roofline denominators measured on the box the bench runs on ("builder measures pinned H2D and D2H
// cudaMemcpyAsync ... and an FMA micro-benchmark ... STREAM-like measurement on the box").  Measurement helpers
// only: none of them is on the training path.
#include <algorithm>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

// 8 independent FMA chains per thread: enough ILP to keep the FP32 pipe of every sub-partition issuing each cycle.
__global__ void __launch_bounds__(1024) fma_chain_kernel(int iters, float seed, float *sink) {
    float a0 = seed + threadIdx.x, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
    float a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 1.0000001f, c = 1e-7f;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456f) sink[0] = s;  // never true; keeps the chains alive
}

}  // namespace

// One launch = grid_ctas x 1024 threads x iters x 16 x 8 FMAs (2 FLOP each).  The caller times it with events.
CLMGS_EXPORT int clmgs_measure_fma(int grid_ctas, int iters, float *sink, void *stream) {
    CLMGS_REQUIRE(grid_ctas > 0 && iters > 0 && sink, "measure_fma: bad arguments");
    fma_chain_kernel<<<grid_ctas, 1024, 0, as_stream(stream)>>>(iters, 0.5f, sink);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// cudaMemcpyAsync between any two pointers (pinned host <-> device for the host-link peak).
CLMGS_EXPORT int clmgs_measure_memcpy(void *dst, const void *src, size_t bytes, void *stream) {
    CLMGS_REQUIRE(dst && src, "measure_memcpy: null pointer");
    CLMGS_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, as_stream(stream)));
    return CLMGS_OK;
}

// STREAM triad a[i] = b[i] + s * c[i] on n_threads host threads (blocking); 12 bytes of DRAM traffic per element
// (16 with write-allocate, which the bench does not count).
CLMGS_EXPORT int clmgs_measure_host_triad(float *host_a, const float *host_b, const float *host_c, int64_t n, float s,
                                          int n_threads) {
    CLMGS_REQUIRE(host_a && host_b && host_c && n > 0, "measure_host_triad: bad arguments");
    if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; ++t)
        th.emplace_back([=] {
            const int64_t a = n * t / n_threads, b = n * (t + 1) / n_threads;
            for (int64_t i = a; i < b; ++i) host_a[i] = host_b[i] + s * host_c[i];
        });
    for (auto &t : th) t.join();
    return CLMGS_OK;
}
