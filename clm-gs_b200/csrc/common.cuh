// Shared host/device helpers for libclmgs.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/clmgs.h"

#define CLMGS_EXPORT extern "C" __attribute__((visibility("default")))

void clmgs_set_error(const char *fmt, ...);

#define CLMGS_REQUIRE(cond, ...)                 \
    do {                                         \
        if (!(cond)) {                           \
            clmgs_set_error(__VA_ARGS__);        \
            return CLMGS_ERR_INVALID;            \
        }                                        \
    } while (0)

#define CLMGS_CUDA(call)                                                              \
    do {                                                                              \
        cudaError_t e__ = (call);                                                     \
        if (e__ != cudaSuccess) {                                                     \
            clmgs_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,              \
                            cudaGetErrorString(e__));                                 \
            return CLMGS_ERR_CUDA;                                                    \
        }                                                                             \
    } while (0)

// after a kernel launch: catches bad launch configurations without syncing
#define CLMGS_LAUNCH_OK() CLMGS_CUDA(cudaGetLastError())

static inline cudaStream_t as_stream(void *s) { return (cudaStream_t)s; }
__host__ __device__ static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

constexpr int kNumSMs = 148;  // B200

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
