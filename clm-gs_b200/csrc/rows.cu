// S2 / S6: movement of the offloaded 48-float (192 B) SH rows between pinned
// host memory and HBM, with cross-camera retention.
//
// Reference: CK/clm_kernels.cu:5-289 moves every float with its own 4-byte
// zero-copy access from only 32 CTAs.  Here a row is 12 x 16-byte accesses by
// 12 consecutive lanes (a row of a pinned [N,48] fp32 table is 64-B aligned, so
// float4 is always legal), the grid is sized to keep all 148 SMs issuing, and
// each thread keeps UNROLL independent loads in flight -- what a PCIe-latency
// bound transfer needs.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int kRowF4 = 12;       // float4 per row
constexpr int kThreads = 256;
constexpr int kUnroll = 4;

__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Transfer kernels run on a side stream next to the render kernels and spend their time waiting on
// PCIe.  Every zero-copy load holds an L1 miss slot of its SM for ~2 us, so a host-bound CTA slows
// whatever else runs on that SM (measured at C4: every render kernel 2-3x slower with one transfer
// CTA on each of the 148 SMs; with all thread slots filled a 0.4 ms Adam took 24 ms).  The link needs
// only bandwidth x latency = ~50 GB/s x 2 us = 100 KB in flight: 32 CTAs x 256 threads x 4 x 16 B
// = 512 KB covers it while touching a fifth of the SMs (the reference also uses 32 CTAs, but with
// 4-byte accesses).
inline int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}
// tunables (read once): CTAs of kernels that LOAD from host / STORE to host
inline int host_load_ctas() { static int v = env_int("CLMGS_HOST_LOAD_CTAS", 48); return v; }
// threads per CTA of the host-LOADING kernels: the same bytes in flight from fewer, fatter CTAs disturb fewer SMs
inline int host_load_threads() {
    static int v = [] { int t = env_int("CLMGS_HOST_LOAD_THREADS", 256); return t < 64 ? 64 : (t > 1024 ? 1024 : (t / 32) * 32); }();
    return v;
}
inline int host_store_ctas() { static int v = env_int("CLMGS_HOST_STORE_CTAS", 64); return v; }
inline unsigned grid_for(int64_t n_rows, bool store = false) {
    int64_t work = ceil_div64(n_rows * kRowF4, (int64_t)kThreads * kUnroll);
    int64_t cap = store ? host_store_ctas() : host_load_ctas();
    return (unsigned)(work < 1 ? 1 : (work > cap ? cap : work));
}

// dst[dst_row(r)] = src[src_row(r)] for r in [0,n); row maps are functors.
template <typename SrcMap, typename DstMap>
__global__ void __launch_bounds__(1024)
copy_rows_kernel(float4 *__restrict__ dst, const float4 *__restrict__ src, int64_t n, SrcMap smap, DstMap dmap) {
    const int64_t total = n * kRowF4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;   // block size is a launch parameter (host_load_threads)
    for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e0 < total; e0 += stride * kUnroll) {
        float4 v[kUnroll];
        int64_t d[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            int64_t e = e0 + u * stride;
            d[u] = -1;
            if (e < total) {
                int64_t r = e / kRowF4;
                int c = (int)(e - r * kRowF4);
                v[u] = ld_stream(src + smap(r) * kRowF4 + c);
                d[u] = dmap(r) * kRowF4 + c;
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (d[u] >= 0) dst[d[u]] = v[u];
    }
}

template <typename SrcMap, typename DstMap>
__global__ void __launch_bounds__(kThreads)
accum_rows_kernel(float4 *__restrict__ dst, const float4 *__restrict__ src, int64_t n, SrcMap smap, DstMap dmap) {
    const int64_t total = n * kRowF4;
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    for (int64_t e0 = (int64_t)blockIdx.x * kThreads + threadIdx.x; e0 < total; e0 += stride * kUnroll) {
        float4 v[kUnroll], o[kUnroll];
        int64_t d[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            int64_t e = e0 + u * stride;
            d[u] = -1;
            if (e < total) {
                int64_t r = e / kRowF4;
                int c = (int)(e - r * kRowF4);
                d[u] = dmap(r) * kRowF4 + c;
                v[u] = ld_stream(src + smap(r) * kRowF4 + c);
                o[u] = dst[d[u]];
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (d[u] >= 0) dst[d[u]] = make_float4(o[u].x + v[u].x, o[u].y + v[u].y, o[u].z + v[u].z, o[u].w + v[u].w);
    }
}

// A transfer kernel sits on every SM for tens of milliseconds.  An SM cannot change its L1/shared
// split while a CTA is resident, so if these kernels ran with the default (L1-heavy) split, any
// render kernel that needs a large shared-memory carve-out -- the radix sort's onesweep passes --
// could not be placed until the transfer finished (measured: a 3.5 ms sort took 22 ms).  They use no
// L1 (streaming loads), so ask for the maximum shared carve-out and let everything co-reside.
template <typename K>
inline void prefer_shared(K kernel) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

struct Map32Via {  // row = table[ids[r]] (slot lookup)
    const int32_t *ids; const int32_t *table;
    __device__ int64_t operator()(int64_t r) const { return table[ids[r]]; }
};

struct Identity { __device__ int64_t operator()(int64_t r) const { return r; } };
struct Map64 { const int64_t *p; __device__ int64_t operator()(int64_t r) const { return p[r]; } };
struct Map32 { const int32_t *p; __device__ int64_t operator()(int64_t r) const { return p[r]; } };

template <typename W>
__device__ __forceinline__ bool test_bit(const void *bitmap, int64_t i, int bit) {
    return (((const W *)bitmap)[i] >> bit) & 1;
}

// position of `id` in the ascending array ids[0..n) (it is known to be present)
__device__ __forceinline__ int32_t rank_of(const int64_t *__restrict__ ids, int64_t n, int64_t id) {
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (__ldg(ids + mid) < id) lo = mid + 1; else hi = mid;
    }
    return (int32_t)lo;
}

// ---------------------------------------------------------------------------
// Retention transfers, two launches each:
//   *_dev_kernel  : rows that stay on the GPU (device -> device), full-occupancy grid, HBM-bound;
//   *_host_kernel : rows that cross the host link, 1 CTA/SM (see grid_for), PCIe-latency-bound.
// A 16-lane group handles 4 rows per iteration: lanes 0..3 classify one row each (bitmap test +
// binary search), lanes 0..11 then move the four 192-B rows with 4 independent 16-B loads in flight.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t load_word(const void *bitmap, int wb, int64_t i) {
    switch (wb) {
        case 1: return ((const uint8_t *)bitmap)[i];
        case 2: return ((const uint16_t *)bitmap)[i];
        case 4: return ((const uint32_t *)bitmap)[i];
        default: return ((const uint64_t *)bitmap)[i];
    }
}

// next rows that the current camera also sees: copy from the current buffer; records src_map for all rows
__global__ void __launch_bounds__(kThreads)
prefetch_dev_kernel(float4 *__restrict__ d_next, const float4 *__restrict__ d_this,
                    const int64_t *__restrict__ ids_this, int64_t n_this, const int64_t *__restrict__ ids_next,
                    int64_t n_next, const void *__restrict__ bitmap, int wb, int bit_this,
                    int32_t *__restrict__ src_map) {
    const int sub = threadIdx.x & 15;
    const int64_t groups = ((int64_t)gridDim.x * kThreads) >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    for (int64_t r0 = (((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 4) * 4; r0 < n_next; r0 += groups * 4) {
        int32_t my_src = -1;
        if (sub < 4 && r0 + sub < n_next) {
            const int64_t id = ids_next[r0 + sub];
            if (n_this > 0 && ((load_word(bitmap, wb, id) >> bit_this) & 1)) my_src = rank_of(ids_this, n_this, id);
            src_map[r0 + sub] = my_src;
        }
        float4 v[4];
        int32_t src[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            src[u] = __shfl_sync(gmask, my_src, u, 16);
            if (sub < kRowF4 && src[u] >= 0) v[u] = ld_stream(d_this + (int64_t)src[u] * kRowF4 + sub);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (sub < kRowF4 && src[u] >= 0) d_next[(r0 + u) * kRowF4 + sub] = v[u];
    }
}

// next rows that are not on the GPU (src_map < 0): zero-copy loads from the pinned table
__global__ void __launch_bounds__(kThreads)
prefetch_host_kernel(float4 *__restrict__ d_next, const float4 *__restrict__ h_rows,
                     const int64_t *__restrict__ ids_next, int64_t n_next, const int32_t *__restrict__ src_map) {
    const int sub = threadIdx.x & 15;
    const int64_t groups = ((int64_t)gridDim.x * kThreads) >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    for (int64_t r0 = (((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 4) * 4; r0 < n_next; r0 += groups * 4) {
        int64_t my_id = -1;
        if (sub < 4 && r0 + sub < n_next && src_map[r0 + sub] < 0) my_id = ids_next[r0 + sub];
        float4 v[4];
        int64_t id[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            id[u] = __shfl_sync(gmask, my_id, u, 16);
            if (sub < kRowF4 && id[u] >= 0) v[u] = ld_stream(h_rows + id[u] * kRowF4 + sub);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (sub < kRowF4 && id[u] >= 0) d_next[(r0 + u) * kRowF4 + sub] = v[u];
    }
}

// next_grad[r'] = grad_this[src_map[r']] for retained rows, 0 for rows new to the GPU
__global__ void __launch_bounds__(kThreads)
handoff_dev_kernel(const float4 *__restrict__ d_grad, float4 *__restrict__ d_next_grad, int64_t n_next,
                   const int32_t *__restrict__ src_map) {
    const int sub = threadIdx.x & 15;
    const int64_t groups = ((int64_t)gridDim.x * kThreads) >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    for (int64_t r0 = (((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 4) * 4; r0 < n_next; r0 += groups * 4) {
        int32_t my_src = (sub < 4 && r0 + sub < n_next) ? src_map[r0 + sub] : -1;
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int32_t src = __shfl_sync(gmask, my_src, u, 16);
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sub < kRowF4 && src >= 0) v[u] = d_grad[(int64_t)src * kRowF4 + sub];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (sub < kRowF4 && r0 + u < n_next) d_next_grad[(r0 + u) * kRowF4 + sub] = v[u];
    }
}

// Rows of this camera that the next one does not see go to the pinned gradient table.
// accum: 0 = store, 1 = read-modify-write over the link (the reference, CK/clm_kernels.cu:201),
//        2 = RMW only for rows that were already flushed once in this batch.  A row was flushed
//            before iff, among the cameras before the current run of consecutive visibility, some
//            camera saw it: with hi = word >> (bit_this+1) that is (hi & (hi+1)) != 0.  Everything
//            else is the first write to a zeroed row (cpu-adam leaves consumed rows zeroed) -> store.
__global__ void __launch_bounds__(kThreads)
handoff_host_kernel(const float4 *__restrict__ d_grad, float4 *__restrict__ h_grad,
                    const int64_t *__restrict__ ids_this, int64_t n_this, int has_next,
                    const void *__restrict__ bitmap, int wb, int bit_this, int bit_next, int accum) {
    const int sub = threadIdx.x & 15;
    const int64_t groups = ((int64_t)gridDim.x * kThreads) >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    for (int64_t r0 = (((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 4) * 4; r0 < n_this; r0 += groups * 4) {
        int64_t my_id = -1;
        int my_rmw = 0;
        if (sub < 4 && r0 + sub < n_this) {
            my_id = ids_this[r0 + sub];
            const uint64_t w = load_word(bitmap, wb, my_id);
            if (has_next && ((w >> bit_next) & 1)) my_id = -1;  // retained on the GPU
            else if (accum == 1) my_rmw = 1;
            else if (accum == 2) { const uint64_t hi = w >> (bit_this + 1); my_rmw = (hi & (hi + 1)) != 0; }
        }
        float4 v[4], o[4];
        int64_t id[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            id[u] = __shfl_sync(gmask, my_id, u, 16);
            const int rmw = __shfl_sync(gmask, my_rmw, u, 16);
            o[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sub < kRowF4 && id[u] >= 0) {
                v[u] = d_grad[(r0 + u) * kRowF4 + sub];
                if (rmw) o[u] = h_grad[id[u] * kRowF4 + sub];
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (sub < kRowF4 && id[u] >= 0)
                h_grad[id[u] * kRowF4 + sub] = make_float4(o[u].x + v[u].x, o[u].y + v[u].y, o[u].z + v[u].z, o[u].w + v[u].w);
    }
}

// dst[idx[r]] += src[r], device memory, indices may repeat (several ranks touching one row): 16-byte
// vector atomics (red.global.add.v4.f32, sm_90+).
__global__ void __launch_bounds__(kThreads)
accumulate_rows_kernel(float4 *__restrict__ dst, const float4 *__restrict__ src, const int64_t *__restrict__ idx,
                       int64_t n) {
    const int64_t total = n * kRowF4;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        const int64_t r = e / kRowF4;
        const int c = (int)(e - r * kRowF4);
        atomicAdd(dst + idx[r] * kRowF4 + c, src[e]);
    }
}

__global__ void set_signal_kernel(volatile int32_t *flags, int index, int value) {
    __threadfence_system();  // everything this stream wrote to host memory is visible first
    flags[index] = value;
    __threadfence_system();
}

void set_carveouts() {
    prefer_shared(copy_rows_kernel<Map64, Identity>);
    prefer_shared(copy_rows_kernel<Map32, Map32>);
    prefer_shared(copy_rows_kernel<Identity, Map64>);
    prefer_shared(accum_rows_kernel<Identity, Map64>);
    prefer_shared(accum_rows_kernel<Map32, Map32>);
    prefer_shared(copy_rows_kernel<Map32, Identity>);
    prefer_shared(copy_rows_kernel<Map32Via, Map32>);
    prefer_shared(prefetch_host_kernel);
    prefer_shared(handoff_host_kernel);
}

int check_align(const void *p, const char *what) {
    CLMGS_REQUIRE((((uintptr_t)p) & 15) == 0, "%s must be 16-byte aligned", what);
    return CLMGS_OK;
}

}  // namespace

#define F4(p) reinterpret_cast<float4 *>(p)
#define CF4(p) reinterpret_cast<const float4 *>(p)

CLMGS_EXPORT int clmgs_rows_gather_i64(float *d_rows, const float *h_rows, const int64_t *ids, int64_t n,
                                       void *stream) {
    set_carveouts();
    if (n == 0) return CLMGS_OK;
    if (int rc = check_align(d_rows, "d_rows")) return rc;
    if (int rc = check_align(h_rows, "h_rows")) return rc;
    copy_rows_kernel<<<grid_for(n), host_load_threads(), 0, as_stream(stream)>>>(F4(d_rows), CF4(h_rows), n, Map64{ids}, Identity{});
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_gather_retention(float *d_rows, const float *h_rows, const float *d_prev_rows,
                                             const int32_t *host_ids, const int32_t *prev_src_rows,
                                             const int32_t *dst_rows_from_host, const int32_t *dst_rows_from_prev,
                                             int64_t n_host, int64_t n_prev, void *stream) {
    set_carveouts();
    if (int rc = check_align(d_rows, "d_rows")) return rc;
    cudaStream_t st = as_stream(stream);
    if (n_host > 0) {
        copy_rows_kernel<<<grid_for(n_host), kThreads, 0, st>>>(F4(d_rows), CF4(h_rows), n_host, Map32{host_ids},
                                                               Map32{dst_rows_from_host});
        CLMGS_LAUNCH_OK();
    }
    if (n_prev > 0) {
        copy_rows_kernel<<<grid_for(n_prev), kThreads, 0, st>>>(F4(d_rows), CF4(d_prev_rows), n_prev,
                                                               Map32{prev_src_rows}, Map32{dst_rows_from_prev});
        CLMGS_LAUNCH_OK();
    }
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_scatter_grad_i64(const float *d_grad, float *h_grad, const int64_t *ids, int64_t n,
                                             int accum, void *stream) {
    set_carveouts();
    if (n == 0) return CLMGS_OK;
    if (int rc = check_align(h_grad, "h_grad")) return rc;
    cudaStream_t st = as_stream(stream);
    if (accum) accum_rows_kernel<<<grid_for(n, true), kThreads, 0, st>>>(F4(h_grad), CF4(d_grad), n, Identity{}, Map64{ids});
    else copy_rows_kernel<<<grid_for(n, true), kThreads, 0, st>>>(F4(h_grad), CF4(d_grad), n, Identity{}, Map64{ids});
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_scatter_grad_retention(const float *d_grad, float *h_grad, float *d_next_grad,
                                                   const int32_t *host_ids, const int32_t *next_dst_rows,
                                                   const int32_t *src_rows_to_host, const int32_t *src_rows_to_next,
                                                   int64_t n_host, int64_t n_next, int accum, void *stream) {
    set_carveouts();
    cudaStream_t st = as_stream(stream);
    if (n_host > 0) {
        if (accum) accum_rows_kernel<<<grid_for(n_host), kThreads, 0, st>>>(F4(h_grad), CF4(d_grad), n_host, Map32{src_rows_to_host}, Map32{host_ids});
        else copy_rows_kernel<<<grid_for(n_host), kThreads, 0, st>>>(F4(h_grad), CF4(d_grad), n_host, Map32{src_rows_to_host}, Map32{host_ids});
        CLMGS_LAUNCH_OK();
    }
    if (n_next > 0) {
        if (accum) accum_rows_kernel<<<grid_for(n_next), kThreads, 0, st>>>(F4(d_next_grad), CF4(d_grad), n_next, Map32{src_rows_to_next}, Map32{next_dst_rows});
        else copy_rows_kernel<<<grid_for(n_next), kThreads, 0, st>>>(F4(d_next_grad), CF4(d_grad), n_next, Map32{src_rows_to_next}, Map32{next_dst_rows});
        CLMGS_LAUNCH_OK();
    }
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_set_signal(int32_t *h_flags, int index, int value, void *stream) {
    set_signal_kernel<<<1, 1, 0, as_stream(stream)>>>(h_flags, index, value);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// One 8-byte word device -> pinned host by a 1-thread kernel.  A cudaMemcpyAsync of the same word queues on the
// device-to-host copy engine BEHIND the gradient flush's bulk DMA (measured at configs[3]: 4.5-7.5 ms per camera for
// 8 bytes); a zero-copy store from an SM does not use the copy engine at all.
__global__ void store_i64_kernel(int64_t *h_dst, const int64_t *d_src) {
    *h_dst = *d_src;
    __threadfence_system();
}

CLMGS_EXPORT int clmgs_store_i64_to_host(int64_t *h_dst, const int64_t *d_src, void *stream) {
    CLMGS_REQUIRE(h_dst && d_src, "store_i64_to_host: null pointer");
    store_i64_kernel<<<1, 1, 0, as_stream(stream)>>>(h_dst, d_src);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_prefetch_next(float *d_next_rows, const float *h_rows, const float *d_this_rows,
                                          const int64_t *ids_this, int64_t n_this, const int64_t *ids_next,
                                          int64_t n_next, const void *bitmap, int word_bytes, int bit_this,
                                          int32_t *src_map, void *stream) {
    set_carveouts();
    if (n_next == 0) return CLMGS_OK;
    CLMGS_REQUIRE(word_bytes == 1 || word_bytes == 2 || word_bytes == 4 || word_bytes == 8, "bad word_bytes");
    cudaStream_t st = as_stream(stream);
    const int64_t blocks = ceil_div64(n_next * 4, kThreads);
    const int64_t cap_dev = (int64_t)kNumSMs * 4, cap_host = host_load_ctas();
    prefetch_dev_kernel<<<(unsigned)(blocks > cap_dev ? cap_dev : blocks), kThreads, 0, st>>>(
        F4(d_next_rows), CF4(d_this_rows), ids_this, n_this, ids_next, n_next, bitmap, word_bytes, bit_this, src_map);
    CLMGS_LAUNCH_OK();
    prefetch_host_kernel<<<(unsigned)(blocks > cap_host ? cap_host : blocks), kThreads, 0, st>>>(
        F4(d_next_rows), CF4(h_rows), ids_next, n_next, src_map);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_grad_handoff(const float *d_grad_this, float *d_next_grad, float *h_grad,
                                         const int64_t *ids_this, int64_t n_this, int64_t n_next,
                                         const int32_t *src_map, const void *bitmap, int word_bytes, int bit_this,
                                         int bit_next, int accum, void *stream) {
    set_carveouts();
    int64_t n = n_this > n_next ? n_this : n_next;
    if (n == 0) return CLMGS_OK;
    CLMGS_REQUIRE(word_bytes == 1 || word_bytes == 2 || word_bytes == 4 || word_bytes == 8, "bad word_bytes");
    CLMGS_REQUIRE(accum >= 0 && accum <= 2, "grad_handoff: accum must be 0, 1 or 2");
    cudaStream_t st = as_stream(stream);
    const int64_t cap_dev = (int64_t)kNumSMs * 4, cap_host = host_store_ctas();
    if (n_next > 0) {
        const int64_t blocks = ceil_div64(n_next * 4, kThreads);
        handoff_dev_kernel<<<(unsigned)(blocks > cap_dev ? cap_dev : blocks), kThreads, 0, st>>>(
            CF4(d_grad_this), F4(d_next_grad), n_next, src_map);
        CLMGS_LAUNCH_OK();
    }
    if (n_this > 0) {
        const int64_t blocks = ceil_div64(n_this * 4, kThreads);
        handoff_host_kernel<<<(unsigned)(blocks > cap_host ? cap_host : blocks), kThreads, 0, st>>>(
            CF4(d_grad_this), F4(h_grad), ids_this, n_this, n_next > 0 ? 1 : 0, bitmap, word_bytes, bit_this, bit_next,
            accum);
    }
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_accumulate(float *d_dst_rows, const float *d_src_rows, const int64_t *dst_index,
                                       int64_t n, void *stream) {
    if (n == 0) return CLMGS_OK;
    int64_t blocks = ceil_div64(n * kRowF4, kThreads);
    int64_t cap = (int64_t)kNumSMs * 8;
    accumulate_rows_kernel<<<(unsigned)(blocks > cap ? cap : blocks), kThreads, 0, as_stream(stream)>>>(
        F4(d_dst_rows), CF4(d_src_rows), dst_index, n);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------
// Batch-resident rows ("slots"): every Gaussian the batch touches owns one row of a device table for
// the whole batch, so a row crosses the host link exactly once each way per batch and nothing is
// copied between cameras (the reference re-packs the retained rows for every camera,
// CK/clm_kernels.cu:71-126,206-268).
// ---------------------------------------------------------------------------
__global__ void slot_map_kernel(const int32_t *__restrict__ ids, int64_t n, int32_t *__restrict__ slot_of) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x)
        slot_of[ids[p]] = (int32_t)p;
}

CLMGS_EXPORT int clmgs_slot_map(const int32_t *ids, int64_t n, int32_t *slot_of, void *stream) {
    if (n == 0) return CLMGS_OK;
    int64_t blocks = ceil_div64(n, 256);
    int64_t cap = (int64_t)kNumSMs * 16;
    slot_map_kernel<<<(unsigned)(blocks > cap ? cap : blocks), 256, 0, as_stream(stream)>>>(ids, n, slot_of);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_gather_i32(float *d_rows, const float *h_rows, const int32_t *ids, int64_t n,
                                       void *stream) {
    set_carveouts();
    if (n == 0) return CLMGS_OK;
    if (int rc = check_align(d_rows, "d_rows")) return rc;
    if (int rc = check_align(h_rows, "h_rows")) return rc;
    copy_rows_kernel<<<grid_for(n), kThreads, 0, as_stream(stream)>>>(F4(d_rows), CF4(h_rows), n, Map32{ids}, Identity{});
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_gather_to_slots(float *d_rows, const float *h_rows, const int32_t *ids, int64_t n,
                                            const int32_t *slot_of, void *stream) {
    set_carveouts();
    if (n == 0) return CLMGS_OK;
    if (int rc = check_align(d_rows, "d_rows")) return rc;
    if (int rc = check_align(h_rows, "h_rows")) return rc;
    copy_rows_kernel<<<grid_for(n), host_load_threads(), 0, as_stream(stream)>>>(F4(d_rows), CF4(h_rows), n, Map32{ids},
                                                                                Map32Via{ids, slot_of});
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rows_flush_slots(const float *d_grad_rows, const int32_t *slot_of, const int32_t *ids,
                                        int64_t n, float *h_grad, void *stream) {
    set_carveouts();
    if (n == 0) return CLMGS_OK;
    if (int rc = check_align(h_grad, "h_grad")) return rc;
    copy_rows_kernel<<<grid_for(n, true), kThreads, 0, as_stream(stream)>>>(F4(h_grad), CF4(d_grad_rows), n,
                                                                           Map32Via{ids, slot_of}, Map32{ids});
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
