// Stable LSD radix sort of (int64 key, int32 value) pairs over key bits
// [0, end_bit) -- the semantics isect_tiles relies on
// (CSRC/isect_tiles.cu:196-203).  Round 1 calls the toolkit's
// cub::DeviceRadixSort with caller-owned scratch; any stable sort over the same
// bit range gives a bit-identical permutation, so this can be replaced by a
// hand-written onesweep without touching parity.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t cub_temp_bytes(int64_t n) {
    size_t tmp = 0;
    cub::DoubleBuffer<int64_t> k(nullptr, nullptr);
    cub::DoubleBuffer<int32_t> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, k, v, n, 0, 64, (cudaStream_t)0);
    return tmp;
}

// scratch layout: alt keys [n] | alt vals [n] | cub temp
size_t clmgs_radix_sort_scratch(int64_t n) {
    if (n <= 0) return 0;
    return align256(sizeof(int64_t) * (size_t)n) + align256(sizeof(int32_t) * (size_t)n) + align256(cub_temp_bytes(n));
}

int clmgs_radix_sort_pairs(int64_t n, int64_t *keys, int32_t *vals, int end_bit, void *scratch,
                           size_t scratch_bytes, cudaStream_t st) {
    if (n <= 1) return CLMGS_OK;
    CLMGS_REQUIRE(n < ((int64_t)1 << 31), "radix_sort: n too large");
    if (scratch == nullptr || scratch_bytes < clmgs_radix_sort_scratch(n)) {
        clmgs_set_error("radix_sort: scratch too small (%zu < %zu)", scratch_bytes, clmgs_radix_sort_scratch(n));
        return CLMGS_ERR_SCRATCH;
    }
    char *base = (char *)scratch;
    int64_t *k_alt = (int64_t *)base;
    int32_t *v_alt = (int32_t *)(base + align256(sizeof(int64_t) * (size_t)n));
    void *tmp = base + align256(sizeof(int64_t) * (size_t)n) + align256(sizeof(int32_t) * (size_t)n);
    size_t tmp_bytes = cub_temp_bytes(n);
    cub::DoubleBuffer<int64_t> dk(keys, k_alt);
    cub::DoubleBuffer<int32_t> dv(vals, v_alt);
    CLMGS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (int)n, 0, end_bit, st));
    if (dk.Current() != keys)
        CLMGS_CUDA(cudaMemcpyAsync(keys, dk.Current(), sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    if (dv.Current() != vals)
        CLMGS_CUDA(cudaMemcpyAsync(vals, dv.Current(), sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    return CLMGS_OK;
}

// ---- 32-bit key variant used by the two-level tile sort (proj.cu) -----------------------------
size_t clmgs_sort_u32_temp_bytes(int64_t n) {
    if (n <= 0) return 0;
    size_t tmp = 0;
    cub::DoubleBuffer<uint32_t> k(nullptr, nullptr);
    cub::DoubleBuffer<int32_t> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, k, v, n, 0, 32, (cudaStream_t)0);
    return align256(tmp);
}

// Sorts (keys, vals) over key bits [0, end_bit); *keys_out / *vals_out tell which of the two
// buffers holds the result (no copy back).
int clmgs_sort_pairs_u32(int64_t n, uint32_t *keys, uint32_t *keys_alt, int32_t *vals, int32_t *vals_alt, int end_bit,
                         void *tmp, size_t tmp_bytes, cudaStream_t st, uint32_t **keys_out, int32_t **vals_out) {
    *keys_out = keys;
    *vals_out = vals;
    if (n <= 1 || end_bit <= 0) return CLMGS_OK;
    cub::DoubleBuffer<uint32_t> dk(keys, keys_alt);
    cub::DoubleBuffer<int32_t> dv(vals, vals_alt);
    CLMGS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (int)n, 0, end_bit, st));
    *keys_out = dk.Current();
    *vals_out = dv.Current();
    return CLMGS_OK;
}
