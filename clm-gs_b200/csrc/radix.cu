// Stable LSD radix sort of (uint32 key, int32 value) pairs, 8-bit digits -- the sort under the tile
// intersection stage (the reference calls cub::DeviceRadixSort on 64-bit keys,
// CSRC/isect_tiles.cu:186-229).  Hand-written so that (a) nothing on the path is a library call and
// (b) the LAST pass can write the stage's real outputs -- isect_ids = (cam|tile) << 32 | depth bits and
// flatten_ids -- directly at their sorted positions instead of a (key, value) pair that a separate kernel
// would have to read back (saves 16 B of the ~52 B per intersection the final pass + finalize moved).
//
// Per pass, three launches over fixed tiles of kTile pairs (no inter-block waiting of any kind):
//   radix_hist_kernel    per-tile digit histogram               -> hist[digit][tile]          (reads 4 B/pair)
//   radix_scan_kernel    one block per digit: exclusive scan over tiles in place, totals[digit]
//   radix_scatter_kernel persistent; tiles stream in through double-buffered cp.async stages; rank (warp match +
//                        per-warp counters), permute in place in shared memory so that each digit's run leaves as
//                        consecutive addresses, write                    (reads 8 B, writes 8 B per pair)
// i.e. 20 B per pair and pass against 16 B (+4 B once) for a chained-scan single-sweep design; with the
// 2 passes the tile sort needs the difference is 10 %, and there is no spin-waiting between blocks.
// HBM-bound: roofline = 20 B x n / pass.
#include "common.cuh"

namespace {

// tunables (tools/radix_sweep.sh builds variants with -D...)
#ifndef CLMGS_RADIX_THREADS
#define CLMGS_RADIX_THREADS 256
#endif
#ifndef CLMGS_RADIX_ITEMS
#define CLMGS_RADIX_ITEMS 16
#endif
#ifndef CLMGS_RADIX_BLOCKS_PER_SM
#define CLMGS_RADIX_BLOCKS_PER_SM 3
#endif
#ifndef CLMGS_RADIX_STAGES
#define CLMGS_RADIX_STAGES 2
#endif
constexpr int kThreads = CLMGS_RADIX_THREADS;
constexpr int kItems = CLMGS_RADIX_ITEMS;
constexpr int kTile = kThreads * kItems;  // 4096 pairs per block
constexpr int kWarps = kThreads / 32;
constexpr int kBins = 256;

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Per-tile digit histogram.  Only the SET of keys of a tile matters here, so full tiles are read as 16-byte
// vectors; per-warp sub-histograms keep same-address shared-memory atomics rare for low-entropy digits.
__global__ void __launch_bounds__(kThreads)
radix_hist_kernel(int64_t n, const uint32_t *__restrict__ keys, int shift, uint32_t digit_mask,
                  uint32_t *__restrict__ hist, int64_t ntiles) {
    __shared__ uint32_t s_hist[kWarps][kBins];
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s_hist[w][threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kTile;
    const int warp = threadIdx.x >> 5;
    uint32_t *mine = s_hist[warp];
    if (base + kTile <= n) {
        const uint4 *src = reinterpret_cast<const uint4 *>(keys + base);
        uint4 k[kItems / 4];
#pragma unroll
        for (int j = 0; j < kItems / 4; ++j) k[j] = src[threadIdx.x + j * kThreads];
#pragma unroll
        for (int j = 0; j < kItems / 4; ++j) {
            atomicAdd(&mine[(k[j].x >> shift) & digit_mask], 1u);
            atomicAdd(&mine[(k[j].y >> shift) & digit_mask], 1u);
            atomicAdd(&mine[(k[j].z >> shift) & digit_mask], 1u);
            atomicAdd(&mine[(k[j].w >> shift) & digit_mask], 1u);
        }
    } else {
        for (int64_t i = base + threadIdx.x; i < n; i += kThreads) atomicAdd(&mine[(keys[i] >> shift) & digit_mask], 1u);
    }
    __syncthreads();
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) c += s_hist[w][threadIdx.x];
    hist[(int64_t)threadIdx.x * ntiles + blockIdx.x] = c;
}

// block d: exclusive scan of hist[d][0..ntiles) in place; totals[d] = sum
__global__ void __launch_bounds__(1024)
radix_scan_kernel(uint32_t *__restrict__ hist, int64_t ntiles, uint32_t *__restrict__ totals) {
    if (blockIdx.x == 0 && threadIdx.x == 0) totals[kBins] = 0;  // tile ticket counter of the scatter kernel
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    uint32_t *row = hist + (int64_t)blockIdx.x * ntiles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t carry = 0;
    for (int64_t b0 = 0; b0 < ntiles; b0 += 1024) {
        int64_t j = b0 + threadIdx.x;
        uint32_t v = j < ntiles ? row[j] : 0, incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
            s_warp[lane] = wi - w;
            if (lane == 31) s_carry = wi;
        }
        __syncthreads();
        if (j < ntiles) row[j] = carry + s_warp[warp] + incl - v;
        carry += s_carry;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

struct FinalOut {  // last pass of the tile sort: write the stage outputs instead of pairs
    int64_t *isect_ids;    // mode 1
    int32_t *flatten_ids;  // modes 1, 2
    const float *depths;   // mode 1
    int32_t *offsets;      // mode 2: [C * n_tiles], pre-filled with n; receives the first position of every key
    int tile_n_bits, n_tiles;
};
// MODE 0: (key, value) pairs.  1: isect_ids + flatten_ids.  2: flatten_ids + tile offsets (what the rasteriser
// consumes; no 8-byte keys are written and no separate offsets pass reads them back).

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

constexpr int kScatterBlocksPerSM = CLMGS_RADIX_BLOCKS_PER_SM;
constexpr size_t kStageBytes = (size_t)kTile * 8;            // keys + values of one tile
constexpr size_t kScatterDynSmem = 2 * kStageBytes;          // double buffered: 64 KB

// Persistent scatter: each block walks tiles blockIdx.x, +gridDim.x, ...  The next tile's keys and values are
// already streaming into the other shared-memory stage (cp.async, no registers) while the current tile is ranked,
// permuted IN PLACE in its stage and written out, so the load latency that dominated a one-tile-per-block version
// (ncu: 50 % of the stall samples were long-scoreboard waits on the tile's own loads) overlaps the rest.
template <int MODE>
__global__ void __launch_bounds__(kThreads, kScatterBlocksPerSM)
radix_scatter_kernel(int64_t n, const uint32_t *__restrict__ keys_in, const int32_t *__restrict__ vals_in,
                     uint32_t *__restrict__ keys_out, int32_t *__restrict__ vals_out, int shift, uint32_t digit_mask,
                     const uint32_t *__restrict__ hist, int64_t ntiles, uint32_t *__restrict__ totals,
                     int aligned16, FinalOut fin) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    __shared__ uint16_t s_cnt[kWarps][kBins];  // per-warp running digit counts, then per-warp digit offsets
    __shared__ uint32_t s_start[kBins];        // first slot of each digit inside the tile
    __shared__ uint32_t s_global[kBins];       // global position of slot 0 of each digit's run, minus its local start
    __shared__ uint32_t s_gbase[kBins];        // first global position of each digit (all tiles)
    __shared__ uint32_t s_scan[kWarps];
    __shared__ uint32_t s_ticket;              // next tile of this block (tiles are handed out dynamically)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int woff = warp * (kItems * 32) + lane;  // this lane's pair in round r is woff + 32 r
    uint32_t *ticket_counter = totals + kBins;

    auto stage_keys = [&](int st) { return reinterpret_cast<uint32_t *>(dyn_smem + st * kStageBytes); };
    auto stage_vals = [&](int st) { return reinterpret_cast<int32_t *>(dyn_smem + st * kStageBytes + (size_t)kTile * 4); };
    auto issue_tile = [&](int64_t tile, int st) {
        const int64_t base = tile * kTile;
        uint32_t *sk = stage_keys(st);
        int32_t *sv = stage_vals(st);
        if (aligned16 && base + kTile <= n) {
#pragma unroll
            for (int j = 0; j < kItems / 4; ++j) {
                const int c = threadIdx.x + j * kThreads;  // 16-byte chunk
                cp_async16(sk + 4 * c, keys_in + base + 4 * c);
                cp_async16(sv + 4 * c, vals_in + base + 4 * c);
            }
        } else {  // last (partial) tile or unaligned arrays: plain loads
            for (int i = threadIdx.x; i < kTile; i += kThreads) {
                const bool ok = base + i < n;
                sk[i] = ok ? keys_in[base + i] : 0u;
                sv[i] = ok ? vals_in[base + i] : 0;
            }
        }
        cp_async_commit();
    };

    {  // global digit bases: exclusive scan of totals[256], once per block
        const uint32_t gtot = totals[threadIdx.x];
        uint32_t gi = gtot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t b = __shfl_up_sync(0xffffffffu, gi, o); if (lane >= o) gi += b; }
        if (lane == 31) s_scan[warp] = gi;
        __syncthreads();
        uint32_t gb = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) if (w < warp) gb += s_scan[w];
        s_gbase[threadIdx.x] = gb + gi - gtot;
        __syncthreads();
    }

    // Tiles are taken from a ticket counter, not by a fixed stride: when part of the grid cannot be resident (the
    // engine runs this next to its host-link transfer kernels) a fixed assignment would leave whole tile sequences
    // to blocks that start late.  Positions come from the histogram prefix, so the assignment does not affect the result.
    if (threadIdx.x == 0) s_ticket = atomicAdd(ticket_counter, 1u);
    __syncthreads();
    int st = 0;
    int64_t tile = s_ticket;
    __syncthreads();
    if (tile < ntiles) issue_tile(tile, 0);
    for (; tile < ntiles; st ^= 1) {
        if (threadIdx.x == 0) s_ticket = atomicAdd(ticket_counter, 1u);
        const uint32_t tile_prefix = hist[(int64_t)threadIdx.x * ntiles + tile];  // used after the ranking
#pragma unroll
        for (int w = 0; w < kWarps; ++w) s_cnt[w][threadIdx.x] = 0;
        cp_async_wait<0>();
        __syncthreads();
        const int64_t next = s_ticket;
        if (next < ntiles) issue_tile(next, st ^ 1);  // streams in while this tile is processed
        const int64_t base = tile * kTile;
        const int count = (int)(n - base < kTile ? n - base : kTile);
        uint32_t *sk = stage_keys(st);
        int32_t *sv = stage_vals(st);
        uint32_t key[kItems];
#pragma unroll
        for (int r = 0; r < kItems; ++r) key[r] = sk[woff + 32 * r];
        // rank inside the warp, round by round (round r holds 32 consecutive pairs): stable by construction
        uint32_t rank2[kItems / 2];  // two 16-bit ranks per register (a tile holds 4096 pairs)
#pragma unroll
        for (int r = 0; r < kItems; ++r) {
            const bool valid = woff + 32 * r < count;
            const uint32_t d = valid ? (key[r] >> shift) & digit_mask : (uint32_t)kBins;  // padding: own group, never counted
            const uint32_t m = __match_any_sync(0xffffffffu, d);
            uint32_t before = 0;
            if (valid) before = s_cnt[warp][d];
            __syncwarp();
            const uint32_t rk = before + __popc(m & lanemask_lt());
            if (valid && (m & lanemask_lt()) == 0) s_cnt[warp][d] = (uint16_t)(before + __popc(m));
            __syncwarp();
            if (r & 1) rank2[r >> 1] |= rk << 16; else rank2[r >> 1] = rk;
        }
        __syncthreads();
        {  // thread t owns digit t: warp counts -> exclusive warp offsets, digit total -> local starts
            uint32_t total = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                uint32_t c = s_cnt[w][threadIdx.x];
                s_cnt[w][threadIdx.x] = (uint16_t)total;
                total += c;
            }
            uint32_t li = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t a = __shfl_up_sync(0xffffffffu, li, o); if (lane >= o) li += a; }
            if (lane == 31) s_scan[warp] = li;
            __syncthreads();
            uint32_t lbase = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) if (w < warp) lbase += s_scan[w];
            const uint32_t lstart = lbase + li - total;
            s_start[threadIdx.x] = lstart;
            s_global[threadIdx.x] = s_gbase[threadIdx.x] + tile_prefix - lstart;
        }
        int32_t val[kItems];
#pragma unroll
        for (int r = 0; r < kItems; ++r) val[r] = sv[woff + 32 * r];
        __syncthreads();  // every key and value of the tile is in registers: the stage can be permuted in place
#pragma unroll
        for (int r = 0; r < kItems; ++r) {
            if (woff + 32 * r < count) {
                const uint32_t d = (key[r] >> shift) & digit_mask;
                const uint32_t rk = (r & 1) ? rank2[r >> 1] >> 16 : rank2[r >> 1] & 0xffffu;
                const uint32_t slot = s_start[d] + s_cnt[warp][d] + rk;
                sk[slot] = key[r];
                sv[slot] = val[r];
            }
        }
        __syncthreads();
#pragma unroll 4
        for (int i = threadIdx.x; i < count; i += kThreads) {
            const uint32_t k = sk[i];
            const int32_t v = sv[i];
            const uint32_t pos = s_global[(k >> shift) & digit_mask] + (uint32_t)i;
            if (MODE == 1) {
                fin.isect_ids[pos] = ((int64_t)k << 32) | (int64_t)(uint32_t)__float_as_int(fin.depths[v]);
                fin.flatten_ids[pos] = v;
            } else if (MODE == 2) {
                fin.flatten_ids[pos] = v;
                // The tile holds a contiguous chunk of a list already sorted on the lower digits, so after the stable
                // permutation equal full keys are adjacent: the first of each run is a candidate for the key's first
                // position overall (other tiles may hold the same key: min over candidates).
                if (i == 0 || sk[i - 1] != k) {
                    const uint32_t cam = k >> fin.tile_n_bits, t = k & ((1u << fin.tile_n_bits) - 1u);
                    atomicMin(fin.offsets + (int64_t)cam * fin.n_tiles + t, (int32_t)pos);
                }
            } else {
                keys_out[pos] = k;
                vals_out[pos] = v;
            }
        }
        __syncthreads();  // stage, counters, digit tables and the ticket are reused by the next iterations
        tile = next;
    }
}

template <int MODE>
int launch_scatter(int64_t n, const uint32_t *keys, const int32_t *vals, uint32_t *keys_out, int32_t *vals_out, int shift,
                   uint32_t mask, const uint32_t *hist, int64_t ntiles, uint32_t *totals, FinalOut fin,
                   cudaStream_t st) {
    static bool configured[64] = {};  // per template instance and device: the attribute is per device function
    int dev = 0;
    CLMGS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        CLMGS_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)kScatterDynSmem));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const int aligned16 = (((uintptr_t)keys | (uintptr_t)vals) & 15) == 0;
    int64_t grid = (int64_t)kNumSMs * kScatterBlocksPerSM;
    if (grid > ntiles) grid = ntiles;
    radix_scatter_kernel<MODE><<<(unsigned)grid, kThreads, kScatterDynSmem, st>>>(n, keys, vals, keys_out, vals_out, shift,
                                                                                  mask, hist, ntiles, totals, aligned16, fin);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

__global__ void __launch_bounds__(256)
fill_i32_kernel(int32_t *__restrict__ p, int64_t n, int32_t v) {
    int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) p[i] = v;
}

// offsets[t] = min(offsets[t], offsets[t+1], ...): keys without entries start where the next key starts.
// One block walks the (<= a few 10^4) slots from the top in chunks of 1024.
__global__ void __launch_bounds__(1024)
suffix_min_kernel(int32_t *__restrict__ offsets, int64_t n_slots) {
    __shared__ int32_t s_warp[32];
    __shared__ int32_t s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t carry = 0x7fffffff;
    for (int64_t hi = n_slots; hi > 0; hi -= 1024) {
        const int64_t j = hi - 1 - threadIdx.x;  // thread 0 takes the top slot of the chunk
        int32_t v = j >= 0 ? offsets[j] : 0x7fffffff;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int32_t t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = min(v, t); }
        if (lane == 31) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            int32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int32_t t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w = min(w, t); }
            s_warp[lane] = w;
            if (lane == 31) s_carry = w;
        }
        __syncthreads();
        if (warp > 0) v = min(v, s_warp[warp - 1]);
        v = min(v, carry);
        if (j >= 0) offsets[j] = v;
        carry = min(carry, s_carry);
        __syncthreads();
    }
}

inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

size_t clmgs_radix_u32_scratch(int64_t n) {
    int64_t ntiles = ceil_div64(n > 0 ? n : 1, kTile);
    return a256((size_t)ntiles * kBins * sizeof(uint32_t)) + a256((kBins + 1) * sizeof(uint32_t));
}

// Sorts on bits [0, end_bit).  Ping-pongs between (keys, vals) and (keys_alt, vals_alt); *keys_out / *vals_out
// say where the result is.  With `isect_ids` set, the last pass writes isect_ids / flatten_ids instead (the
// caller arranges the buffers so that the last pass does not read flatten_ids: clmgs_radix_passes()).
int clmgs_radix_passes(int end_bit) { return end_bit <= 0 ? 0 : (end_bit + 7) / 8; }

int clmgs_radix_sort_u32(int64_t n, uint32_t *keys, uint32_t *keys_alt, int32_t *vals, int32_t *vals_alt, int end_bit,
                         void *scratch, size_t scratch_bytes, cudaStream_t st, uint32_t **keys_out, int32_t **vals_out,
                         int64_t *isect_ids, int32_t *flatten_ids, const float *depths, int32_t *offsets,
                         int64_t n_slots, int tile_n_bits, int n_tiles) {
    CLMGS_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "radix sort: n=%lld out of range", (long long)n);
    CLMGS_REQUIRE(end_bit >= 0 && end_bit <= 32, "radix sort: end_bit=%d", end_bit);
    if (keys_out) *keys_out = keys;
    if (vals_out) *vals_out = vals;
    const int passes = clmgs_radix_passes(end_bit);
    if (n == 0) return CLMGS_OK;
    CLMGS_REQUIRE(scratch && scratch_bytes >= clmgs_radix_u32_scratch(n), "radix sort: scratch too small");
    const int64_t ntiles = ceil_div64(n, kTile);
    uint32_t *hist = (uint32_t *)scratch;
    uint32_t *totals = (uint32_t *)((char *)scratch + a256((size_t)ntiles * kBins * sizeof(uint32_t)));
    FinalOut fin{isect_ids, flatten_ids, depths, offsets, tile_n_bits, n_tiles};
    const int mode = isect_ids ? 1 : (offsets ? 2 : 0);
    CLMGS_REQUIRE(mode == 0 || passes > 0, "radix sort: fused outputs need at least one key bit");
    if (mode == 2) {
        fill_i32_kernel<<<(unsigned)ceil_div64(n_slots, 256), 256, 0, st>>>(offsets, n_slots, (int32_t)n);
        CLMGS_LAUNCH_OK();
    }
    // Balanced digits: 13 key bits go as 7 + 6 rather than 8 + 5.  Fewer bins in the wider pass means longer runs per
    // bin and tile (4096 / 128 = 32 pairs = 128 B instead of 64 B), i.e. fewer partially written 32-byte sectors.
    const int base_bits = passes ? end_bit / passes : 0, extra = passes ? end_bit % passes : 0;
    int shift = 0;
    for (int p = 0; p < passes; ++p) {
        const int bits = base_bits + (p < extra ? 1 : 0);
        const uint32_t mask = (1u << bits) - 1u;
        const bool last = p == passes - 1;
        radix_hist_kernel<<<(unsigned)ntiles, kThreads, 0, st>>>(n, keys, shift, mask, hist, ntiles);
        CLMGS_LAUNCH_OK();
        radix_scan_kernel<<<kBins, 1024, 0, st>>>(hist, ntiles, totals);
        CLMGS_LAUNCH_OK();
        if (last && mode != 0) {
            CLMGS_REQUIRE(vals != flatten_ids, "radix sort: final pass would read the buffer it writes");
            if (mode == 1) {
                if (int rc = launch_scatter<1>(n, keys, vals, nullptr, nullptr, shift, mask, hist, ntiles, totals, fin, st)) return rc;
            } else {
                if (int rc = launch_scatter<2>(n, keys, vals, nullptr, nullptr, shift, mask, hist, ntiles, totals, fin, st)) return rc;
                suffix_min_kernel<<<1, 1024, 0, st>>>(offsets, n_slots);
                CLMGS_LAUNCH_OK();
            }
        } else {
            if (int rc = launch_scatter<0>(n, keys, vals, keys_alt, vals_alt, shift, mask, hist, ntiles, totals, fin, st)) return rc;
        }
        uint32_t *tk = keys; keys = keys_alt; keys_alt = tk;
        int32_t *tv = vals; vals = vals_alt; vals_alt = tv;
        shift += bits;
    }
    if (keys_out) *keys_out = keys;
    if (vals_out) *vals_out = vals;
    return CLMGS_OK;
}

// ---- C ABI: the sort on its own (tests, and hosts that want the pairs) ------------------------------------------
CLMGS_EXPORT size_t clmgs_radix_sort_pairs_u32_scratch_bytes(int64_t n) { return clmgs_radix_u32_scratch(n); }

CLMGS_EXPORT int clmgs_radix_sort_pairs_u32(int64_t n, uint32_t *keys, uint32_t *keys_alt, int32_t *vals,
                                            int32_t *vals_alt, int end_bit, void *scratch, size_t scratch_bytes,
                                            int32_t *result_in_alt, void *stream) {
    CLMGS_REQUIRE(n == 0 || (keys && keys_alt && vals && vals_alt), "radix_sort_pairs_u32: null pointer");
    uint32_t *ko = keys;
    int32_t *vo = vals;
    int rc = clmgs_radix_sort_u32(n, keys, keys_alt, vals, vals_alt, end_bit, scratch, scratch_bytes, as_stream(stream),
                                  &ko, &vo, nullptr, nullptr, nullptr, nullptr, 0, 0, 0);
    if (rc) return rc;
    if (result_in_alt) *result_in_alt = (ko == keys_alt) ? 1 : 0;
    return CLMGS_OK;
}
