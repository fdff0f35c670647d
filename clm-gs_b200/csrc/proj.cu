// S1 frustum cull + compaction, A4 gather/activate, S3 dense projection,
// S4 tile counting / key emission / offsets.  Compiled with -fmad=false (see
// gs_math.cuh): these kernels make the integer decisions that must be
// bit-exact against the oracle.
#include "common.cuh"
#include "gs_math.cuh"
#include <stdlib.h>

using namespace gsm;

constexpr int kCullThreads = 256;
constexpr int kMaxCams = 64;

struct CullParams {
    int width, height;
    float eps2d, near_plane, far_plane, radius_clip;
};

// Shared-memory camera table: 16 floats per camera (R row-major, t, fx fy cx cy).
__device__ __forceinline__ void stage_cameras(float *s_cam, const float *viewmats, const float *Ks, int B) {
    for (int i = threadIdx.x; i < B * 16; i += blockDim.x) {
        int c = i >> 4, k = i & 15;
        float v;
        if (k < 9) v = viewmats[c * 16 + (k / 3) * 4 + (k % 3)];
        else if (k < 12) v = viewmats[c * 16 + (k - 9) * 4 + 3];
        else if (k == 12) v = Ks[c * 9 + 0];
        else if (k == 13) v = Ks[c * 9 + 4];
        else if (k == 14) v = Ks[c * 9 + 2];
        else v = Ks[c * 9 + 5];
        s_cam[i] = v;
    }
}
__device__ __forceinline__ void camera_from_smem(const float *s, Camera &cam) {
#pragma unroll
    for (int k = 0; k < 9; ++k) cam.R[k] = s[k];
    cam.t[0] = s[9]; cam.t[1] = s[10]; cam.t[2] = s[11];
    cam.fx = s[12]; cam.fy = s[13]; cam.cx = s[14]; cam.cy = s[15];
}

__device__ __forceinline__ void load_gaussian(const float *means, const float *scales, const float *quats,
                                              int activated, int64_t i, float mean[3], float covar[9],
                                              float *smax2 = nullptr) {
    float s[3], q[4];
#pragma unroll
    for (int k = 0; k < 3; ++k) { mean[k] = __ldg(means + 3 * i + k); s[k] = __ldg(scales + 3 * i + k); }
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = __ldg(quats + 4 * i + k);
    if (!activated) {
#pragma unroll
        for (int k = 0; k < 3; ++k) s[k] = det_expf(s[k]);
        float qn[4];
        normalize_quat(q, qn);
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = qn[k];
    }
    quat_scale_to_covar(q, s, covar);
    if (smax2) { float m = fmaxf(fmaxf(fabsf(s[0]), fabsf(s[1])), fabsf(s[2])); *smax2 = m * m; }
}

// Cheap sufficient condition for project_one<true>() == false, evaluated before the covariance is carried into the
// camera frame (most (camera, Gaussian) pairs of a large scene fail here: ~98 % at C4).  It only ever rejects what
// the full test rejects, so the masks stay bit-exact; it is itself NOT exact arithmetic (approximate reciprocal /
// square root, fused multiply-adds) and pays for that with explicit slack:
//  * depth planes: the depth is computed with project_one's own operations, so the comparison is the same;
//  * disc test with an upper bound on the radius.  lambda_max(J cov_c J^T + eps I) <= |J|_F^2 smax^2 + eps, and with
//    |tx / z| <= lim_x:  |J|_F^2 <= (fx^2 (1 + lim_x^2) + fy^2 (1 + lim_y^2)) / z^2  -- a per-camera constant over z^2
//    (the view rotation is taken as a similarity: |R|_F^2 / 3 = its squared scale).  The packed radius rule uses at
//    most lambda_max + sqrt(0.1), radius = ceil(3 sqrt(.)) <= 3 sqrt(.) + 1; 2 % + 0.07 px cover the rounding;
//  * the image position is one fused dot product per axis with the per-camera rows  fx R0 + cx R2,  fy R1 + cy R2
//    divided by the depth.  It differs from project_one's mean2d by rounding only: at most
//    2^-19 (f + |c|)(max|R| |mean|_1 + max|t|) / z  +  1e-5 |mean2d|, which is added to the radius bound.
// Non-finite values make every comparison false and fall through to the full test.
struct PreCam {      // 16 floats per camera in shared memory
    float4 zrow;     // R[2][0..2], t[2]
    float4 xrow;     // fx R[0] + cx R[2],  fx t[0] + cx t[2]
    float4 yrow;     // fy R[1] + cy R[2],  fy t[1] + cy t[2]
    float4 k;        // x: radius constant, y / z: rounding allowance per unit |mean|_1 / constant, w: unused
};
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rsqrt_approx(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ void make_precam(const float *vm, const float *K, int width, int height, PreCam &o) {
    const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    const float lim_x = 1.3f * (0.5f * (float)width / fx), lim_y = 1.3f * (0.5f * (float)height / fy);
    float rs = 0.f, rmax = 0.f, tmax = 0.f;
    for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < 3; ++k) { rs += vm[r * 4 + k] * vm[r * 4 + k]; rmax = fmaxf(rmax, fabsf(vm[r * 4 + k])); }
        tmax = fmaxf(tmax, fabsf(vm[r * 4 + 3]));
    }
    o.zrow = make_float4(vm[8], vm[9], vm[10], vm[11]);
    o.xrow = make_float4(fx * vm[0] + cx * vm[8], fx * vm[1] + cx * vm[9], fx * vm[2] + cx * vm[10], fx * vm[3] + cx * vm[11]);
    o.yrow = make_float4(fy * vm[4] + cy * vm[8], fy * vm[5] + cy * vm[9], fy * vm[6] + cy * vm[10], fy * vm[7] + cy * vm[11]);
    const float g = 1.9073486e-6f * (fmaxf(fabsf(fx), fabsf(fy)) + fmaxf(fabsf(cx), fabsf(cy)));  // 2^-19 (f + |c|)
    o.k = make_float4(1.001f * (rs * (1.0f / 3.0f)) * (fx * fx * (1.f + lim_x * lim_x) + fy * fy * (1.f + lim_y * lim_y)),
                      g * rmax, g * tmax, 0.f);
}

__device__ __forceinline__ bool surely_culled(const float mean[3], float smax2, float mean_l1, const PreCam *pc,
                                              float fwidth, float fheight, float c0, float near_plane,
                                              float far_plane) {
    const float4 zr = pc->zrow;
    const float z = ((zr.x * mean[0] + zr.y * mean[1]) + zr.z * mean[2]) + zr.w;  // = project_one's depth, same operations
    if (z < near_plane || z > far_plane) return true;
    const float4 xr = pc->xrow, yr = pc->yrow, k = pc->k;
    const float rz = rcp_approx(z);
    const float u = __fmaf_rn(k.x * smax2, rz * rz, c0);              // >= eps2d + 0.32 > 0
    const float Rb = __fmaf_rn(3.06f, u * rsqrt_approx(u), 1.07f);
    const float mx = __fmaf_rn(xr.x, mean[0], __fmaf_rn(xr.y, mean[1], __fmaf_rn(xr.z, mean[2], xr.w))) * rz;
    const float my = __fmaf_rn(yr.x, mean[0], __fmaf_rn(yr.y, mean[1], __fmaf_rn(yr.z, mean[2], yr.w))) * rz;
    const float Rr = __fmaf_rn(__fmaf_rn(mean_l1, k.y, k.z), rz, Rb);
    const float Rx = __fmaf_rn(1e-5f, fabsf(mx), Rr), Ry = __fmaf_rn(1e-5f, fabsf(my), Rr);
    return mx + Rx <= 0.f || mx - Rx >= fwidth || my + Ry <= 0.f || my - Ry >= fheight;
}

// ---------------------------------------------------------------------------
// S1.  One thread per Gaussian, all B cameras: xyz/scale/quat are read once
// (40 B/Gaussian) instead of 2*B times as in the reference's two-pass kernel.
// PRETEST (B > 8): pass 1 runs the rejection above for every camera (no divergence) and leaves each lane a bit
// list of candidate cameras; pass 2 runs the reference's full test on each lane's own candidates (pooling the
// warp's candidates and dealing them out evenly was tried: 3 % faster at configs[3], 25 % slower on dense scenes).
// ---------------------------------------------------------------------------
template <typename W, bool PRETEST>
__global__ void __launch_bounds__(kCullThreads)
cull_count_kernel(const float *__restrict__ means, const float *__restrict__ scales,
                  const float *__restrict__ quats, int activated, int64_t N,
                  const float *__restrict__ viewmats, const float *__restrict__ Ks, int B, CullParams p,
                  W *__restrict__ bitmap, int32_t *__restrict__ block_counts /* [B][gridDim.x] */) {
    __shared__ __align__(16) float s_cam[kMaxCams * 16];
    __shared__ int s_cnt[kMaxCams];
    __shared__ __align__(16) PreCam s_pre[PRETEST ? kMaxCams : 1];
    stage_cameras(s_cam, viewmats, Ks, B);
    if (threadIdx.x < kMaxCams) s_cnt[threadIdx.x] = 0;
    if (PRETEST && threadIdx.x < B)
        make_precam(viewmats + threadIdx.x * 16, Ks + threadIdx.x * 9, p.width, p.height, s_pre[threadIdx.x]);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t i = (int64_t)blockIdx.x * kCullThreads + threadIdx.x;
    uint64_t word = 0;
    if (!PRETEST) {
        if (i < N) {
            float mean[3], covar[9];
            load_gaussian(means, scales, quats, activated, i, mean, covar);
            for (int c = 0; c < B; ++c) {
                Camera cam;
                camera_from_smem(s_cam + c * 16, cam);
                Projected o;
                bool ok = project_one<true>(mean, covar, cam, p.width, p.height, p.eps2d, p.near_plane,
                                            p.far_plane, p.radius_clip, o);
                word |= (uint64_t)ok << c;
            }
            bitmap[i] = (W)word;
        }
    } else {
        float mean[3], covar[9], smax2 = 0.f;
        uint64_t cand = 0;
        if (i < N) {
            load_gaussian(means, scales, quats, activated, i, mean, covar, &smax2);
            const float l1 = (fabsf(mean[0]) + fabsf(mean[1])) + fabsf(mean[2]);
            const float fw = (float)p.width, fh = (float)p.height, c0 = p.eps2d + 0.32f;
            for (int cb = 0; cb < B; cb += 32) {   // 32 cameras per mask word: no 64-bit shifts in the loop
                unsigned m = 0;
                const int nb = min(32, B - cb);
                for (int j = 0; j < nb; ++j)
                    if (!surely_culled(mean, smax2, l1, &s_pre[cb + j], fw, fh, c0, p.near_plane, p.far_plane)) m |= 1u << j;
                cand |= (uint64_t)m << cb;
            }
        }
        while (cand) {
            const int c = __ffsll((long long)cand) - 1;
            cand &= cand - 1;
            Camera cam;
            camera_from_smem(s_cam + c * 16, cam);
            Projected o;
            bool ok = project_one<true>(mean, covar, cam, p.width, p.height, p.eps2d, p.near_plane,
                                        p.far_plane, p.radius_clip, o);
            word |= (uint64_t)ok << c;
        }
        if (i < N) bitmap[i] = (W)word;
    }
    // per-camera block totals: warp ballot + one shared atomic per warp, only for the cameras some lane is visible to
    unsigned any_lo = __reduce_or_sync(0xffffffffu, (unsigned)word), any_hi = __reduce_or_sync(0xffffffffu, (unsigned)(word >> 32));
    while (any_lo | any_hi) {
        int c;
        if (any_lo) { c = __ffs(any_lo) - 1; any_lo &= any_lo - 1; }
        else { c = 32 + __ffs(any_hi) - 1; any_hi &= any_hi - 1; }
        const unsigned m = __ballot_sync(0xffffffffu, (word >> c) & 1);
        if (lane == 0) atomicAdd(&s_cnt[c], __popc(m));
    }
    __syncthreads();
    if (threadIdx.x < B) block_counts[(int64_t)threadIdx.x * gridDim.x + blockIdx.x] = s_cnt[threadIdx.x];
}

// Block-wide exclusive scan of one value per thread (1024 threads); returns
// the exclusive prefix, *total gets the block total.  s_warp: 33 entries.
template <typename T>
__device__ __forceinline__ T block_excl_scan_1024(T v, T *s_warp, T *total) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        s_warp[lane] = wi - w;
        if (lane == 31) s_warp[32] = wi;
    }
    __syncthreads();
    T excl = s_warp[warp] + (incl - v);
    *total = s_warp[32];
    __syncthreads();  // s_warp may be reused by the caller's next round
    return excl;
}

// One block per camera: exclusive scan of its block counts in place; total -> counts[c].
__global__ void __launch_bounds__(1024)
cull_scan_kernel(int32_t *__restrict__ block_counts, int64_t nblocks, int64_t *__restrict__ counts_a,
                 int64_t *__restrict__ counts_b) {
    __shared__ int64_t s_warp[33];
    int32_t *row = block_counts + (int64_t)blockIdx.x * nblocks;
    int64_t carry = 0;
    for (int64_t base = 0; base < nblocks; base += 1024) {
        int64_t j = base + threadIdx.x;
        int64_t v = j < nblocks ? row[j] : 0, total;
        int64_t excl = block_excl_scan_1024<int64_t>(v, s_warp, &total);
        if (j < nblocks) row[j] = (int32_t)(carry + excl);
        carry += total;
    }
    if (threadIdx.x == 0) {
        counts_a[blockIdx.x] = carry;
        counts_b[blockIdx.x] = carry;
    }
}

// s_base[c] = counts[0] + .. + counts[c-1] for c < kMaxCams, by the first two warps (one load per thread: a serial
// loop in thread 0 is 64 dependent global loads at the head of every block).  Caller syncs before reading.
__device__ __forceinline__ void camera_bases(const int64_t *__restrict__ counts, int B, int64_t *s_base) {
    static_assert(kMaxCams == 64, "camera_bases: two warps");
    __shared__ int64_t s_first_half;
    const int t = threadIdx.x, lane = t & 31;
    int64_t v = 0, incl = 0;
    if (t < kMaxCams) {
        v = t < B ? counts[t] : 0;
        incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int64_t u = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += u;
        }
        if (t == 31) s_first_half = incl;
    }
    __syncthreads();
    if (t < kMaxCams) s_base[t] = (incl - v) + (t >= 32 ? s_first_half : 0);
}

// Ids only (what the training engine asks for).  The block's bitmap words are transposed into per-(camera, warp)
// ballots in shared memory (only for the cameras some lane of the warp is visible to).  Sparse blocks (large batches
// over large scenes: 1-2 % of the pairs set): thread (camera c, word w) writes the ids of its 32-Gaussian word --
// eight neighbouring threads fill one camera's consecutive positions -- instead of every warp walking all B cameras
// twice.  Dense blocks (more than 4 cameras per Gaussian on average): every thread writes its own Gaussian's bits,
// ranked by the same ballots.
template <typename W>
__global__ void __launch_bounds__(kCullThreads)
cull_emit_ids_kernel(int64_t N, int B, const W *__restrict__ bitmap, const int32_t *__restrict__ block_offsets,
                     const int64_t *__restrict__ counts, int64_t *__restrict__ gaussian_ids,
                     int64_t *__restrict__ camera_ids) {
    constexpr int kWarps = kCullThreads / 32;
    static_assert(kWarps == 8, "cull_emit_ids: 8 ballot words per camera");
    __shared__ int64_t s_base[kMaxCams];
    __shared__ unsigned s_bal[kMaxCams][kWarps];
    __shared__ int s_pref[kMaxCams][kWarps];   // dense path: set bits of camera c in the warps before w
    __shared__ int s_total;
    camera_bases(counts, B, s_base);   // syncs once
    for (int k = threadIdx.x; k < kMaxCams * kWarps; k += kCullThreads) (&s_bal[0][0])[k] = 0;
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * kCullThreads + threadIdx.x;
    const uint64_t word = i < N ? (uint64_t)bitmap[i] : 0;
    unsigned any_lo = __reduce_or_sync(0xffffffffu, (unsigned)word), any_hi = __reduce_or_sync(0xffffffffu, (unsigned)(word >> 32));
    int warp_bits = 0;
    while (any_lo | any_hi) {
        int c;
        if (any_lo) { c = __ffs(any_lo) - 1; any_lo &= any_lo - 1; }
        else { c = 32 + __ffs(any_hi) - 1; any_hi &= any_hi - 1; }
        const unsigned m = __ballot_sync(0xffffffffu, (word >> c) & 1);
        if (lane == 0) { s_bal[c][warp] = m; warp_bits += __popc(m); }
    }
    if (lane == 0 && warp_bits) atomicAdd(&s_total, warp_bits);
    __syncthreads();
    const bool dense = s_total > 4 * kCullThreads;   // block-uniform
    const int w = threadIdx.x & (kWarps - 1);
    for (int cb = 0; cb < B; cb += kCullThreads / kWarps) {
        const int c = cb + (threadIdx.x >> 3);
        unsigned m = c < B ? s_bal[c][w] : 0;
        // set bits of this camera's words before word w (prefix over the 8 lanes that share the camera)
        const int n = __popc(m);
        int incl = n;
#pragma unroll
        for (int d = 1; d < kWarps; d <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, incl, d, kWarps);
            if (w >= d) incl += u;
        }
        if (dense) {
            if (c < B) s_pref[c][w] = incl - n;
            continue;
        }
        if (!m) continue;
        int64_t pos = s_base[c] + block_offsets[(int64_t)c * gridDim.x + blockIdx.x] + (incl - n);
        const int64_t id0 = (int64_t)blockIdx.x * kCullThreads + w * 32;
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            gaussian_ids[pos] = id0 + b;
            if (camera_ids) camera_ids[pos] = c;
            ++pos;
        }
    }
    if (!dense) return;
    __syncthreads();
    const unsigned lt_mask = (1u << lane) - 1u;
    for (uint64_t wd = word; wd; wd &= wd - 1) {
        const int c = __ffsll((long long)wd) - 1;
        const int r = s_pref[c][warp] + __popc(s_bal[c][warp] & lt_mask);
        const int64_t pos = s_base[c] + block_offsets[(int64_t)c * gridDim.x + blockIdx.x] + r;
        gaussian_ids[pos] = i;
        if (camera_ids) camera_ids[pos] = c;
    }
}

// Compaction from the bitmap: camera c's ascending ids land at
// [sum(counts[<c]) + block offset + rank in block).
template <typename W>
__global__ void __launch_bounds__(kCullThreads)
cull_emit_kernel(const float *__restrict__ means, const float *__restrict__ scales,
                 const float *__restrict__ quats, int activated, int64_t N,
                 const float *__restrict__ viewmats, const float *__restrict__ Ks, int B, CullParams p,
                 const W *__restrict__ bitmap, const int32_t *__restrict__ block_offsets,
                 const int64_t *__restrict__ counts, int64_t *__restrict__ gaussian_ids,
                 int64_t *__restrict__ camera_ids, int32_t *__restrict__ radii, float *__restrict__ means2d,
                 float *__restrict__ depths, float *__restrict__ conics) {
    __shared__ float s_cam[kMaxCams * 16];
    __shared__ int64_t s_base[kMaxCams];
    __shared__ int s_wcnt[kMaxCams][kCullThreads / 32];
    const bool extra = radii != nullptr;
    if (extra) stage_cameras(s_cam, viewmats, Ks, B);
    camera_bases(counts, B, s_base);
    int64_t i = (int64_t)blockIdx.x * kCullThreads + threadIdx.x;
    uint64_t word = i < N ? (uint64_t)bitmap[i] : 0;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = 0; c < B; ++c) {
        unsigned m = __ballot_sync(0xffffffffu, (word >> c) & 1);
        if (lane == 0) s_wcnt[c][warp] = __popc(m);
    }
    __syncthreads();
    float mean[3], covar[9];
    if (extra && word) load_gaussian(means, scales, quats, activated, i, mean, covar);
    const unsigned lt_mask = (1u << lane) - 1u;
    for (int c = 0; c < B; ++c) {
        bool bit = (word >> c) & 1;
        unsigned m = __ballot_sync(0xffffffffu, bit);
        if (!bit) continue;
        int r = __popc(m & lt_mask);
        for (int w = 0; w < warp; ++w) r += s_wcnt[c][w];
        int64_t pos = s_base[c] + block_offsets[(int64_t)c * gridDim.x + blockIdx.x] + r;
        gaussian_ids[pos] = i;
        if (camera_ids) camera_ids[pos] = c;
        if (extra) {
            Camera cam;
            camera_from_smem(s_cam + c * 16, cam);
            Projected o;
            project_one<true>(mean, covar, cam, p.width, p.height, p.eps2d, p.near_plane, p.far_plane,
                              p.radius_clip, o);
            radii[pos] = (int32_t)o.radius;
            means2d[2 * pos] = o.mean2d[0];
            means2d[2 * pos + 1] = o.mean2d[1];
            depths[pos] = o.depth;
            conics[3 * pos] = o.conic[0];
            conics[3 * pos + 1] = o.conic[1];
            conics[3 * pos + 2] = o.conic[2];
        }
    }
}

// scratch layout: int64 counts[64] | int32 block_offsets[B][nblocks]
constexpr size_t kCullHead = kMaxCams * sizeof(int64_t);

CLMGS_EXPORT size_t clmgs_cull_scratch_bytes(int64_t N, int B) {
    int64_t nblocks = ceil_div64(N > 0 ? N : 1, kCullThreads);
    return kCullHead + (size_t)(nblocks * B) * sizeof(int32_t);
}

static int check_cull_args(int B, int word_bytes, int64_t N) {
    CLMGS_REQUIRE(B >= 1 && B <= kMaxCams, "cull: B=%d must be in [1,64]", B);
    CLMGS_REQUIRE(word_bytes == 1 || word_bytes == 2 || word_bytes == 4 || word_bytes == 8,
                  "cull: word_bytes=%d must be 1,2,4,8", word_bytes);
    CLMGS_REQUIRE(8 * word_bytes >= B, "cull: %d-byte bitmap words cannot hold %d cameras", word_bytes, B);
    CLMGS_REQUIRE(N >= 0 && ceil_div64(N, kCullThreads) < ((int64_t)1 << 31), "cull: bad N");
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_cull_count(const float *means, const float *scales, const float *quats, int activated,
                                  int64_t N, const float *viewmats, const float *Ks, int B, int width, int height,
                                  float eps2d, float near_plane, float far_plane, float radius_clip, void *bitmap,
                                  int word_bytes, int64_t *counts, void *scratch, size_t scratch_bytes,
                                  void *stream) {
    if (int rc = check_cull_args(B, word_bytes, N)) return rc;
    cudaStream_t st = as_stream(stream);
    if (scratch_bytes < clmgs_cull_scratch_bytes(N, B)) {
        clmgs_set_error("cull: scratch too small (%zu < %zu)", scratch_bytes, clmgs_cull_scratch_bytes(N, B));
        return CLMGS_ERR_SCRATCH;
    }
    int64_t *head = (int64_t *)scratch;
    int32_t *blk = (int32_t *)((char *)scratch + kCullHead);
    if (N == 0) {
        CLMGS_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * B, st));
        CLMGS_CUDA(cudaMemsetAsync(head, 0, kCullHead, st));
        return CLMGS_OK;
    }
    CullParams p{width, height, eps2d, near_plane, far_plane, radius_clip};
    unsigned nblocks = (unsigned)ceil_div64(N, kCullThreads);
    // The rejection pass pays off when most (camera, Gaussian) pairs fail (large batches over large scenes: 1-2 %
    // visible at C4); with a handful of cameras that each see most of the scene it only adds work (C2: +25 %).
#define LAUNCH(W)                                                                                                  \
    do {                                                                                                           \
        if (B > 8)                                                                                                 \
            cull_count_kernel<W, true><<<nblocks, kCullThreads, 0, st>>>(means, scales, quats, activated, N,       \
                                                                        viewmats, Ks, B, p, (W *)bitmap, blk);    \
        else                                                                                                       \
            cull_count_kernel<W, false><<<nblocks, kCullThreads, 0, st>>>(means, scales, quats, activated, N,      \
                                                                         viewmats, Ks, B, p, (W *)bitmap, blk);   \
    } while (0)
    switch (word_bytes) {
        case 1: LAUNCH(uint8_t); break;
        case 2: LAUNCH(uint16_t); break;
        case 4: LAUNCH(uint32_t); break;
        default: LAUNCH(uint64_t); break;
    }
#undef LAUNCH
    CLMGS_LAUNCH_OK();
    cull_scan_kernel<<<B, 1024, 0, st>>>(blk, (int64_t)nblocks, head, counts);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_cull_emit(const float *means, const float *scales, const float *quats, int activated,
                                 int64_t N, const float *viewmats, const float *Ks, int B, int width, int height,
                                 float eps2d, float near_plane, float far_plane, float radius_clip,
                                 const void *bitmap, int word_bytes, const void *scratch, int64_t *gaussian_ids,
                                 int64_t *camera_ids, int32_t *radii, float *means2d, float *depths, float *conics,
                                 void *stream) {
    if (int rc = check_cull_args(B, word_bytes, N)) return rc;
    CLMGS_REQUIRE((radii == nullptr) == (means2d == nullptr) && (radii == nullptr) == (depths == nullptr) &&
                      (radii == nullptr) == (conics == nullptr),
                  "cull_emit: radii/means2d/depths/conics must be all NULL or all set");
    if (N == 0) return CLMGS_OK;
    CullParams p{width, height, eps2d, near_plane, far_plane, radius_clip};
    cudaStream_t st = as_stream(stream);
    unsigned nblocks = (unsigned)ceil_div64(N, kCullThreads);
    const int64_t *head = (const int64_t *)scratch;
    const int32_t *blk = (const int32_t *)((const char *)scratch + kCullHead);
#define LAUNCH(W)                                                                                              \
    do {                                                                                                       \
        if (radii == nullptr)                                                                                  \
            cull_emit_ids_kernel<W><<<nblocks, kCullThreads, 0, st>>>(N, B, (const W *)bitmap, blk, head,      \
                                                                     gaussian_ids, camera_ids);               \
        else                                                                                                   \
            cull_emit_kernel<W><<<nblocks, kCullThreads, 0, st>>>(means, scales, quats, activated, N, viewmats, \
                                                                 Ks, B, p, (const W *)bitmap, blk, head,       \
                                                                 gaussian_ids, camera_ids, radii, means2d,     \
                                                                 depths, conics);                              \
    } while (0)
    switch (word_bytes) {
        case 1: LAUNCH(uint8_t); break;
        case 2: LAUNCH(uint16_t); break;
        case 4: LAUNCH(uint32_t); break;
        default: LAUNCH(uint64_t); break;
    }
#undef LAUNCH
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------
// A4.  Filter gather + activations (engine.py:650-666).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_activate_kernel(const float *__restrict__ xyz, const float *__restrict__ log_scale,
                       const float *__restrict__ raw_quat, const float *__restrict__ opacity_logit,
                       const int64_t *__restrict__ ids, int64_t n, float *__restrict__ means_out,
                       float *__restrict__ scales_out, float *__restrict__ quats_out,
                       float *__restrict__ opac_out) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    int64_t g = ids[r];
    float q[4], qn[4];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        means_out[3 * r + k] = __ldg(xyz + 3 * g + k);
        scales_out[3 * r + k] = det_expf(__ldg(log_scale + 3 * g + k));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = __ldg(raw_quat + 4 * g + k);
    normalize_quat(q, qn);
#pragma unroll
    for (int k = 0; k < 4; ++k) quats_out[4 * r + k] = qn[k];
    opac_out[r] = det_sigmoidf(__ldg(opacity_logit + g));
}

CLMGS_EXPORT int clmgs_gather_activate(const float *xyz, const float *log_scale, const float *raw_quat,
                                       const float *opacity_logit, const int64_t *ids, int64_t n,
                                       float *means_out, float *scales_out, float *quats_out,
                                       float *opacities_out, void *stream) {
    if (n == 0) return CLMGS_OK;
    gather_activate_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, as_stream(stream)>>>(
        xyz, log_scale, raw_quat, opacity_logit, ids, n, means_out, scales_out, quats_out, opacities_out);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------
// S3.  Dense projection, one thread per (camera, Gaussian).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
projection_fwd_kernel(int C, int64_t N, const float *__restrict__ means, const float *__restrict__ quats,
                      const float *__restrict__ scales, const float *__restrict__ viewmats,
                      const float *__restrict__ Ks, CullParams p, int32_t *__restrict__ radii,
                      float *__restrict__ means2d, float *__restrict__ depths, float *__restrict__ conics,
                      float *__restrict__ compensations) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)C * N) return;
    int c = (int)(idx / N);
    int64_t i = idx % N;
    Camera cam;
    load_camera(viewmats + 16 * c, Ks + 9 * c, cam);
    float mean[3], covar[9];
    load_gaussian(means, scales, quats, 1, i, mean, covar);
    Projected o;
    bool ok = project_one<false>(mean, covar, cam, p.width, p.height, p.eps2d, p.near_plane, p.far_plane,
                                 p.radius_clip, o);
    if (!ok) {
        radii[idx] = 0;
        means2d[2 * idx] = 0.f; means2d[2 * idx + 1] = 0.f;
        depths[idx] = 0.f;
        conics[3 * idx] = 0.f; conics[3 * idx + 1] = 0.f; conics[3 * idx + 2] = 0.f;
        if (compensations) compensations[idx] = 0.f;
        return;
    }
    radii[idx] = (int32_t)o.radius;
    means2d[2 * idx] = o.mean2d[0];
    means2d[2 * idx + 1] = o.mean2d[1];
    depths[idx] = o.depth;
    conics[3 * idx] = o.conic[0];
    conics[3 * idx + 1] = o.conic[1];
    conics[3 * idx + 2] = o.conic[2];
    if (compensations) compensations[idx] = o.compensation;
}

CLMGS_EXPORT int clmgs_projection_fwd(int C, int64_t N, const float *means, const float *quats,
                                      const float *scales, const float *viewmats, const float *Ks, int width,
                                      int height, float eps2d, float near_plane, float far_plane,
                                      float radius_clip, int32_t *radii, float *means2d, float *depths,
                                      float *conics, float *compensations, void *stream) {
    CLMGS_REQUIRE(C >= 0 && N >= 0, "projection_fwd: negative size");
    int64_t total = (int64_t)C * N;
    if (total == 0) return CLMGS_OK;
    CullParams p{width, height, eps2d, near_plane, far_plane, radius_clip};
    projection_fwd_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, as_stream(stream)>>>(
        C, N, means, quats, scales, viewmats, Ks, p, radii, means2d, depths, conics, compensations);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------
// S4.  Tile counting with an inclusive int64 prefix sum (three phases, 1024
// elements per block), key emission, per-tile offsets.
// ---------------------------------------------------------------------------
constexpr int kScanThreads = 1024;

__device__ __forceinline__ int tiles_of(const float *means2d, const int32_t *radii, int64_t idx, int tile_size,
                                        int tw, int th) {
    int r = radii[idx];
    if (r <= 0) return 0;
    int x0, y0, x1, y1;
    tile_rect(means2d[2 * idx], means2d[2 * idx + 1], r, tile_size, tw, th, x0, y0, x1, y1);
    return (y1 - y0) * (x1 - x0);
}

__global__ void __launch_bounds__(kScanThreads)
isect_count_kernel(int64_t total, const float *__restrict__ means2d, const int32_t *__restrict__ radii,
                   int tile_size, int tw, int th, int32_t *__restrict__ tiles_per_gauss,
                   int64_t *__restrict__ block_sums) {
    __shared__ int64_t s_warp[33];
    int64_t idx = (int64_t)blockIdx.x * kScanThreads + threadIdx.x;
    int64_t v = 0, tot;
    if (idx < total) {
        v = tiles_of(means2d, radii, idx, tile_size, tw, th);
        tiles_per_gauss[idx] = (int32_t)v;
    }
    block_excl_scan_1024<int64_t>(v, s_warp, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kScanThreads)
scan_block_sums_kernel(int64_t *__restrict__ block_sums, int64_t nblocks) {
    __shared__ int64_t s_warp[33];
    int64_t carry = 0;
    for (int64_t base = 0; base < nblocks; base += kScanThreads) {
        int64_t j = base + threadIdx.x;
        int64_t v = j < nblocks ? block_sums[j] : 0, tot;
        int64_t excl = block_excl_scan_1024<int64_t>(v, s_warp, &tot);
        if (j < nblocks) block_sums[j] = carry + excl;
        carry += tot;
    }
}

__global__ void __launch_bounds__(kScanThreads)
isect_cumsum_kernel(int64_t total, const int32_t *__restrict__ tiles_per_gauss,
                    const int64_t *__restrict__ block_sums, int64_t *__restrict__ cum_tiles) {
    __shared__ int64_t s_warp[33];
    int64_t idx = (int64_t)blockIdx.x * kScanThreads + threadIdx.x;
    int64_t v = idx < total ? tiles_per_gauss[idx] : 0, tot;
    int64_t excl = block_excl_scan_1024<int64_t>(v, s_warp, &tot);
    if (idx < total) cum_tiles[idx] = block_sums[blockIdx.x] + excl + v;
}

CLMGS_EXPORT size_t clmgs_isect_count_scratch_bytes(int64_t total) {
    return (size_t)(ceil_div64(total > 0 ? total : 1, kScanThreads)) * sizeof(int64_t);
}

CLMGS_EXPORT int clmgs_isect_count(int64_t total, const float *means2d, const int32_t *radii, int tile_size,
                                   int tile_width, int tile_height, int32_t *tiles_per_gauss, int64_t *cum_tiles,
                                   void *scratch, size_t scratch_bytes, void *stream) {
    CLMGS_REQUIRE(total >= 0 && tile_size > 0, "isect_count: bad sizes");
    if (total == 0) return CLMGS_OK;
    if (scratch_bytes < clmgs_isect_count_scratch_bytes(total)) {
        clmgs_set_error("isect_count: scratch too small");
        return CLMGS_ERR_SCRATCH;
    }
    cudaStream_t st = as_stream(stream);
    int64_t nblocks = ceil_div64(total, kScanThreads);
    int64_t *bs = (int64_t *)scratch;
    isect_count_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, means2d, radii, tile_size, tile_width,
                                                                  tile_height, tiles_per_gauss, bs);
    CLMGS_LAUNCH_OK();
    scan_block_sums_kernel<<<1, kScanThreads, 0, st>>>(bs, nblocks);
    CLMGS_LAUNCH_OK();
    isect_cumsum_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, tiles_per_gauss, bs, cum_tiles);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

__global__ void __launch_bounds__(256)
isect_emit_kernel(int64_t total, int64_t N, const int64_t *__restrict__ camera_ids,
                  const float *__restrict__ means2d, const int32_t *__restrict__ radii,
                  const float *__restrict__ depths, const int64_t *__restrict__ cum_tiles, int tile_size,
                  int tw, int th, int tile_n_bits, int64_t *__restrict__ isect_ids,
                  int32_t *__restrict__ flatten_ids) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int r = radii[idx];
    if (r <= 0) return;
    int x0, y0, x1, y1;
    tile_rect(means2d[2 * idx], means2d[2 * idx + 1], r, tile_size, tw, th, x0, y0, x1, y1);
    int64_t cid = camera_ids ? camera_ids[idx] : idx / N;
    int64_t cid_enc = cid << (32 + tile_n_bits);
    int64_t depth_enc = (int64_t)__float_as_int(depths[idx]);  // sign-extending cast, as the reference
    int64_t cur = idx == 0 ? 0 : cum_tiles[idx - 1];
    for (int i = y0; i < y1; ++i)
        for (int j = x0; j < x1; ++j) {
            int64_t tile_id = (int64_t)i * tw + j;
            isect_ids[cur] = cid_enc | (tile_id << 32) | depth_enc;
            flatten_ids[cur] = (int32_t)idx;
            ++cur;
        }
}

// radix.cu: hand-written stable LSD radix sort of (u32, i32) pairs with an optional fused last pass
size_t clmgs_radix_u32_scratch(int64_t n);
int clmgs_radix_passes(int end_bit);
int clmgs_radix_sort_u32(int64_t n, uint32_t *keys, uint32_t *keys_alt, int32_t *vals, int32_t *vals_alt, int end_bit,
                         void *scratch, size_t scratch_bytes, cudaStream_t st, uint32_t **keys_out, int32_t **vals_out,
                         int64_t *isect_ids, int32_t *flatten_ids, const float *depths, int32_t *offsets,
                         int64_t n_slots, int tile_n_bits, int n_tiles);

static int n_bits_of(int64_t v) {  // floor(log2(v)) + 1, CSRC/isect_tiles.cu:143-144
    int b = 0;
    while (((int64_t)1 << b) <= v) ++b;
    return b;
}

// ---------------------------------------------------------------------------
// Two-level tile sort.  The reference sorts I = n_isects pairs on a 64-bit key (cam | tile | depth)
// with a stable LSD radix sort: ceil((32 + tile + cam bits)/8) = 6 passes over 12-byte pairs.  The
// same permutation comes from (1) sorting the n Gaussians by depth bits (n << I, 32-bit keys), (2)
// emitting intersections in that order and (3) one STABLE sort on the (cam | tile) bits alone
// (13-14 bits = 2 passes over 8-byte pairs): within a tile the entries keep the order of step (1),
// i.e. ascending depth with ties in ascending flatten index -- exactly what the 64-bit stable sort
// leaves.  Traffic per intersection drops from ~6 x 24 B to ~2 x 20 B + 12 B: the last pass of (3) writes
// isect_ids / flatten_ids itself (radix.cu), there is no separate finalize kernel.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
depth_keys_kernel(int64_t total, const int32_t *__restrict__ radii, const float *__restrict__ depths,
                  uint32_t *__restrict__ keys, int32_t *__restrict__ perm) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    keys[i] = radii[i] > 0 ? (uint32_t)__float_as_int(depths[i]) : 0xffffffffu;  // culled entries last (emit nothing)
    perm[i] = (int32_t)i;
}

__global__ void __launch_bounds__(kScanThreads)
perm_count_kernel(int64_t total, const int32_t *__restrict__ perm, const float *__restrict__ means2d,
                  const int32_t *__restrict__ radii, int tile_size, int tw, int th, int64_t *__restrict__ block_sums) {
    __shared__ int64_t s_warp[33];
    int64_t j = (int64_t)blockIdx.x * kScanThreads + threadIdx.x;
    int64_t v = j < total ? tiles_of(means2d, radii, perm[j], tile_size, tw, th) : 0, tot;
    block_excl_scan_1024<int64_t>(v, s_warp, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// emits the (cam|tile) key and the flatten index of every intersection, Gaussians in depth order
__global__ void __launch_bounds__(kScanThreads)
perm_emit_kernel(int64_t total, int64_t N, const int64_t *__restrict__ camera_ids, const int32_t *__restrict__ perm,
                 const float *__restrict__ means2d, const int32_t *__restrict__ radii,
                 const int64_t *__restrict__ block_sums, int tile_size, int tw, int th, int tile_n_bits,
                 uint32_t *__restrict__ tkeys, int32_t *__restrict__ vals) {
    __shared__ int64_t s_warp[33];
    int64_t j = (int64_t)blockIdx.x * kScanThreads + threadIdx.x;
    int idx = j < total ? perm[j] : 0;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    int64_t v = 0, tot;
    if (j < total && radii[idx] > 0) {
        tile_rect(means2d[2 * (int64_t)idx], means2d[2 * (int64_t)idx + 1], radii[idx], tile_size, tw, th, x0, y0, x1, y1);
        v = (int64_t)(y1 - y0) * (x1 - x0);
    }
    const int64_t cur = block_sums[blockIdx.x] + block_excl_scan_1024<int64_t>(v, s_warp, &tot);
    // Warp-cooperative expansion: the 32 runs of a warp are consecutive in the output, so the warp writes them as
    // one stream -- entry e of the warp goes to lane e % 32 (coalesced 128-byte stores) and finds its Gaussian by
    // a 5-step search over the lanes' start offsets.  (One thread per run wrote 2 x 4 bytes at 32 scattered places.)
    const int lane = threadIdx.x & 31;
    const int64_t warp_base = __shfl_sync(0xffffffffu, cur, 0);
    const int p = (int)(cur - warp_base);                       // start of this lane's run inside the warp's stream
    const int total_w = __shfl_sync(0xffffffffu, p + (int)v, 31);
    const uint32_t cid = (j < total && v > 0) ? (uint32_t)(camera_ids ? camera_ids[idx] : idx / N) : 0u;
    const uint32_t cid_enc = cid << tile_n_bits;
    const int wdt = x1 - x0;
    for (int e0 = 0; e0 < total_w; e0 += 32) {
        const int e = e0 + lane;
        int owner = 0, p_owner = __shfl_sync(0xffffffffu, p, 0);
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {  // largest lane whose run starts at or before e
            const int cand = owner + step;
            const int pc = __shfl_sync(0xffffffffu, p, cand & 31);
            if (cand < 32 && pc <= e) { owner = cand; p_owner = pc; }
        }
        const int ox0 = __shfl_sync(0xffffffffu, x0, owner), oy0 = __shfl_sync(0xffffffffu, y0, owner);
        const int ow = __shfl_sync(0xffffffffu, wdt, owner), oidx = __shfl_sync(0xffffffffu, idx, owner);
        const uint32_t oenc = __shfl_sync(0xffffffffu, cid_enc, owner);
        if (e < total_w) {
            const int r = e - p_owner, row = r / ow, colm = r - row * ow;
            tkeys[warp_base + e] = oenc | (uint32_t)((oy0 + row) * tw + (ox0 + colm));
            vals[warp_base + e] = oidx;
        }
    }
}

static inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

struct TwoLevelLayout {
    size_t dkeys, dkeys_alt, perm, perm_alt, block_sums, tkeys, tkeys_alt, vals_alt, tmp, total_bytes, tmp_bytes;
};
static TwoLevelLayout two_level_layout(int64_t total, int64_t n_isects) {
    TwoLevelLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += a256(bytes); return r; };
    L.dkeys = take(4 * (size_t)total); L.dkeys_alt = take(4 * (size_t)total);
    L.perm = take(4 * (size_t)total); L.perm_alt = take(4 * (size_t)total);
    L.block_sums = take(8 * (size_t)ceil_div64(total > 0 ? total : 1, kScanThreads));
    L.tkeys = take(4 * (size_t)n_isects); L.tkeys_alt = take(4 * (size_t)n_isects);
    L.vals_alt = take(4 * (size_t)n_isects);
    size_t t1 = clmgs_radix_u32_scratch(total), t2 = clmgs_radix_u32_scratch(n_isects);
    L.tmp_bytes = t1 > t2 ? t1 : t2;
    L.tmp = take(L.tmp_bytes);
    L.total_bytes = o;
    return L;
}

CLMGS_EXPORT size_t clmgs_isect_sort_scratch_bytes(int64_t total, int64_t n_isects) {
    if (total <= 0 || n_isects <= 0) return 0;
    return two_level_layout(total, n_isects).total_bytes;
}

static int isect_emit_sort_impl(int C, int64_t N, const int64_t *camera_ids, const float *means2d,
                                const int32_t *radii, const float *depths, const int64_t *cum_tiles,
                                int tile_size, int tile_width, int tile_height, int64_t n_isects, int sort,
                                int64_t *isect_ids, int32_t *flatten_ids, int32_t *tile_offsets, void *scratch,
                                size_t scratch_bytes, void *stream) {
    int64_t total = camera_ids ? N : (int64_t)C * N;
    if (total == 0 || n_isects == 0) {
        if (tile_offsets)  // every tile is empty
            CLMGS_CUDA(cudaMemsetAsync(tile_offsets, 0, sizeof(int32_t) * (size_t)C * tile_width * tile_height, as_stream(stream)));
        return CLMGS_OK;
    }
    CLMGS_REQUIRE(n_isects < ((int64_t)1 << 31) && total < ((int64_t)1 << 31),
                  "isect: n_isects=%lld / total=%lld do not fit int32 indices", (long long)n_isects, (long long)total);
    int tile_n_bits = n_bits_of((int64_t)tile_width * tile_height);
    int cam_n_bits = n_bits_of(C);
    CLMGS_REQUIRE(tile_n_bits + cam_n_bits <= 32, "isect: too many tiles/cameras for a 64-bit key");
    cudaStream_t st = as_stream(stream);
    if (!sort) {  // unsorted emission in flatten-index order (gsplat's sort=False)
        isect_emit_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, st>>>(total, N, camera_ids, means2d, radii, depths,
                                                                           cum_tiles, tile_size, tile_width,
                                                                           tile_height, tile_n_bits, isect_ids,
                                                                           flatten_ids);
        CLMGS_LAUNCH_OK();
        return CLMGS_OK;
    }
    TwoLevelLayout L = two_level_layout(total, n_isects);
    if (scratch == nullptr || scratch_bytes < L.total_bytes) {
        clmgs_set_error("isect_emit_sort: scratch too small (%zu < %zu)", scratch_bytes, L.total_bytes);
        return CLMGS_ERR_SCRATCH;
    }
    char *base = (char *)scratch;
    uint32_t *dkeys = (uint32_t *)(base + L.dkeys), *dkeys_alt = (uint32_t *)(base + L.dkeys_alt);
    int32_t *perm = (int32_t *)(base + L.perm), *perm_alt = (int32_t *)(base + L.perm_alt);
    int64_t *bs = (int64_t *)(base + L.block_sums);
    uint32_t *tkeys = (uint32_t *)(base + L.tkeys), *tkeys_alt = (uint32_t *)(base + L.tkeys_alt);
    int32_t *vals_alt = (int32_t *)(base + L.vals_alt);
    void *tmp = base + L.tmp;
    // (1) Gaussians by depth
    depth_keys_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, st>>>(total, radii, depths, dkeys, perm);
    CLMGS_LAUNCH_OK();
    uint32_t *dk_sorted; int32_t *perm_sorted;
    if (int rc = clmgs_radix_sort_u32(total, dkeys, dkeys_alt, perm, perm_alt, 32, tmp, L.tmp_bytes, st, &dk_sorted,
                                      &perm_sorted, nullptr, nullptr, nullptr, nullptr, 0, 0, 0)) return rc;
    // (2) offsets in depth order, emission
    const int64_t nblocks = ceil_div64(total, kScanThreads);
    perm_count_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, perm_sorted, means2d, radii, tile_size, tile_width,
                                                                 tile_height, bs);
    CLMGS_LAUNCH_OK();
    scan_block_sums_kernel<<<1, kScanThreads, 0, st>>>(bs, nblocks);
    CLMGS_LAUNCH_OK();
    // the last pass of (3) writes flatten_ids, so it must not read it: with an even number of passes the
    // emission goes to flatten_ids (A -> alt -> A), with an odd number to the scratch buffer
    const int end_bit = tile_n_bits + cam_n_bits;
    int32_t *v_first = (clmgs_radix_passes(end_bit) % 2 == 0) ? flatten_ids : vals_alt;
    int32_t *v_other = (v_first == flatten_ids) ? vals_alt : flatten_ids;
    perm_emit_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, N, camera_ids, perm_sorted, means2d, radii, bs,
                                                                tile_size, tile_width, tile_height, tile_n_bits, tkeys,
                                                                v_first);
    CLMGS_LAUNCH_OK();
    // (3) stable partition by (cam | tile); its last pass emits isect_ids / flatten_ids at the sorted positions
    return clmgs_radix_sort_u32(n_isects, tkeys, tkeys_alt, v_first, v_other, end_bit, tmp, L.tmp_bytes, st, nullptr,
                                nullptr, isect_ids, flatten_ids, depths, tile_offsets,
                                (int64_t)C * tile_width * tile_height, tile_n_bits, tile_width * tile_height);
}

CLMGS_EXPORT int clmgs_isect_emit_sort(int C, int64_t N, const int64_t *camera_ids, const float *means2d,
                                       const int32_t *radii, const float *depths, const int64_t *cum_tiles,
                                       int tile_size, int tile_width, int tile_height, int64_t n_isects, int sort,
                                       int64_t *isect_ids, int32_t *flatten_ids, void *scratch,
                                       size_t scratch_bytes, void *stream) {
    CLMGS_REQUIRE(n_isects == 0 || (isect_ids && flatten_ids), "isect_emit_sort: null output");
    return isect_emit_sort_impl(C, N, camera_ids, means2d, radii, depths, cum_tiles, tile_size, tile_width, tile_height,
                                n_isects, sort, isect_ids, flatten_ids, nullptr, scratch, scratch_bytes, stream);
}

CLMGS_EXPORT int clmgs_isect_sorted_offsets(int C, int64_t N, const int64_t *camera_ids, const float *means2d,
                                            const int32_t *radii, const float *depths, const int64_t *cum_tiles,
                                            int tile_size, int tile_width, int tile_height, int64_t n_isects,
                                            int32_t *flatten_ids, int32_t *tile_offsets, void *scratch,
                                            size_t scratch_bytes, void *stream) {
    CLMGS_REQUIRE(tile_offsets && (n_isects == 0 || flatten_ids), "isect_sorted_offsets: null output");
    return isect_emit_sort_impl(C, N, camera_ids, means2d, radii, depths, cum_tiles, tile_size, tile_width, tile_height,
                                n_isects, 1, nullptr, flatten_ids, tile_offsets, scratch, scratch_bytes, stream);
}

__global__ void __launch_bounds__(256)
isect_offsets_kernel(int64_t n_isects, const int64_t *__restrict__ isect_ids, int64_t n_slots, int n_tiles,
                     int tile_n_bits, int32_t *__restrict__ offsets) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_isects) return;
    int64_t hi = isect_ids[idx] >> 32;
    int64_t cur = (hi >> tile_n_bits) * n_tiles + (hi & (((int64_t)1 << tile_n_bits) - 1));
    int64_t prev = -1;
    if (idx > 0) {
        int64_t hp = isect_ids[idx - 1] >> 32;
        prev = (hp >> tile_n_bits) * n_tiles + (hp & (((int64_t)1 << tile_n_bits) - 1));
    }
    for (int64_t t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)idx;  // first entry of every slot in (prev, cur]
    if (idx == n_isects - 1)
        for (int64_t t = cur + 1; t < n_slots; ++t) offsets[t] = (int32_t)n_isects;
}

CLMGS_EXPORT int clmgs_isect_offset_encode(int64_t n_isects, const int64_t *isect_ids, int C, int tile_width,
                                           int tile_height, int32_t *offsets, void *stream) {
    cudaStream_t st = as_stream(stream);
    int n_tiles = tile_width * tile_height;
    int64_t n_slots = (int64_t)C * n_tiles;
    if (n_slots == 0) return CLMGS_OK;
    if (n_isects == 0) {
        CLMGS_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int32_t) * n_slots, st));
        return CLMGS_OK;
    }
    isect_offsets_kernel<<<(unsigned)ceil_div64(n_isects, 256), 256, 0, st>>>(n_isects, isect_ids, n_slots, n_tiles,
                                                                             n_bits_of(n_tiles), offsets);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------
// Tile counting for callers that must not allocate.
//
// clmgs_isect_count needs a few KiB of temporary device memory (one int64 per
// 1024 inputs) and asks the caller for it.  Some callers have no way to provide
// that: a C or ctypes host that owns no device allocator, a trainer whose whole
// HBM budget was carved up at start-up, or a latency-sensitive loop in which a
// cudaMalloc / cudaFree (both may synchronise the device) is not acceptable.
// For them this variant keeps the temporary in a fixed array that is part of
// the loaded library image: nothing is allocated, here or in any routine this
// function calls.
//
// Limits and rules
//   * at most kStaticMaxTotal (16 Mi) inputs per call; larger calls are
//     rejected with CLMGS_ERR_INVALID and a message, nothing is launched;
//   * the array costs 128 KiB of device memory on every GPU that loads the
//     library, until the process exits;
//   * the first call on a device looks up the array's device address (once per
//     device, with that device current).  Make that first call before starting
//     any stream capture;
//   * all calls on a device share the one array and it keeps whatever the last
//     call left in it.  Nothing in it has to start at zero: each call writes
//     the entries it is going to read before reading them.  Two calls that
//     overlap on the GPU (different streams, no ordering between them) race on
//     the array and produce garbage -- order them.
// ---------------------------------------------------------------------------
#include <mutex>

constexpr int64_t kStaticMaxTotal = (int64_t)16 * 1024 * 1024;
constexpr int64_t kStaticBlocks = kStaticMaxTotal / kScanThreads;  // 16384 block sums = 128 KiB
static __device__ __align__(256) int64_t g_isect_static_block_sums[kStaticBlocks];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mutex;
static int64_t *g_static_dev_ptr[kStaticMaxDevices] = {nullptr};

// The array's NAME is only a host-side handle: passing it to a kernel would hand the kernel a
// host address.  The device address comes from cudaGetSymbolAddress, per device.
static int static_block_sums(int64_t **out) {
    int dev = -1;
    CLMGS_CUDA(cudaGetDevice(&dev));
    CLMGS_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "isect_count_static: device index %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_dev_ptr[dev] == nullptr) {
        void *p = nullptr;
        CLMGS_CUDA(cudaGetSymbolAddress(&p, g_isect_static_block_sums));
        CLMGS_REQUIRE(p != nullptr, "isect_count_static: no device address for the static scratch array");
        g_static_dev_ptr[dev] = (int64_t *)p;
    }
    *out = g_static_dev_ptr[dev];
    return CLMGS_OK;
}

CLMGS_EXPORT int64_t clmgs_isect_count_static_max(void) { return kStaticMaxTotal; }

CLMGS_EXPORT int clmgs_isect_count_static(int64_t total, const float *means2d, const int32_t *radii, int tile_size,
                                          int tile_width, int tile_height, int32_t *tiles_per_gauss,
                                          int64_t *cum_tiles, void *stream) {
    CLMGS_REQUIRE(total >= 0 && tile_size > 0, "isect_count_static: bad sizes");
    CLMGS_REQUIRE(total <= kStaticMaxTotal,
                  "isect_count_static: %lld inputs exceed the built-in scratch (max %lld per call); "
                  "split the call or use clmgs_isect_count with caller scratch",
                  (long long)total, (long long)kStaticMaxTotal);
    if (total == 0) return CLMGS_OK;
    int64_t *bs = nullptr;
    if (int rc = static_block_sums(&bs)) return rc;
    cudaStream_t st = as_stream(stream);
    const int64_t nblocks = ceil_div64(total, kScanThreads);
    isect_count_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, means2d, radii, tile_size, tile_width,
                                                                  tile_height, tiles_per_gauss, bs);
    CLMGS_LAUNCH_OK();
    scan_block_sums_kernel<<<1, kScanThreads, 0, st>>>(bs, nblocks);
    CLMGS_LAUNCH_OK();
    isect_cumsum_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(total, tiles_per_gauss, bs, cum_tiles);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
