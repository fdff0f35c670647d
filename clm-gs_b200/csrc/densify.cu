// Densify / prune / opacity reset on the split GPU + pinned-host state.
//
// Reference flow (strategies/base_gaussian_model.py:364-388 densify_and_prune;
// strategies/clm_offload/gaussian_model.py:741-831 split / clone, :522-611 prune, :488-520 reset):
// clone (append copies) -> split (append 2 children per parent, prune the parents) -> prune
// (opacity / world-size), each step re-concatenating every tensor and compacting the pinned table
// on the CPU.  The three steps only ever (a) keep a row, (b) copy a row or (c) derive a child from
// a row, and every decision depends on the ORIGINAL row alone, so the whole step is
//     classify (1 pass over N) -> plan (one stable 4-way compaction = the final row order)
//     -> apply (one gather kernel writes all new GPU tensors and moments)
// plus one in-place compaction of the pinned tables on the host (cpu_adam.cpp).
// Final row order = the reference's: kept originals, kept clones, kept children copy 0, copy 1.
//
// Compiled with -fmad=false: decisions use det_expf / det_sigmoidf / det_logf so that the plan is
// reproducible bit for bit by the oracle.
#include "common.cuh"
#include "gs_math.cuh"

namespace {

constexpr int kT = 256;

// cls bits
constexpr uint8_t kClone = 1, kSplit = 2, kDropSelf = 4, kDropChild = 8;

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

__global__ void __launch_bounds__(kT)
classify_kernel(int64_t N, const float *__restrict__ grad_accum, const float *__restrict__ denom,
                const float *__restrict__ log_scale, const float *__restrict__ opacity_logit, float grad_threshold,
                float dense_threshold, float min_opacity, float world_size_threshold, int use_world_size,
                uint8_t *__restrict__ cls) {
    int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= N) return;
    float g = grad_accum[i] / denom[i];  // base_gaussian_model.py:366-367
    if (g != g) g = 0.0f;
    float s0 = gsm::det_expf(log_scale[3 * i]), s1 = gsm::det_expf(log_scale[3 * i + 1]),
          s2 = gsm::det_expf(log_scale[3 * i + 2]);
    float smax = max3(s0, s1, s2);
    uint8_t c = 0;
    if (fabsf(g) >= grad_threshold && smax <= dense_threshold) c |= kClone;  // GM:803-810
    if (g >= grad_threshold && smax > dense_threshold) c |= kSplit;          // GM:745-759
    bool low = gsm::det_sigmoidf(opacity_logit[i]) < min_opacity;            // base:372
    if (low || (use_world_size && smax > world_size_threshold)) c |= kDropSelf;  // base:382-385
    if (c & kSplit) {
        // the child's stored scaling is log(s / (0.8 * 2)) (GM:773-775); the prune looks at exp of it
        float c0 = gsm::det_expf(gsm::det_logf(s0 / 1.6f)), c1 = gsm::det_expf(gsm::det_logf(s1 / 1.6f)),
              c2 = gsm::det_expf(gsm::det_logf(s2 / 1.6f));
        if (low || (use_world_size && max3(c0, c1, c2) > world_size_threshold)) c |= kDropChild;
    }
    cls[i] = c;
}

__global__ void __launch_bounds__(kT)
mask_to_cls_kernel(int64_t N, const uint8_t *__restrict__ prune_mask, uint8_t *__restrict__ cls) {
    int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i < N) cls[i] = prune_mask[i] ? kDropSelf : 0;
}

// predicate p of a class byte: 0 keep-self (A), 1 kept clone (B), 2 split parent (S), 3 kept split parent (C)
__device__ __forceinline__ bool pred(uint8_t c, int p) {
    switch (p) {
        case 0: return !(c & kSplit) && !(c & kDropSelf);
        case 1: return (c & kClone) && !(c & kDropSelf);
        case 2: return (c & kSplit) != 0;
        default: return (c & kSplit) && !(c & kDropChild);
    }
}

__global__ void __launch_bounds__(kT)
plan_count_kernel(int64_t N, const uint8_t *__restrict__ cls, int32_t *__restrict__ counts, int64_t nblocks) {
    int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    bool valid = i < N;
    uint8_t c = valid ? cls[i] : 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        int n = __syncthreads_count(valid && pred(c, p));
        if (threadIdx.x == 0) counts[(int64_t)p * nblocks + blockIdx.x] = n;
    }
}

// exclusive scan of each predicate's per-block counts; totals[p] = number of rows with predicate p
__global__ void __launch_bounds__(1024)
plan_scan_kernel(const int32_t *__restrict__ counts, int64_t nblocks, int64_t *__restrict__ offsets,
                 int64_t *__restrict__ totals) {
    __shared__ int64_t s_warp[33];
    __shared__ int64_t s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = 0; p < 4; ++p) {
        int64_t carry = 0;
        for (int64_t base = 0; base < nblocks; base += 1024) {
            int64_t j = base + threadIdx.x;
            int64_t v = j < nblocks ? counts[(int64_t)p * nblocks + j] : 0, incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int64_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) s_warp[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                int64_t w = s_warp[lane], wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { int64_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
                s_warp[lane] = wi - w;
                if (lane == 31) s_carry = wi;
            }
            __syncthreads();
            if (j < nblocks) offsets[(int64_t)p * nblocks + j] = carry + s_warp[warp] + incl - v;
            carry += s_carry;
            __syncthreads();
        }
        if (threadIdx.x == 0) totals[p] = carry;
    }
}

__global__ void __launch_bounds__(kT)
plan_emit_kernel(int64_t N, const uint8_t *__restrict__ cls, const int64_t *__restrict__ offsets, int64_t nblocks,
                 const int64_t *__restrict__ totals, int32_t *__restrict__ src, int32_t *__restrict__ noise_row) {
    __shared__ int s_w[4][kT / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    bool valid = i < N;
    uint8_t c = valid ? cls[i] : 0;
    unsigned b[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        b[p] = __ballot_sync(0xffffffffu, valid && pred(c, p));
        if (lane == 0) s_w[p][warp] = __popc(b[p]);
    }
    __syncthreads();
    int64_t pos[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        int before = 0;
        for (int w = 0; w < warp; ++w) before += s_w[p][w];
        pos[p] = offsets[(int64_t)p * nblocks + blockIdx.x] + before + __popc(b[p] & ((1u << lane) - 1u));
    }
    if (!valid) return;
    const int64_t nA = totals[0], nB = totals[1], nC = totals[3];
    if (pred(c, 0)) src[pos[0]] = (int32_t)i;
    if (pred(c, 1)) src[nA + pos[1]] = (int32_t)i;
    if (pred(c, 3)) {
        src[nA + nB + pos[3]] = (int32_t)i;
        src[nA + nB + nC + pos[3]] = (int32_t)i;
        noise_row[pos[3]] = (int32_t)pos[2];  // rank among ALL split parents: the reference draws noise for all of them
    }
}

struct Attrs {
    const float *xyz, *opacity, *log_scale, *quat;
    const float *m_xyz, *m_opacity, *m_log_scale, *m_quat;
    const float *v_xyz, *v_opacity, *v_log_scale, *v_quat;
};
struct AttrsOut {
    float *xyz, *opacity, *log_scale, *quat;
    float *m_xyz, *m_opacity, *m_log_scale, *m_quat;
    float *v_xyz, *v_opacity, *v_log_scale, *v_quat;
};

__device__ __forceinline__ void copy_n(float *dst, const float *src, int n) {
    for (int k = 0; k < n; ++k) dst[k] = src[k];
}
__device__ __forceinline__ void zero_n(float *dst, int n) {
    for (int k = 0; k < n; ++k) dst[k] = 0.0f;
}

__global__ void __launch_bounds__(kT)
apply_kernel(int64_t n_out, int64_t nA, int64_t nB, int64_t nC, int64_t nS, const int32_t *__restrict__ src,
             const int32_t *__restrict__ noise_row, const float *__restrict__ noise, Attrs in, AttrsOut out) {
    int64_t j = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (j >= n_out) return;
    const int64_t i = src[j];
    if (j < nA) {  // kept row: parameters and Adam moments move together (GM:522-576)
        copy_n(out.m_xyz + 3 * j, in.m_xyz + 3 * i, 3); copy_n(out.v_xyz + 3 * j, in.v_xyz + 3 * i, 3);
        out.m_opacity[j] = in.m_opacity[i]; out.v_opacity[j] = in.v_opacity[i];
        copy_n(out.m_log_scale + 3 * j, in.m_log_scale + 3 * i, 3); copy_n(out.v_log_scale + 3 * j, in.v_log_scale + 3 * i, 3);
        copy_n(out.m_quat + 4 * j, in.m_quat + 4 * i, 4); copy_n(out.v_quat + 4 * j, in.v_quat + 4 * i, 4);
    } else {  // appended rows start with zero moments (GM:613-700)
        zero_n(out.m_xyz + 3 * j, 3); zero_n(out.v_xyz + 3 * j, 3);
        out.m_opacity[j] = 0.0f; out.v_opacity[j] = 0.0f;
        zero_n(out.m_log_scale + 3 * j, 3); zero_n(out.v_log_scale + 3 * j, 3);
        zero_n(out.m_quat + 4 * j, 4); zero_n(out.v_quat + 4 * j, 4);
    }
    out.opacity[j] = in.opacity[i];
    copy_n(out.quat + 4 * j, in.quat + 4 * i, 4);
    if (j < nA + nB) {
        copy_n(out.xyz + 3 * j, in.xyz + 3 * i, 3);
        copy_n(out.log_scale + 3 * j, in.log_scale + 3 * i, 3);
        return;
    }
    // split child (GM:761-776): sample ~ N(0, diag(s^2)) in the parent's frame
    const int64_t c = j - nA - nB, which = c / nC, r = c - which * nC;
    const float *z = noise + 3 * (which * nS + noise_row[r]);
    float s[3], smp[3];
    for (int k = 0; k < 3; ++k) {
        s[k] = gsm::det_expf(in.log_scale[3 * i + k]);
        smp[k] = z[k] * s[k];
        out.log_scale[3 * j + k] = gsm::det_logf(s[k] / 1.6f);
    }
    // utils/general_utils.py:311-334 build_rotation
    float qr = in.quat[4 * i], qx = in.quat[4 * i + 1], qy = in.quat[4 * i + 2], qz = in.quat[4 * i + 3];
    float nrm = sqrtf(qr * qr + qx * qx + qy * qy + qz * qz);
    qr /= nrm; qx /= nrm; qy /= nrm; qz /= nrm;
    float R[9] = {1.f - 2.f * (qy * qy + qz * qz), 2.f * (qx * qy - qr * qz), 2.f * (qx * qz + qr * qy),
                  2.f * (qx * qy + qr * qz), 1.f - 2.f * (qx * qx + qz * qz), 2.f * (qy * qz - qr * qx),
                  2.f * (qx * qz - qr * qy), 2.f * (qy * qz + qr * qx), 1.f - 2.f * (qx * qx + qy * qy)};
    for (int k = 0; k < 3; ++k)
        out.xyz[3 * j + k] = (R[3 * k] * smp[0] + R[3 * k + 1] * smp[1] + R[3 * k + 2] * smp[2]) + in.xyz[3 * i + k];
}

// GM:488-520: opacity <- inverse_sigmoid(min(sigmoid(opacity), 0.01)); its Adam moments <- 0
__global__ void __launch_bounds__(kT)
reset_opacity_kernel(int64_t N, float *__restrict__ opacity_logit, float *__restrict__ m, float *__restrict__ v,
                     float max_opacity) {
    int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= N) return;
    float o = fminf(gsm::det_sigmoidf(opacity_logit[i]), max_opacity);
    opacity_logit[i] = gsm::det_logf(o / (1.0f - o));
    m[i] = 0.0f;
    v[i] = 0.0f;
}

inline unsigned grid_n(int64_t n) { return (unsigned)(n > 0 ? ceil_div64(n, kT) : 1); }

}  // namespace

CLMGS_EXPORT int clmgs_densify_classify(int64_t N, const float *grad_accum, const float *denom, const float *log_scale,
                                        const float *opacity_logit, float grad_threshold, float dense_threshold,
                                        float min_opacity, float world_size_threshold, int use_world_size,
                                        uint8_t *cls, void *stream) {
    CLMGS_REQUIRE(N >= 0 && N < (int64_t)1 << 31, "densify_classify: N=%lld out of range", (long long)N);
    if (N == 0) return CLMGS_OK;
    CLMGS_REQUIRE(grad_accum && denom && log_scale && opacity_logit && cls, "densify_classify: null pointer");
    classify_kernel<<<grid_n(N), kT, 0, as_stream(stream)>>>(N, grad_accum, denom, log_scale, opacity_logit,
                                                            grad_threshold, dense_threshold, min_opacity,
                                                            world_size_threshold, use_world_size, cls);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_prune_mask_to_class(int64_t N, const uint8_t *prune_mask, uint8_t *cls, void *stream) {
    CLMGS_REQUIRE(N >= 0 && N < (int64_t)1 << 31, "prune_mask_to_class: N=%lld out of range", (long long)N);
    if (N == 0) return CLMGS_OK;
    CLMGS_REQUIRE(prune_mask && cls, "prune_mask_to_class: null pointer");
    mask_to_cls_kernel<<<grid_n(N), kT, 0, as_stream(stream)>>>(N, prune_mask, cls);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT size_t clmgs_densify_plan_scratch_bytes(int64_t N) {
    int64_t nblocks = ceil_div64(N > 0 ? N : 1, kT);
    return (size_t)nblocks * 4 * (sizeof(int64_t) + sizeof(int32_t)) + 512;
}

CLMGS_EXPORT int clmgs_densify_plan(int64_t N, const uint8_t *cls, int32_t *src, int32_t *noise_row, int64_t *totals,
                                    void *scratch, size_t scratch_bytes, void *stream) {
    CLMGS_REQUIRE(N >= 0 && N < (int64_t)1 << 31, "densify_plan: N=%lld out of range", (long long)N);
    CLMGS_REQUIRE(totals, "densify_plan: totals is null");
    cudaStream_t st = as_stream(stream);
    if (N == 0) { CLMGS_CUDA(cudaMemsetAsync(totals, 0, 4 * sizeof(int64_t), st)); return CLMGS_OK; }
    CLMGS_REQUIRE(cls && src && noise_row && scratch, "densify_plan: null pointer");
    CLMGS_REQUIRE(scratch_bytes >= clmgs_densify_plan_scratch_bytes(N), "densify_plan: scratch too small");
    int64_t nblocks = ceil_div64(N, kT);
    int64_t *offsets = (int64_t *)scratch;
    int32_t *counts = (int32_t *)((char *)scratch + ((size_t)nblocks * 4 * sizeof(int64_t) + 255) / 256 * 256);
    plan_count_kernel<<<(unsigned)nblocks, kT, 0, st>>>(N, cls, counts, nblocks);
    CLMGS_LAUNCH_OK();
    plan_scan_kernel<<<1, 1024, 0, st>>>(counts, nblocks, offsets, totals);
    CLMGS_LAUNCH_OK();
    plan_emit_kernel<<<(unsigned)nblocks, kT, 0, st>>>(N, cls, offsets, nblocks, totals, src, noise_row);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_densify_apply(int64_t nA, int64_t nB, int64_t nC, int64_t nS, const int32_t *src,
                                     const int32_t *noise_row, const float *noise, const float *const *in12,
                                     float *const *out12, void *stream) {
    CLMGS_REQUIRE(nA >= 0 && nB >= 0 && nC >= 0 && nS >= nC, "densify_apply: bad counts");
    int64_t n_out = nA + nB + 2 * nC;
    if (n_out == 0) return CLMGS_OK;
    CLMGS_REQUIRE(n_out < (int64_t)1 << 31, "densify_apply: %lld rows exceed int32 ids", (long long)n_out);
    CLMGS_REQUIRE(src && in12 && out12 && (nC == 0 || (noise && noise_row)), "densify_apply: null pointer");
    for (int k = 0; k < 12; ++k) CLMGS_REQUIRE(in12[k] && out12[k], "densify_apply: null tensor %d", k);
    Attrs in{in12[0], in12[1], in12[2], in12[3], in12[4], in12[5], in12[6], in12[7], in12[8], in12[9], in12[10], in12[11]};
    AttrsOut out{out12[0], out12[1], out12[2], out12[3], out12[4], out12[5], out12[6], out12[7], out12[8], out12[9], out12[10], out12[11]};
    apply_kernel<<<grid_n(n_out), kT, 0, as_stream(stream)>>>(n_out, nA, nB, nC, nS, src, noise_row, noise, in, out);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_reset_opacity(int64_t N, float *opacity_logit, float *exp_avg, float *exp_avg_sq,
                                     float max_opacity, void *stream) {
    CLMGS_REQUIRE(N >= 0, "reset_opacity: N<0");
    if (N == 0) return CLMGS_OK;
    CLMGS_REQUIRE(opacity_logit && exp_avg && exp_avg_sq, "reset_opacity: null pointer");
    CLMGS_REQUIRE(max_opacity > 0.0f && max_opacity < 1.0f, "reset_opacity: max_opacity must be in (0,1)");
    reset_opacity_kernel<<<grid_n(N), kT, 0, as_stream(stream)>>>(N, opacity_logit, exp_avg, exp_avg_sq, max_opacity);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
