// A16 GPU Adam for the GPU-resident attributes (xyz, opacity, scaling, rotation).
//  * selective: masked update without bias correction (CK/adam.cu:12-41)
//  * dense: torch.optim.Adam(fused=True) rule with bias correction (optimizer.py:127-129)
// `grad_scale` folds the `param.grad /= bsz` pass of engine.py:870-872 into the update.
// HBM-bound: 28 B per updated float, one thread per float, grid-stride.
#include "common.cuh"

namespace {
constexpr int kT = 256;
inline unsigned grid_n(int64_t n) {
    int64_t b = ceil_div64(n, kT);
    int64_t cap = (int64_t)kNumSMs * 16;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

__global__ void __launch_bounds__(kT)
selective_adam_kernel(float *__restrict__ param, const float *__restrict__ grad, float *__restrict__ m,
                      float *__restrict__ v, const uint8_t *__restrict__ mask, float lr, float b1, float b2,
                      float eps, float gs, int64_t total, int M) {
    for (int64_t p = (int64_t)blockIdx.x * kT + threadIdx.x; p < total; p += (int64_t)gridDim.x * kT) {
        if (!mask[p / M]) continue;
        const float g = grad[p] * gs;
        const float mm = b1 * m[p] + (1.0f - b1) * g;
        const float vv = b2 * v[p] + (1.0f - b2) * g * g;
        param[p] += -lr * mm / (sqrtf(vv) + eps);
        m[p] = mm;
        v[p] = vv;
    }
}

__global__ void __launch_bounds__(kT)
adam_dense_kernel(float *__restrict__ param, const float *__restrict__ grad, float *__restrict__ m,
                  float *__restrict__ v, float step_size, float b1, float b2, float eps, float bc2_sqrt, float gs,
                  int64_t n) {
    for (int64_t p = (int64_t)blockIdx.x * kT + threadIdx.x; p < n; p += (int64_t)gridDim.x * kT) {
        const float g = grad[p] * gs;
        const float m0 = m[p];
        const float mm = m0 + (g - m0) * (1.0f - b1);  // torch lerp form
        const float vv = b2 * v[p] + (1.0f - b2) * g * g;
        const float denom = sqrtf(vv) / bc2_sqrt + eps;
        param[p] -= step_size * (mm / denom);
        m[p] = mm;
        v[p] = vv;
    }
}
}  // namespace

CLMGS_EXPORT int clmgs_selective_adam(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                                      const uint8_t *mask, float lr, float b1, float b2, float eps,
                                      float grad_scale, int64_t N, int M, void *stream) {
    CLMGS_REQUIRE(M >= 1, "selective_adam: M must be >= 1");
    int64_t total = N * M;
    if (total == 0) return CLMGS_OK;
    selective_adam_kernel<<<grid_n(total), kT, 0, as_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq, mask, lr, b1,
                                                                      b2, eps, grad_scale, total, M);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_adam_dense(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, float lr,
                                  float b1, float b2, float eps, float grad_scale, int64_t step, int64_t n,
                                  void *stream) {
    CLMGS_REQUIRE(step >= 1, "adam_dense: step must be >= 1 (1-based, after increment)");
    if (n == 0) return CLMGS_OK;
    const double bc1 = 1.0 - pow((double)b1, (double)step), bc2 = 1.0 - pow((double)b2, (double)step);
    adam_dense_kernel<<<grid_n(n), kT, 0, as_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq,
                                                              (float)((double)lr / bc1), b1, b2, eps,
                                                              (float)sqrt(bc2), grad_scale, n);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
