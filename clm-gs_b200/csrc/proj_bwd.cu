// S3 backward: VJP of the dense projection (fully_fused_projection_bwd_kernel,
// CSRC/fully_fused_projection_bwd.cu:19-185, with the VJPs of CSRC/utils.cuh)
// and the backward of the gather + activations with scatter into the dense
// gradient tensors (engine.py:650-666, 725-742).  Float-tolerance code: normal
// FMA contraction is fine here.
#include "common.cuh"

namespace {

struct M3 { float m[9]; };  // row-major

__device__ __forceinline__ M3 mul(const M3 &a, const M3 &b) {
    M3 c;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            c.m[i * 3 + j] = a.m[i * 3] * b.m[j] + a.m[i * 3 + 1] * b.m[3 + j] + a.m[i * 3 + 2] * b.m[6 + j];
    return c;
}
__device__ __forceinline__ M3 transpose(const M3 &a) {
    M3 c;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) c.m[i * 3 + j] = a.m[j * 3 + i];
    return c;
}

__device__ __forceinline__ void quat_to_rotmat(const float q[4], M3 &R, float qn[4], float &inv_norm) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    inv_norm = rsqrtf(x * x + y * y + z * z + w * w);
    x *= inv_norm; y *= inv_norm; z *= inv_norm; w *= inv_norm;
    qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z;
    float x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    R.m[0] = 1.f - 2.f * (y2 + z2); R.m[1] = 2.f * (xy - wz); R.m[2] = 2.f * (xz + wy);
    R.m[3] = 2.f * (xy + wz); R.m[4] = 1.f - 2.f * (x2 + z2); R.m[5] = 2.f * (yz - wx);
    R.m[6] = 2.f * (xz - wy); R.m[7] = 2.f * (yz + wx); R.m[8] = 1.f - 2.f * (x2 + y2);
}

__global__ void __launch_bounds__(256)
projection_bwd_kernel(int C, int64_t N, const float *__restrict__ means, const float *__restrict__ quats,
                      const float *__restrict__ scales, const float *__restrict__ viewmats,
                      const float *__restrict__ Ks, int width, int height, const int32_t *__restrict__ radii,
                      const float *__restrict__ conics, const float *__restrict__ v_means2d,
                      const float *__restrict__ v_depths, const float *__restrict__ v_conics,
                      float *__restrict__ v_means, float *__restrict__ v_quats, float *__restrict__ v_scales) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)C * N || radii[idx] <= 0) return;
    const int c = (int)(idx / N);
    const int64_t i = idx % N;
    const float *vm = viewmats + 16 * c, *K = Ks + 9 * c;
    M3 R;
    float t[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int k = 0; k < 3; ++k) R.m[r * 3 + k] = vm[r * 4 + k];
        t[r] = vm[r * 4 + 3];
    }
    // inverse_vjp: v_cov2d = -P vP P
    const float p00 = conics[3 * idx], p01 = conics[3 * idx + 1], p11 = conics[3 * idx + 2];
    const float g00 = v_conics[3 * idx], g01 = v_conics[3 * idx + 1] * .5f, g11 = v_conics[3 * idx + 2];
    const float a00 = p00 * g00 + p01 * g01, a01 = p00 * g01 + p01 * g11;
    const float a10 = p01 * g00 + p11 * g01, a11 = p01 * g01 + p11 * g11;
    const float V00 = -(a00 * p00 + a01 * p01), V01 = -(a00 * p01 + a01 * p11);
    const float V10 = -(a10 * p00 + a11 * p01), V11 = -(a10 * p01 + a11 * p11);

    float q[4] = {quats[4 * i], quats[4 * i + 1], quats[4 * i + 2], quats[4 * i + 3]};
    float s[3] = {scales[3 * i], scales[3 * i + 1], scales[3 * i + 2]};
    float mean[3] = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    M3 Rq, M;
    float qn[4], inv_norm;
    quat_to_rotmat(q, Rq, qn, inv_norm);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int k = 0; k < 3; ++k) M.m[r * 3 + k] = Rq.m[r * 3 + k] * s[k];
    M3 covar = mul(M, transpose(M));
    float mc[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) mc[r] = R.m[r * 3] * mean[0] + R.m[r * 3 + 1] * mean[1] + R.m[r * 3 + 2] * mean[2] + t[r];
    M3 Rt = transpose(R);
    M3 cc = mul(mul(R, covar), Rt);

    // persp_proj_vjp (CSRC/utils.cuh:174-232)
    const float fx = K[0], fy = K[4];
    const float x = mc[0], y = mc[1], z = mc[2];
    const float lim_x = 1.3f * (0.5f * (float)width / fx), lim_y = 1.3f * (0.5f * (float)height / fy);
    const float rz = 1.f / z, rz2 = rz * rz, rz3 = rz2 * rz;
    const float tx = z * fminf(lim_x, fmaxf(-lim_x, x * rz));
    const float ty = z * fminf(lim_y, fmaxf(-lim_y, y * rz));
    const float J[2][3] = {{fx * rz, 0.f, -fx * tx * rz2}, {0.f, fy * rz, -fy * ty * rz2}};
    const float Vm[2][2] = {{V00, V01}, {V10, V11}};
    M3 v_cc;  // J^T V J
    float JtV[3][2];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) JtV[a][b] = J[0][a] * Vm[0][b] + J[1][a] * Vm[1][b];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) v_cc.m[a * 3 + b] = JtV[a][0] * J[0][b] + JtV[a][1] * J[1][b];
    const float vm2x = v_means2d[2 * idx], vm2y = v_means2d[2 * idx + 1];
    float v_mc[3] = {fx * rz * vm2x, fy * rz * vm2y, -(fx * x * vm2x + fy * y * vm2y) * rz2};
    float VJ[2][3], VtJ[2][3], vJ02, vJ12, vJ00, vJ11;
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            VJ[a][b] = Vm[a][0] * J[0][b] + Vm[a][1] * J[1][b];
            VtJ[a][b] = Vm[0][a] * J[0][b] + Vm[1][a] * J[1][b];
        }
    auto vJ = [&](int a, int b) {
        return (VJ[a][0] * cc.m[b * 3] + VJ[a][1] * cc.m[b * 3 + 1] + VJ[a][2] * cc.m[b * 3 + 2]) +
               (VtJ[a][0] * cc.m[b] + VtJ[a][1] * cc.m[3 + b] + VtJ[a][2] * cc.m[6 + b]);
    };
    vJ00 = vJ(0, 0); vJ02 = vJ(0, 2); vJ11 = vJ(1, 1); vJ12 = vJ(1, 2);
    if (x * rz <= lim_x && x * rz >= -lim_x) v_mc[0] += -fx * rz2 * vJ02;
    else v_mc[2] += -fx * rz3 * vJ02 * tx;
    if (y * rz <= lim_y && y * rz >= -lim_y) v_mc[1] += -fy * rz2 * vJ12;
    else v_mc[2] += -fy * rz3 * vJ12 * ty;
    v_mc[2] += -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.f * fx * tx * rz3 * vJ02 + 2.f * fy * ty * rz3 * vJ12;
    if (v_depths) v_mc[2] += v_depths[idx];

    // world->camera VJPs (CSRC/utils.cuh:241-280)
    float v_mean[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) v_mean[a] = R.m[a] * v_mc[0] + R.m[3 + a] * v_mc[1] + R.m[6 + a] * v_mc[2];
    M3 v_covar = mul(mul(Rt, v_cc), R);

    // quat_scale_to_covar_vjp (CSRC/utils.cuh:75-107) + quat_to_rotmat_vjp (:31-52)
    M3 sym;
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) sym.m[a * 3 + b] = v_covar.m[a * 3 + b] + v_covar.m[b * 3 + a];
    M3 v_M = mul(sym, M);
    float vR[9];  // row-major d/dR
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) vR[a * 3 + b] = v_M.m[a * 3 + b] * s[b];
    float v_scale[3];
#pragma unroll
    for (int b = 0; b < 3; ++b) v_scale[b] = Rq.m[b] * v_M.m[b] + Rq.m[3 + b] * v_M.m[3 + b] + Rq.m[6 + b] * v_M.m[6 + b];
    const float w = qn[0], qx = qn[1], qy = qn[2], qz = qn[3];
    // glm v_R[c][r] == vR[r*3+c]
#define VR(c, r) vR[(r) * 3 + (c)]
    float vq[4];
    vq[0] = 2.f * (qx * (VR(1, 2) - VR(2, 1)) + qy * (VR(2, 0) - VR(0, 2)) + qz * (VR(0, 1) - VR(1, 0)));
    vq[1] = 2.f * (-2.f * qx * (VR(1, 1) + VR(2, 2)) + qy * (VR(0, 1) + VR(1, 0)) + qz * (VR(0, 2) + VR(2, 0)) +
                   w * (VR(1, 2) - VR(2, 1)));
    vq[2] = 2.f * (qx * (VR(0, 1) + VR(1, 0)) - 2.f * qy * (VR(0, 0) + VR(2, 2)) + qz * (VR(1, 2) + VR(2, 1)) +
                   w * (VR(2, 0) - VR(0, 2)));
    vq[3] = 2.f * (qx * (VR(0, 2) + VR(2, 0)) + qy * (VR(1, 2) + VR(2, 1)) - 2.f * qz * (VR(0, 0) + VR(1, 1)) +
                   w * (VR(0, 1) - VR(1, 0)));
#undef VR
    const float dotq = vq[0] * qn[0] + vq[1] * qn[1] + vq[2] * qn[2] + vq[3] * qn[3];
    if (C == 1) {  // one contribution per Gaussian: plain read-modify-write
#pragma unroll
        for (int k = 0; k < 4; ++k) v_quats[4 * i + k] += (vq[k] - dotq * qn[k]) * inv_norm;
#pragma unroll
        for (int k = 0; k < 3; ++k) { v_means[3 * i + k] += v_mean[k]; v_scales[3 * i + k] += v_scale[k]; }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) atomicAdd(v_quats + 4 * i + k, (vq[k] - dotq * qn[k]) * inv_norm);
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicAdd(v_means + 3 * i + k, v_mean[k]); atomicAdd(v_scales + 3 * i + k, v_scale[k]); }
    }
}

// deterministic activations, same op order as gs_math.cuh (kept local: this TU
// is compiled with FMA contraction, the explicit fmaf chain is unaffected)
__device__ __forceinline__ float det_expf(float x) {
    if (x > 88.72283f) return __int_as_float(0x7f800000);
    if (x < -103.0f) return 0.0f;
    float n = rintf(__fmul_rn(x, 1.442695040888963f));
    float r = fmaf(-n, 0.693145751953125f, x);
    r = fmaf(-n, 1.428606765330187e-06f, r);
    float p = 1.9841270e-4f;
    p = fmaf(p, r, 1.3888889e-3f); p = fmaf(p, r, 8.3333333e-3f); p = fmaf(p, r, 4.1666668e-2f);
    p = fmaf(p, r, 1.6666667e-1f); p = fmaf(p, r, 0.5f); p = fmaf(p, r, 1.0f); p = fmaf(p, r, 1.0f);
    int ni = (int)n, n1 = ni / 2, n2 = ni - n1;
    return __fmul_rn(__fmul_rn(p, __int_as_float((n1 + 127) << 23)), __int_as_float((n2 + 127) << 23));
}

__global__ void __launch_bounds__(256)
activate_bwd_scatter_kernel(const float *__restrict__ log_scale, const float *__restrict__ raw_quat,
                            const float *__restrict__ opacity_logit, const int64_t *__restrict__ ids, int64_t n,
                            const float *__restrict__ v_means, const float *__restrict__ v_scales,
                            const float *__restrict__ v_quats, const float *__restrict__ v_opac,
                            float *__restrict__ g_xyz, float *__restrict__ g_ls, float *__restrict__ g_q,
                            float *__restrict__ g_o) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int64_t g = ids[r];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        g_xyz[3 * g + k] += v_means[3 * r + k];
        g_ls[3 * g + k] += v_scales[3 * r + k] * det_expf(log_scale[3 * g + k]);  // d exp(x) = exp(x)
    }
    // q_n = q / max(|q|, eps):  v_q = (v - (v . q_n) q_n) / |q|
    float q[4] = {raw_quat[4 * g], raw_quat[4 * g + 1], raw_quat[4 * g + 2], raw_quat[4 * g + 3]};
    float nn = fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
    float inv = 1.0f / nn, dot = 0.f, v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[k] = v_quats[4 * r + k]; q[k] *= inv; dot += v[k] * q[k]; }
#pragma unroll
    for (int k = 0; k < 4; ++k) g_q[4 * g + k] += (v[k] - dot * q[k]) * inv;
    const float o = 1.0f / (1.0f + det_expf(-opacity_logit[g]));
    g_o[g] += v_opac[r] * o * (1.0f - o);
}

// A17 densification statistics (densification.py:59-102 -> gaussian_model.py:833-851): one thread per
// visible Gaussian; ids are unique within a camera, so plain read-modify-write.
__global__ void __launch_bounds__(256)
densification_stats_kernel(const int64_t *__restrict__ ids, int64_t n, const float *__restrict__ v_means2d,
                           const int32_t *__restrict__ radii, float half_w, float half_h,
                           float *__restrict__ max_radii2d, float *__restrict__ grad_accum, float *__restrict__ denom) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int64_t g = ids[r];
    max_radii2d[g] = fmaxf(max_radii2d[g], (float)radii[r]);
    const float gx = v_means2d[2 * r] * half_w, gy = v_means2d[2 * r + 1] * half_h;
    grad_accum[g] += sqrtf(gx * gx + gy * gy);
    denom[g] += 1.0f;
}

}  // namespace

CLMGS_EXPORT int clmgs_densification_stats(const int64_t *ids, int64_t n, const float *v_means2d,
                                           const int32_t *radii, int width, int height, float *max_radii2d,
                                           float *xyz_gradient_accum, float *denom, void *stream) {
    if (n == 0) return CLMGS_OK;
    densification_stats_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, as_stream(stream)>>>(
        ids, n, v_means2d, radii, 0.5f * (float)width, 0.5f * (float)height, max_radii2d, xyz_gradient_accum, denom);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_projection_bwd(int C, int64_t N, const float *means, const float *quats,
                                      const float *scales, const float *viewmats, const float *Ks, int width,
                                      int height, float eps2d, const int32_t *radii, const float *conics,
                                      const float *v_means2d, const float *v_depths, const float *v_conics,
                                      float *v_means, float *v_quats, float *v_scales, void *stream) {
    (void)eps2d;  // only needed for the compensation term, which CLM does not use
    int64_t total = (int64_t)C * N;
    if (total == 0) return CLMGS_OK;
    projection_bwd_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, as_stream(stream)>>>(
        C, N, means, quats, scales, viewmats, Ks, width, height, radii, conics, v_means2d, v_depths, v_conics,
        v_means, v_quats, v_scales);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_activate_bwd_scatter(const float *log_scale, const float *raw_quat,
                                            const float *opacity_logit, const int64_t *ids, int64_t n,
                                            const float *v_means, const float *v_scales, const float *v_quats,
                                            const float *v_opacities, float *g_xyz, float *g_log_scale,
                                            float *g_raw_quat, float *g_opacity_logit, void *stream) {
    if (n == 0) return CLMGS_OK;
    activate_bwd_scatter_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, as_stream(stream)>>>(
        log_scale, raw_quat, opacity_logit, ids, n, v_means, v_scales, v_quats, v_opacities, g_xyz, g_log_scale,
        g_raw_quat, g_opacity_logit);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
