// Decision-critical projection math (cull masks, radii, tile rectangles, depth
// keys).  Translation units that include this header are compiled with
// -fmad=false and without fast-math: every operation is a single IEEE-754
// binary32 op in the order glm expands the reference's expressions
// (CSRC/utils.cuh:10-73,146-172,234-312), so the results are reproducible
// bit-for-bit by a plain C program.  rsqrt(x) is 1/sqrt(x).
#pragma once
#include <math_constants.h>
#include <stdint.h>

#if defined(__FAST_MATH__) || defined(__USE_FAST_MATH__)
#error "gs_math.cuh must not be compiled with --use_fast_math"
#endif

namespace gsm {

// Deterministic exp: only rint / fma / mul, identical on any IEEE machine.
// <= 2 ulp from the correctly rounded result.
__device__ __forceinline__ float det_expf(float x) {
    if (x > 88.72283f) return CUDART_INF_F;
    if (x < -103.0f) return 0.0f;
    const float LOG2E = 1.442695040888963f;
    const float LN2_HI = 0.693145751953125f;
    const float LN2_LO = 1.428606765330187e-06f;
    float n = rintf(x * LOG2E);
    float r = fmaf(-n, LN2_HI, x);
    r = fmaf(-n, LN2_LO, r);
    float p = 1.9841270e-4f;
    p = fmaf(p, r, 1.3888889e-3f);
    p = fmaf(p, r, 8.3333333e-3f);
    p = fmaf(p, r, 4.1666668e-2f);
    p = fmaf(p, r, 1.6666667e-1f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    int ni = (int)n;
    int n1 = ni / 2, n2 = ni - n1;
    float s1 = __int_as_float((n1 + 127) << 23);
    float s2 = __int_as_float((n2 + 127) << 23);
    return (p * s1) * s2;
}

__device__ __forceinline__ float det_sigmoidf(float x) { return 1.0f / (1.0f + det_expf(-x)); }

// Deterministic natural log (only integer ops, +, *, /, fma): x = m 2^e with m in [sqrt(1/2), sqrt(2)),
// log m = 2 atanh(s), s = (m-1)/(m+1), odd series to s^9 (|s| <= 0.1716 -> truncation < 1e-9).  ~2 ulp.
// The test oracle restates the same operation order on the CPU.
__device__ __forceinline__ float det_logf(float x) {
    if (!(x >= 0.0f)) return CUDART_NAN_F;  // negative or NaN
    if (x == 0.0f) return -CUDART_INF_F;
    if (x == CUDART_INF_F) return x;
    int bias = 0;
    if (x < 1.17549435e-38f) { x *= 8388608.0f; bias = -23; }  // subnormal
    int ix = __float_as_int(x);
    int e = (ix - 0x3f3504f3) >> 23;
    float m = __int_as_float(ix - (e << 23));
    float f = m - 1.0f;
    float s = f / (2.0f + f);
    float z = s * s;
    float p = 0.22222222f;
    p = fmaf(p, z, 0.2857143f);
    p = fmaf(p, z, 0.4f);
    p = fmaf(p, z, 0.6666667f);
    float logm = fmaf(s * z, p, 2.0f * s);
    float fe = (float)(e + bias);
    return fmaf(fe, 0.693145751953125f, fmaf(fe, 1.428606765330187e-06f, logm));
}

// F.normalize(q, dim=-1) with eps 1e-12 (strategies/base_gaussian_model.py:50)
__device__ __forceinline__ void normalize_quat(const float q[4], float out[4]) {
    float nn = sqrtf(((q[0] * q[0] + q[1] * q[1]) + q[2] * q[2]) + q[3] * q[3]);
    float d = fmaxf(nn, 1e-12f);
    out[0] = q[0] / d; out[1] = q[1] / d; out[2] = q[2] / d; out[3] = q[3] / d;
}

// CSRC/utils.cuh:11-29.  Row-major R[r*3+c].
__device__ __forceinline__ void quat_to_rotmat(const float quat[4], float R[9]) {
    float w = quat[0], x = quat[1], y = quat[2], z = quat[3];
    float inv_norm = 1.0f / sqrtf(((x * x + y * y) + z * z) + w * w);
    x *= inv_norm; y *= inv_norm; z *= inv_norm; w *= inv_norm;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.f - 2.f * (y2 + z2); R[1] = 2.f * (xy - wz); R[2] = 2.f * (xz + wy);
    R[3] = 2.f * (xy + wz); R[4] = 1.f - 2.f * (x2 + z2); R[5] = 2.f * (yz - wx);
    R[6] = 2.f * (xz - wy); R[7] = 2.f * (yz + wx); R[8] = 1.f - 2.f * (x2 + y2);
}

// CSRC/utils.cuh:55-65: covar = (R S)(R S)^T, all 9 entries (the float results
// of [r][c] and [c][r] are identical, so 6 are computed and mirrored).
__device__ __forceinline__ void quat_scale_to_covar(const float quat[4], const float scale[3],
                                                    float covar[9]) {
    float R[9], M[9];
    quat_to_rotmat(quat, R);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) M[r * 3 + c] = R[r * 3 + c] * scale[c];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = r; c < 3; ++c) {
            float v = (M[r * 3 + 0] * M[c * 3 + 0] + M[r * 3 + 1] * M[c * 3 + 1]) + M[r * 3 + 2] * M[c * 3 + 2];
            covar[r * 3 + c] = v;
            covar[c * 3 + r] = v;
        }
}

struct Camera {
    float R[9];  // row-major world->camera rotation
    float t[3];
    float fx, fy, cx, cy;
};

__device__ __forceinline__ void load_camera(const float *viewmat, const float *K, Camera &cam) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) cam.R[r * 3 + c] = viewmat[r * 4 + c];
        cam.t[r] = viewmat[r * 4 + 3];
    }
    cam.fx = K[0]; cam.fy = K[4]; cam.cx = K[2]; cam.cy = K[5];
}

struct Projected {
    float depth;
    float mean2d[2];
    float conic[3];
    float compensation;
    float radius;
};

// Returns true when the Gaussian passes every keep test of the reference.
// PACKED_RULE selects the radius formula: true  = CSRC/fully_fused_projection_packed_fwd.cu:117-119
//                                         false = CSRC/fully_fused_projection_fwd.cu:100-102
template <bool PACKED_RULE>
__device__ __forceinline__ bool project_one(const float mean[3], const float covar[9],
                                            const Camera &cam, int width, int height, float eps2d,
                                            float near_plane, float far_plane, float radius_clip,
                                            Projected &o) {
    const float *R = cam.R;
    float mc[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        mc[r] = ((R[r * 3 + 0] * mean[0] + R[r * 3 + 1] * mean[1]) + R[r * 3 + 2] * mean[2]) + cam.t[r];
    o.depth = mc[2];
    if (mc[2] < near_plane || mc[2] > far_plane) return false;

    // covar_c = (R * covar) * R^T   (CSRC/utils.cuh:257-261)
    float T1[9], cc[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            T1[i * 3 + j] = (R[i * 3 + 0] * covar[0 * 3 + j] + R[i * 3 + 1] * covar[1 * 3 + j]) + R[i * 3 + 2] * covar[2 * 3 + j];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            cc[i * 3 + j] = (T1[i * 3 + 0] * R[j * 3 + 0] + T1[i * 3 + 1] * R[j * 3 + 1]) + T1[i * 3 + 2] * R[j * 3 + 2];

    // persp_proj (CSRC/utils.cuh:146-172)
    float fx = cam.fx, fy = cam.fy;
    float x = mc[0], y = mc[1], z = mc[2];
    float tan_fovx = 0.5f * (float)width / fx;
    float tan_fovy = 0.5f * (float)height / fy;
    float lim_x = 1.3f * tan_fovx, lim_y = 1.3f * tan_fovy;
    float rz = 1.f / z, rz2 = rz * rz;
    float tx = z * fminf(lim_x, fmaxf(-lim_x, x * rz));
    float ty = z * fminf(lim_y, fmaxf(-lim_y, y * rz));
    float J00 = fx * rz, J11 = fy * rz;
    float J02 = (-fx * tx) * rz2, J12 = (-fy * ty) * rz2;
    float T00 = J00 * cc[0] + J02 * cc[6];
    float T01 = J00 * cc[1] + J02 * cc[7];
    float T02 = J00 * cc[2] + J02 * cc[8];
    float T10 = J11 * cc[3] + J12 * cc[6];
    float T11 = J11 * cc[4] + J12 * cc[7];
    float T12 = J11 * cc[5] + J12 * cc[8];
    float c00 = T00 * J00 + T02 * J02;
    float c01 = T01 * J11 + T02 * J12;
    float c10 = T10 * J00 + T12 * J02;
    float c11 = T11 * J11 + T12 * J12;
    o.mean2d[0] = (fx * x) * rz + cam.cx;
    o.mean2d[1] = (fy * y) * rz + cam.cy;

    // add_blur (CSRC/utils.cuh:303-312)
    float det_orig = c00 * c11 - c10 * c01;
    c00 += eps2d; c11 += eps2d;
    float det = c00 * c11 - c10 * c01;
    o.compensation = sqrtf(fmaxf(0.f, det_orig / det));
    if (det <= 0.f) return false;
    float inv_det = 1.f / det;
    o.conic[0] = c11 * inv_det;
    o.conic[1] = -c10 * inv_det;
    o.conic[2] = c00 * inv_det;

    float b = 0.5f * (c00 + c11);
    float radius;
    if (PACKED_RULE) {
        float sq = sqrtf(fmaxf(0.1f, b * b - det));
        float v1 = b + sq, v2 = b - sq;
        radius = ceilf(3.f * sqrtf(fmaxf(v1, v2)));
    } else {
        float v1 = b + sqrtf(fmaxf(0.01f, b * b - det));
        radius = ceilf(3.f * sqrtf(v1));
    }
    o.radius = radius;
    if (radius <= radius_clip) return false;
    if (o.mean2d[0] + radius <= 0 || o.mean2d[0] - radius >= (float)width ||
        o.mean2d[1] + radius <= 0 || o.mean2d[1] - radius >= (float)height)
        return false;
    return true;
}

// Tile rectangle of CSRC/isect_tiles.cu:55-64 ((uint32_t)floor(x) of a negative
// float saturates to 0 on the GPU; written here as a signed clamp).
__device__ __forceinline__ int clamp_tile(float v, int hi) {
    if (!(v > 0.f)) return 0;
    if (v >= (float)hi) return hi;
    return (int)v;
}
__device__ __forceinline__ void tile_rect(float mx, float my, int radius_i, int tile_size, int tw,
                                          int th, int &x0, int &y0, int &x1, int &y1) {
    float radius = (float)radius_i;
    float ts = (float)tile_size;
    float tile_radius = radius / ts;
    float tile_x = mx / ts, tile_y = my / ts;
    x0 = clamp_tile(floorf(tile_x - tile_radius), tw);
    y0 = clamp_tile(floorf(tile_y - tile_radius), th);
    x1 = clamp_tile(ceilf(tile_x + tile_radius), tw);
    y1 = clamp_tile(ceilf(tile_y + tile_radius), th);
}

}  // namespace gsm
