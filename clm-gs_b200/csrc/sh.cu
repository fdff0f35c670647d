// Spherical-harmonics colour, forward and backward.
// Math: Sloan, "Efficient Spherical Harmonic Evaluation" (JCGT 2013), the same
// polynomial basis the reference evaluates (CSRC/spherical_harmonics.cuh:11-91,
// CK/compute_sh_bwd.cu:14-273), restated as basis values b[k] and their
// partial derivatives.
//
// Layout: a row is K*3 floats (192 B for K=16).  A block stages ROWS rows
// through shared memory with coalesced 16-byte loads (the reference gives each
// of 3 threads per Gaussian a strided walk over the row), then one thread per
// Gaussian evaluates all three channels.
#include "common.cuh"

constexpr int kShRows = 128;  // Gaussians per block

template <int MAXDEG>
__device__ __forceinline__ void sh_basis(int degree, float x, float y, float z, float *b) {
    b[0] = 0.2820947917738781f;
    if (MAXDEG < 1 || degree < 1) return;
    b[1] = -0.48860251190292f * y; b[2] = 0.48860251190292f * z; b[3] = -0.48860251190292f * x;
    if (MAXDEG < 2 || degree < 2) return;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    float fC1 = x * x - y * y, fS1 = 2.f * x * y;
    b[4] = 0.5462742152960395f * fS1; b[5] = fTmp0B * y;
    b[6] = 0.9461746957575601f * z2 - 0.3153915652525201f;
    b[7] = fTmp0B * x; b[8] = 0.5462742152960395f * fC1;
    if (MAXDEG < 3 || degree < 3) return;
    float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    float fTmp1B = 1.445305721320277f * z;
    float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
    b[9] = -0.5900435899266435f * fS2; b[10] = fTmp1B * fS1; b[11] = fTmp0C * y;
    b[12] = z * (1.865881662950577f * z2 - 1.119528997770346f);
    b[13] = fTmp0C * x; b[14] = fTmp1B * fC1; b[15] = -0.5900435899266435f * fC2;
    if (MAXDEG < 4 || degree < 4) return;
    float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    float fTmp2B = -1.770130769779931f * z;
    float fC3 = x * fC2 - y * fS2, fS3 = x * fS2 + y * fC2;
    b[16] = 0.6258357354491763f * fS3; b[17] = fTmp2B * fS2; b[18] = fTmp1C * fS1;
    b[19] = fTmp0D * y;
    b[20] = 1.984313483298443f * z * b[12] - 1.006230589874905f * b[6];
    b[21] = fTmp0D * x; b[22] = fTmp1C * fC1; b[23] = fTmp2B * fC2;
    b[24] = 0.6258357354491763f * fC3;
}

// basis + d/dx, d/dy, d/dz (treating x,y,z as independent)
__device__ __forceinline__ void sh_basis_grad(int degree, float x, float y, float z, float *b, float *dx,
                                              float *dy, float *dz) {
#pragma unroll
    for (int k = 0; k < 25; ++k) dx[k] = dy[k] = dz[k] = 0.f;
    b[0] = 0.2820947917738781f;
    if (degree < 1) return;
    b[1] = -0.48860251190292f * y; b[2] = 0.48860251190292f * z; b[3] = -0.48860251190292f * x;
    dy[1] = -0.48860251190292f; dz[2] = 0.48860251190292f; dx[3] = -0.48860251190292f;
    if (degree < 2) return;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    float fC1 = x * x - y * y, fS1 = 2.f * x * y;
    float fC1_x = 2.f * x, fC1_y = -2.f * y, fS1_x = 2.f * y, fS1_y = 2.f * x;
    b[4] = 0.5462742152960395f * fS1; b[5] = fTmp0B * y;
    b[6] = 0.9461746957575601f * z2 - 0.3153915652525201f;
    b[7] = fTmp0B * x; b[8] = 0.5462742152960395f * fC1;
    dx[4] = 0.5462742152960395f * fS1_x; dy[4] = 0.5462742152960395f * fS1_y;
    dy[5] = fTmp0B; dz[5] = -1.092548430592079f * y;
    dz[6] = 2.f * 0.9461746957575601f * z;
    dx[7] = fTmp0B; dz[7] = -1.092548430592079f * x;
    dx[8] = 0.5462742152960395f * fC1_x; dy[8] = 0.5462742152960395f * fC1_y;
    if (degree < 3) return;
    float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    float fTmp1B = 1.445305721320277f * z;
    float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
    float fC2_x = fC1 + x * fC1_x - y * fS1_x, fC2_y = x * fC1_y - fS1 - y * fS1_y;
    float fS2_x = fS1 + x * fS1_x + y * fC1_x, fS2_y = x * fS1_y + fC1 + y * fC1_y;
    float fTmp0C_z = -2.285228997322329f * 2.f * z;
    float b12_z = 3.f * 1.865881662950577f * z2 - 1.119528997770346f;
    b[9] = -0.5900435899266435f * fS2; b[10] = fTmp1B * fS1; b[11] = fTmp0C * y;
    b[12] = z * (1.865881662950577f * z2 - 1.119528997770346f);
    b[13] = fTmp0C * x; b[14] = fTmp1B * fC1; b[15] = -0.5900435899266435f * fC2;
    dx[9] = -0.5900435899266435f * fS2_x; dy[9] = -0.5900435899266435f * fS2_y;
    dx[10] = fTmp1B * fS1_x; dy[10] = fTmp1B * fS1_y; dz[10] = 1.445305721320277f * fS1;
    dy[11] = fTmp0C; dz[11] = fTmp0C_z * y;
    dz[12] = b12_z;
    dx[13] = fTmp0C; dz[13] = fTmp0C_z * x;
    dx[14] = fTmp1B * fC1_x; dy[14] = fTmp1B * fC1_y; dz[14] = 1.445305721320277f * fC1;
    dx[15] = -0.5900435899266435f * fC2_x; dy[15] = -0.5900435899266435f * fC2_y;
    if (degree < 4) return;
    float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    float fTmp2B = -1.770130769779931f * z;
    float fC3 = x * fC2 - y * fS2, fS3 = x * fS2 + y * fC2;
    float fTmp0D_z = 3.f * -4.683325804901025f * z2 + 2.007139630671868f;
    float fTmp1C_z = 2.f * 3.31161143515146f * z;
    float fC3_x = fC2 + x * fC2_x - y * fS2_x, fC3_y = x * fC2_y - fS2 - y * fS2_y;
    float fS3_x = fS2 + y * fC2_x + x * fS2_x, fS3_y = x * fS2_y + fC2 + y * fC2_y;
    b[16] = 0.6258357354491763f * fS3; b[17] = fTmp2B * fS2; b[18] = fTmp1C * fS1;
    b[19] = fTmp0D * y;
    b[20] = 1.984313483298443f * z * b[12] - 1.006230589874905f * b[6];
    b[21] = fTmp0D * x; b[22] = fTmp1C * fC1; b[23] = fTmp2B * fC2;
    b[24] = 0.6258357354491763f * fC3;
    dx[16] = 0.6258357354491763f * fS3_x; dy[16] = 0.6258357354491763f * fS3_y;
    dx[17] = fTmp2B * fS2_x; dy[17] = fTmp2B * fS2_y; dz[17] = -1.770130769779931f * fS2;
    dx[18] = fTmp1C * fS1_x; dy[18] = fTmp1C * fS1_y; dz[18] = fTmp1C_z * fS1;
    dy[19] = fTmp0D; dz[19] = fTmp0D_z * y;
    dz[20] = 1.984313483298443f * (b[12] + z * b12_z) - 1.006230589874905f * dz[6];
    dx[21] = fTmp0D; dz[21] = fTmp0D_z * x;
    dx[22] = fTmp1C * fC1_x; dy[22] = fTmp1C * fC1_y; dz[22] = fTmp1C_z * fC1;
    dx[23] = fTmp2B * fC2_x; dy[23] = fTmp2B * fC2_y; dz[23] = -1.770130769779931f * fC2;
    dx[24] = 0.6258357354491763f * fC3_x; dy[24] = 0.6258357354491763f * fC3_y;
}

// Visitor form: f(k, b_k, db_k/dx, db_k/dy, db_k/dz) is called for k = 0 .. (degree+1)^2-1 as soon as the
// value is known, so a consumer that uses each basis once keeps ~20 live registers instead of 4 x 25.
template <typename F>
__device__ __forceinline__ void sh_visit(int degree, float x, float y, float z, F &&f) {
    f(0, 0.2820947917738781f, 0.f, 0.f, 0.f);
    if (degree < 1) return;
    f(1, -0.48860251190292f * y, 0.f, -0.48860251190292f, 0.f);
    f(2, 0.48860251190292f * z, 0.f, 0.f, 0.48860251190292f);
    f(3, -0.48860251190292f * x, -0.48860251190292f, 0.f, 0.f);
    if (degree < 2) return;
    const float z2 = z * z;
    const float fTmp0B = -1.092548430592079f * z;
    const float fC1 = x * x - y * y, fS1 = 2.f * x * y;
    const float fC1_x = 2.f * x, fC1_y = -2.f * y, fS1_x = 2.f * y, fS1_y = 2.f * x;
    const float b6 = 0.9461746957575601f * z2 - 0.3153915652525201f;
    const float b6_z = 2.f * 0.9461746957575601f * z;
    f(4, 0.5462742152960395f * fS1, 0.5462742152960395f * fS1_x, 0.5462742152960395f * fS1_y, 0.f);
    f(5, fTmp0B * y, 0.f, fTmp0B, -1.092548430592079f * y);
    f(6, b6, 0.f, 0.f, b6_z);
    f(7, fTmp0B * x, fTmp0B, 0.f, -1.092548430592079f * x);
    f(8, 0.5462742152960395f * fC1, 0.5462742152960395f * fC1_x, 0.5462742152960395f * fC1_y, 0.f);
    if (degree < 3) return;
    const float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    const float fTmp1B = 1.445305721320277f * z;
    const float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
    const float fC2_x = fC1 + x * fC1_x - y * fS1_x, fC2_y = x * fC1_y - fS1 - y * fS1_y;
    const float fS2_x = fS1 + x * fS1_x + y * fC1_x, fS2_y = x * fS1_y + fC1 + y * fC1_y;
    const float fTmp0C_z = -2.285228997322329f * 2.f * z;
    const float b12 = z * (1.865881662950577f * z2 - 1.119528997770346f);
    const float b12_z = 3.f * 1.865881662950577f * z2 - 1.119528997770346f;
    f(9, -0.5900435899266435f * fS2, -0.5900435899266435f * fS2_x, -0.5900435899266435f * fS2_y, 0.f);
    f(10, fTmp1B * fS1, fTmp1B * fS1_x, fTmp1B * fS1_y, 1.445305721320277f * fS1);
    f(11, fTmp0C * y, 0.f, fTmp0C, fTmp0C_z * y);
    f(12, b12, 0.f, 0.f, b12_z);
    f(13, fTmp0C * x, fTmp0C, 0.f, fTmp0C_z * x);
    f(14, fTmp1B * fC1, fTmp1B * fC1_x, fTmp1B * fC1_y, 1.445305721320277f * fC1);
    f(15, -0.5900435899266435f * fC2, -0.5900435899266435f * fC2_x, -0.5900435899266435f * fC2_y, 0.f);
    if (degree < 4) return;
    const float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    const float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    const float fTmp2B = -1.770130769779931f * z;
    const float fC3 = x * fC2 - y * fS2, fS3 = x * fS2 + y * fC2;
    const float fTmp0D_z = 3.f * -4.683325804901025f * z2 + 2.007139630671868f;
    const float fTmp1C_z = 2.f * 3.31161143515146f * z;
    const float fC3_x = fC2 + x * fC2_x - y * fS2_x, fC3_y = x * fC2_y - fS2 - y * fS2_y;
    const float fS3_x = fS2 + y * fC2_x + x * fS2_x, fS3_y = x * fS2_y + fC2 + y * fC2_y;
    f(16, 0.6258357354491763f * fS3, 0.6258357354491763f * fS3_x, 0.6258357354491763f * fS3_y, 0.f);
    f(17, fTmp2B * fS2, fTmp2B * fS2_x, fTmp2B * fS2_y, -1.770130769779931f * fS2);
    f(18, fTmp1C * fS1, fTmp1C * fS1_x, fTmp1C * fS1_y, fTmp1C_z * fS1);
    f(19, fTmp0D * y, 0.f, fTmp0D, fTmp0D_z * y);
    f(20, 1.984313483298443f * z * b12 - 1.006230589874905f * b6, 0.f, 0.f,
      1.984313483298443f * (b12 + z * b12_z) - 1.006230589874905f * b6_z);
    f(21, fTmp0D * x, fTmp0D, 0.f, fTmp0D_z * x);
    f(22, fTmp1C * fC1, fTmp1C * fC1_x, fTmp1C * fC1_y, fTmp1C_z * fC1);
    f(23, fTmp2B * fC2, fTmp2B * fC2_x, fTmp2B * fC2_y, -1.770130769779931f * fC2);
    f(24, 0.6258357354491763f * fC3, 0.6258357354491763f * fC3_x, 0.6258357354491763f * fC3_y, 0.f);
}

// Stage `rows` consecutive rows of `row_len` floats into smem with padded stride.
__device__ __forceinline__ void stage_rows(float *s, int stride, const float *g, int rows, int row_len) {
    int total = rows * row_len;
    if ((((uintptr_t)g) & 15) == 0 && (row_len & 3) == 0) {
        const float4 *g4 = reinterpret_cast<const float4 *>(g);
        for (int i = threadIdx.x; i < total / 4; i += blockDim.x) {
            float4 v = __ldg(g4 + i);
            int e = i * 4, r = e / row_len, c = e - r * row_len;
            float *d = s + r * stride + c;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
    } else {
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            int r = i / row_len, c = i - r * row_len;
            s[r * stride + c] = __ldg(g + i);
        }
    }
}

template <bool ENGINE>
__global__ void __launch_bounds__(kShRows)
sh_fwd_kernel(int64_t N, int K, int degree, const float *__restrict__ dirs, const float *__restrict__ campos,
              const float *__restrict__ coeffs, const uint8_t *__restrict__ masks, float *__restrict__ colors) {
    extern __shared__ float s_rows[];
    const int row_len = K * 3, stride = row_len | 1;
    int64_t base = (int64_t)blockIdx.x * kShRows;
    int rows = (int)min((int64_t)kShRows, N - base);
    stage_rows(s_rows, stride, coeffs + base * row_len, rows, row_len);
    __syncthreads();
    if ((int)threadIdx.x >= rows) return;
    int64_t i = base + threadIdx.x;
    if (masks && !masks[i]) return;
    float x = 0.f, y = 0.f, z = 0.f;
    if (degree >= 1) {
        float dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
        if (ENGINE) { dx -= campos[0]; dy -= campos[1]; dz -= campos[2]; }
        float inorm = rsqrtf(dx * dx + dy * dy + dz * dz);
        x = dx * inorm; y = dy * inorm; z = dz * inorm;
    }
    float b[25];
    sh_basis<4>(degree, x, y, z, b);
    const int nb = (degree + 1) * (degree + 1);
    const float *row = s_rows + threadIdx.x * stride;
    float r = 0.f, g = 0.f, bl = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
        if (k < nb) {
            r += b[k] * row[3 * k];
            g += b[k] * row[3 * k + 1];
            bl += b[k] * row[3 * k + 2];
        }
    }
    if (ENGINE) { r = fmaxf(r + 0.5f, 0.f); g = fmaxf(g + 0.5f, 0.f); bl = fmaxf(bl + 0.5f, 0.f); }
    colors[3 * i] = r; colors[3 * i + 1] = g; colors[3 * i + 2] = bl;
}

CLMGS_EXPORT int clmgs_sh_fwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                              const uint8_t *masks, float *colors, void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 4, "sh_fwd: degree %d not in [0,4]", degree);
    CLMGS_REQUIRE(K >= (degree + 1) * (degree + 1) && K <= 25, "sh_fwd: K=%d too small for degree %d (or > 25)", K, degree);
    if (N == 0) return CLMGS_OK;
    size_t smem = (size_t)kShRows * ((K * 3) | 1) * sizeof(float);
    sh_fwd_kernel<false><<<(unsigned)ceil_div64(N, kShRows), kShRows, smem, as_stream(stream)>>>(
        N, K, degree, dirs, nullptr, coeffs, masks, colors);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

__global__ void sh_colors_fwd_lanes_kernel(int64_t N, int degree, const float *__restrict__ means,
                                           const float *__restrict__ campos, const float4 *__restrict__ coeffs,
                                           const int64_t *__restrict__ ids, const int32_t *__restrict__ slot_of,
                                           float *__restrict__ colors);

CLMGS_EXPORT int clmgs_sh_colors_fwd(int64_t N, int K, int degree, const float *means, const float *campos,
                                     const float *coeffs, float *colors, void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 4, "sh_colors_fwd: degree %d not in [0,4]", degree);
    CLMGS_REQUIRE(K >= (degree + 1) * (degree + 1) && K <= 25, "sh_colors_fwd: K=%d too small for degree %d (or > 25)", K, degree);
    if (N == 0) return CLMGS_OK;
    if (K == 16 && degree <= 3 && (((uintptr_t)coeffs) & 15) == 0) {
        // same 4-lane evaluation order as the backward kernel, so both agree on the clamp mask bit for bit
        sh_colors_fwd_lanes_kernel<<<(unsigned)ceil_div64(N * 4, 256), 256, 0, as_stream(stream)>>>(
            N, degree, means, campos, reinterpret_cast<const float4 *>(coeffs), nullptr, nullptr, colors);
        CLMGS_LAUNCH_OK();
        return CLMGS_OK;
    }
    size_t smem = (size_t)kShRows * ((K * 3) | 1) * sizeof(float);
    sh_fwd_kernel<true><<<(unsigned)ceil_div64(N, kShRows), kShRows, smem, as_stream(stream)>>>(
        N, K, degree, means, campos, coeffs, nullptr, colors);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// Backward: rows of coeffs AND of v_coeffs go through shared memory so that
// both the read and the read-modify-write of the 192-B rows are coalesced.
template <bool ENGINE>
__global__ void __launch_bounds__(kShRows, 8)  // <= 64 registers: 8 CTAs/SM (8 x 24.5 KB smem) keeps enough rows in flight
sh_bwd_kernel(int64_t N, int K, int degree, const float *__restrict__ dirs, const float *__restrict__ campos,
              const float *__restrict__ coeffs, const uint8_t *__restrict__ masks,
              const float *__restrict__ v_colors, float *__restrict__ v_coeffs, int accumulate,
              float *__restrict__ v_dirs) {
    extern __shared__ float s_rows[];
    const int row_len = K * 3, stride = row_len | 1;
    int64_t base = (int64_t)blockIdx.x * kShRows;
    int rows = (int)min((int64_t)kShRows, N - base);
    stage_rows(s_rows, stride, coeffs + base * row_len, rows, row_len);
    __syncthreads();
    const int nb = (degree + 1) * (degree + 1);
    bool active = (int)threadIdx.x < rows;
    int64_t i = base + threadIdx.x;
    if (active && masks && !masks[i]) active = false;
    float *row = s_rows + threadIdx.x * stride;
    if ((int)threadIdx.x < rows) {
        if (active) {
            float x = 0.f, y = 0.f, z = 0.f, inorm = 0.f;
            if (degree >= 1) {
                float dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
                if (ENGINE) { dx -= campos[0]; dy -= campos[1]; dz -= campos[2]; }
                inorm = rsqrtf(dx * dx + dy * dy + dz * dz);
                x = dx * inorm; y = dy * inorm; z = dz * inorm;
            }
            float vr = v_colors[3 * i], vg = v_colors[3 * i + 1], vb = v_colors[3 * i + 2];
            if (ENGINE) {  // gradient of clamp_min(c + 0.5, 0): passes where c + 0.5 >= 0
                float cr = 0.f, cg = 0.f, cb = 0.f;
                sh_visit(degree, x, y, z, [&](int k, float b, float, float, float) {
                    cr += b * row[3 * k]; cg += b * row[3 * k + 1]; cb += b * row[3 * k + 2];
                });
                if (cr + 0.5f < 0.f) vr = 0.f;
                if (cg + 0.5f < 0.f) vg = 0.f;
                if (cb + 0.5f < 0.f) vb = 0.f;
            }
            float vx = 0.f, vy = 0.f, vz = 0.f;
            sh_visit(degree, x, y, z, [&](int k, float b, float dx, float dy, float dz) {
                const float cv = row[3 * k] * vr + row[3 * k + 1] * vg + row[3 * k + 2] * vb;
                vx += dx * cv; vy += dy * cv; vz += dz * cv;
                // overwrite the staged coefficient with its gradient contribution
                row[3 * k] = b * vr; row[3 * k + 1] = b * vg; row[3 * k + 2] = b * vb;
            });
            for (int k = nb; k < K; ++k) row[3 * k] = row[3 * k + 1] = row[3 * k + 2] = 0.f;
            if (v_dirs) {
                if (degree >= 1) {
                    float dot = vx * x + vy * y + vz * z;
                    if (ENGINE) {  // dirs = means - campos: v_means += v_dirs
                        v_dirs[3 * i] += (vx - dot * x) * inorm;
                        v_dirs[3 * i + 1] += (vy - dot * y) * inorm;
                        v_dirs[3 * i + 2] += (vz - dot * z) * inorm;
                    } else {
                        v_dirs[3 * i] = (vx - dot * x) * inorm;
                        v_dirs[3 * i + 1] = (vy - dot * y) * inorm;
                        v_dirs[3 * i + 2] = (vz - dot * z) * inorm;
                    }
                } else if (ENGINE) {
                } else {
                    v_dirs[3 * i] = v_dirs[3 * i + 1] = v_dirs[3 * i + 2] = 0.f;
                }
            }
        } else {
            for (int k = 0; k < row_len; ++k) row[k] = 0.f;
            if (v_dirs && !ENGINE) v_dirs[3 * i] = v_dirs[3 * i + 1] = v_dirs[3 * i + 2] = 0.f;
        }
    }
    __syncthreads();
    // coalesced write-back of the staged gradient rows
    float *out = v_coeffs + base * row_len;
    int total = rows * row_len;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        int r = e / row_len, c = e - r * row_len;
        float v = s_rows[r * stride + c];
        if (accumulate) out[e] += v;
        else if (!masks || masks[base + r]) out[e] = v;
    }
}

CLMGS_EXPORT int clmgs_sh_bwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                              const uint8_t *masks, const float *v_colors, float *v_coeffs, int accumulate,
                              float *v_dirs, void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 4, "sh_bwd: degree %d not in [0,4]", degree);
    CLMGS_REQUIRE(K >= (degree + 1) * (degree + 1) && K <= 25, "sh_bwd: K=%d too small for degree %d (or > 25)", K, degree);
    if (N == 0) return CLMGS_OK;
    size_t smem = (size_t)kShRows * ((K * 3) | 1) * sizeof(float);
    sh_bwd_kernel<false><<<(unsigned)ceil_div64(N, kShRows), kShRows, smem, as_stream(stream)>>>(
        N, K, degree, dirs, nullptr, coeffs, masks, v_colors, v_coeffs, accumulate, v_dirs);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// Engine backward for K = 16 (degree <= 3): four lanes per Gaussian, lane j owns bases 4j..4j+3 = 12
// consecutive floats = three float4 of the 192-B row.  No shared memory, no block barriers: the row of
// coefficients is read and the row of gradients read-modified-written with 16-byte accesses that are
// contiguous across the warp (8 Gaussians = 1536 B), the colour (for the clamp mask) and the direction
// gradient are 4-lane shuffle reductions.  Each lane evaluates the whole basis but keeps only its four.
__global__ void __launch_bounds__(256)
sh_colors_bwd_lanes_kernel(int64_t N, int degree, const float *__restrict__ means, const float *__restrict__ campos,
                           const float4 *__restrict__ coeffs, const float *__restrict__ v_colors,
                           float4 *__restrict__ v_coeffs, float *__restrict__ v_means,
                           const int64_t *__restrict__ ids, const int32_t *__restrict__ slot_of) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t g = tid >> 2;
    const int j = (int)(tid & 3);
    if (g >= N) return;  // N*4 threads: a whole 4-lane group leaves together
    const unsigned gmask = 0xfu << ((threadIdx.x & 31) & ~3);
    float x = 0.f, y = 0.f, z = 0.f, inorm = 0.f;
    if (degree >= 1) {
        float dx = __ldg(means + 3 * g) - campos[0], dy = __ldg(means + 3 * g + 1) - campos[1],
              dz = __ldg(means + 3 * g + 2) - campos[2];
        inorm = rsqrtf(dx * dx + dy * dy + dz * dz);
        x = dx * inorm; y = dy * inorm; z = dz * inorm;
    }
    float mb[4] = {0.f, 0.f, 0.f, 0.f}, mx[4] = {0.f, 0.f, 0.f, 0.f}, my[4] = {0.f, 0.f, 0.f, 0.f},
          mz[4] = {0.f, 0.f, 0.f, 0.f};
    sh_visit(degree, x, y, z, [&](int k, float b, float bx, float by, float bz) {
        if ((k >> 2) == j) { mb[k & 3] = b; mx[k & 3] = bx; my[k & 3] = by; mz[k & 3] = bz; }
    });
    const int64_t row = slot_of ? (int64_t)slot_of[ids[g]] : g;  // batch-resident table: row = slot of the Gaussian
    const float4 *crow = coeffs + row * 12 + j * 3;
    float cf[12];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const float4 v = __ldg(crow + q);
        cf[4 * q] = v.x; cf[4 * q + 1] = v.y; cf[4 * q + 2] = v.z; cf[4 * q + 3] = v.w;
    }
    // colour -> clamp mask (gradient of clamp_min(c + 0.5, 0) passes where c + 0.5 >= 0)
    float cr = 0.f, cg = 0.f, cb = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) { cr += mb[t] * cf[3 * t]; cg += mb[t] * cf[3 * t + 1]; cb += mb[t] * cf[3 * t + 2]; }
#pragma unroll
    for (int o = 1; o < 4; o <<= 1) {
        cr += __shfl_xor_sync(gmask, cr, o); cg += __shfl_xor_sync(gmask, cg, o); cb += __shfl_xor_sync(gmask, cb, o);
    }
    float vr = __ldg(v_colors + 3 * g), vg = __ldg(v_colors + 3 * g + 1), vb = __ldg(v_colors + 3 * g + 2);
    if (cr + 0.5f < 0.f) vr = 0.f;
    if (cg + 0.5f < 0.f) vg = 0.f;
    if (cb + 0.5f < 0.f) vb = 0.f;
    float vx = 0.f, vy = 0.f, vz = 0.f, gf[12];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const float cv = cf[3 * t] * vr + cf[3 * t + 1] * vg + cf[3 * t + 2] * vb;
        vx += mx[t] * cv; vy += my[t] * cv; vz += mz[t] * cv;
        gf[3 * t] = mb[t] * vr; gf[3 * t + 1] = mb[t] * vg; gf[3 * t + 2] = mb[t] * vb;
    }
    float4 *grow = v_coeffs + row * 12 + j * 3;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        float4 o = grow[q];
        o.x += gf[4 * q]; o.y += gf[4 * q + 1]; o.z += gf[4 * q + 2]; o.w += gf[4 * q + 3];
        grow[q] = o;
    }
    if (degree >= 1) {
#pragma unroll
        for (int o = 1; o < 4; o <<= 1) {
            vx += __shfl_xor_sync(gmask, vx, o); vy += __shfl_xor_sync(gmask, vy, o); vz += __shfl_xor_sync(gmask, vz, o);
        }
        if (j == 0) {
            const float dot = vx * x + vy * y + vz * z;
            v_means[3 * g] += (vx - dot * x) * inorm;
            v_means[3 * g + 1] += (vy - dot * y) * inorm;
            v_means[3 * g + 2] += (vz - dot * z) * inorm;
        }
    }
}

CLMGS_EXPORT int clmgs_sh_colors_bwd(int64_t N, int K, int degree, const float *means, const float *campos,
                                     const float *coeffs, const float *v_colors, float *v_coeffs, float *v_means,
                                     void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 4, "sh_colors_bwd: degree %d not in [0,4]", degree);
    CLMGS_REQUIRE(K >= (degree + 1) * (degree + 1) && K <= 25, "sh_colors_bwd: K=%d too small for degree %d (or > 25)", K, degree);
    if (N == 0) return CLMGS_OK;
    if (K == 16 && degree <= 3 && ((((uintptr_t)coeffs) | ((uintptr_t)v_coeffs)) & 15) == 0) {
        sh_colors_bwd_lanes_kernel<<<(unsigned)ceil_div64(N * 4, 256), 256, 0, as_stream(stream)>>>(
            N, degree, means, campos, reinterpret_cast<const float4 *>(coeffs), v_colors,
            reinterpret_cast<float4 *>(v_coeffs), v_means, nullptr, nullptr);
        CLMGS_LAUNCH_OK();
        return CLMGS_OK;
    }
    size_t smem = (size_t)kShRows * ((K * 3) | 1) * sizeof(float);
    sh_bwd_kernel<true><<<(unsigned)ceil_div64(N, kShRows), kShRows, smem, as_stream(stream)>>>(
        N, K, degree, means, campos, coeffs, nullptr, v_colors, v_coeffs, 1, v_means);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

// Forward in the same 4-lanes-per-Gaussian form (no shared memory), rows addressed through the
// batch's slot table.
__global__ void __launch_bounds__(256)
sh_colors_fwd_lanes_kernel(int64_t N, int degree, const float *__restrict__ means, const float *__restrict__ campos,
                           const float4 *__restrict__ coeffs, const int64_t *__restrict__ ids,
                           const int32_t *__restrict__ slot_of, float *__restrict__ colors) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t g = tid >> 2;
    const int j = (int)(tid & 3);
    if (g >= N) return;
    const unsigned gmask = 0xfu << ((threadIdx.x & 31) & ~3);
    float x = 0.f, y = 0.f, z = 0.f;
    if (degree >= 1) {
        float dx = __ldg(means + 3 * g) - campos[0], dy = __ldg(means + 3 * g + 1) - campos[1],
              dz = __ldg(means + 3 * g + 2) - campos[2];
        const float inorm = rsqrtf(dx * dx + dy * dy + dz * dz);
        x = dx * inorm; y = dy * inorm; z = dz * inorm;
    }
    float mb[4] = {0.f, 0.f, 0.f, 0.f};
    sh_visit(degree, x, y, z, [&](int k, float b, float, float, float) {
        if ((k >> 2) == j) mb[k & 3] = b;
    });
    const int64_t row = slot_of ? (int64_t)slot_of[ids[g]] : g;
    const float4 *crow = coeffs + row * 12 + j * 3;
    float cf[12];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const float4 v = __ldg(crow + q);
        cf[4 * q] = v.x; cf[4 * q + 1] = v.y; cf[4 * q + 2] = v.z; cf[4 * q + 3] = v.w;
    }
    float cr = 0.f, cg = 0.f, cb = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) { cr += mb[t] * cf[3 * t]; cg += mb[t] * cf[3 * t + 1]; cb += mb[t] * cf[3 * t + 2]; }
#pragma unroll
    for (int o = 1; o < 4; o <<= 1) {
        cr += __shfl_xor_sync(gmask, cr, o); cg += __shfl_xor_sync(gmask, cg, o); cb += __shfl_xor_sync(gmask, cb, o);
    }
    if (j < 3) colors[3 * g + j] = fmaxf((j == 0 ? cr : (j == 1 ? cg : cb)) + 0.5f, 0.f);
}

CLMGS_EXPORT int clmgs_sh_colors_fwd_slots(int64_t N, int degree, const float *means, const float *campos,
                                           const float *d_rows, const int64_t *ids, const int32_t *slot_of,
                                           float *colors, void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 3, "sh_colors_fwd_slots: degree %d not in [0,3] (K = 16 rows)", degree);
    CLMGS_REQUIRE((((uintptr_t)d_rows) & 15) == 0, "sh_colors_fwd_slots: rows must be 16-byte aligned");
    if (N == 0) return CLMGS_OK;
    sh_colors_fwd_lanes_kernel<<<(unsigned)ceil_div64(N * 4, 256), 256, 0, as_stream(stream)>>>(
        N, degree, means, campos, reinterpret_cast<const float4 *>(d_rows), ids, slot_of, colors);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_sh_colors_bwd_slots(int64_t N, int degree, const float *means, const float *campos,
                                           const float *d_rows, const float *v_colors, float *d_grad_rows,
                                           const int64_t *ids, const int32_t *slot_of, float *v_means, void *stream) {
    CLMGS_REQUIRE(degree >= 0 && degree <= 3, "sh_colors_bwd_slots: degree %d not in [0,3] (K = 16 rows)", degree);
    CLMGS_REQUIRE(((((uintptr_t)d_rows) | ((uintptr_t)d_grad_rows)) & 15) == 0, "sh_colors_bwd_slots: rows must be 16-byte aligned");
    if (N == 0) return CLMGS_OK;
    sh_colors_bwd_lanes_kernel<<<(unsigned)ceil_div64(N * 4, 256), 256, 0, as_stream(stream)>>>(
        N, degree, means, campos, reinterpret_cast<const float4 *>(d_rows), v_colors,
        reinterpret_cast<float4 *>(d_grad_rows), v_means, ids, slot_of);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
