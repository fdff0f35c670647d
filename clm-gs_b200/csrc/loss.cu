// A9 loss: 0.8*L1 + 0.2*(1 - SSIM) forward and backward in two tiled kernels.
// SSIM = 11-tap separable Gaussian (sigma 1.5), zero "same" padding, C1=1e-4,
// C2=9e-4: the fused kernels of clm_kernels (CK/ssim.cu:187-366) plus the
// torch.compile'd loss_combined (strategies/base_engine.py:79-94) in one place.
//
// Pass 1 (per plane, 32x32 output tile + 5-pixel halo in shared memory; all five moments in one horizontal and
//   one vertical sweep):
//   mu1, mu2, E[x^2], E[y^2], E[xy] -> ssim value and the three partial
//   derivative maps, already multiplied by d loss / d ssim_map; block-reduced
//   SSIM and L1 sums go to two double accumulators.
// Pass 2: convolves the three maps back and adds the L1 sign term.
#include "common.cuh"

namespace {

constexpr int TB = 32;          // output tile edge
constexpr int HALO = 5;
constexpr int TS = TB + 2 * HALO;  // 42

__constant__ float c_gauss[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f,
                                  0.10936068743467331f,  0.21300552785396576f,   0.26601171493530273f,
                                  0.21300552785396576f,  0.10936068743467331f,   0.036000773310661316f,
                                  0.0075987582094967365f, 0.001028380123898387f};

// A "plane" is one channel of one image: element (y,x) lives at base[y*row_stride + x*pix_stride].
struct Plane {
    const float *base;
    int64_t row_stride;
    int pix_stride;
};
struct PlaneOut {
    float *base;
    int64_t row_stride;
    int pix_stride;
};

__device__ __forceinline__ float conv_h(const float (*tile)[TS + 1], int y, int x) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) v += c_gauss[k] * tile[y][x + k];
    return v;
}

template <typename GT>
__device__ __forceinline__ float load_gt(const GT *p);
template <>
__device__ __forceinline__ float load_gt<uint8_t>(const uint8_t *p) {
    return fminf(fmaxf((float)(*p) / 255.0f, 0.f), 1.f);  // base_engine.py:92
}
template <>
__device__ __forceinline__ float load_gt<float>(const float *p) { return *p; }

// grid: (tiles_x, tiles_y, n_planes)
template <typename GT>
__global__ void __launch_bounds__(TB *TB)
ssim_pass1_kernel(int H, int W, const float *img, int64_t img_plane_stride, int64_t img_row_stride,
                  int img_pix_stride, const GT *gt /* planar [P,H,W] */, float dmap /* dL/d ssim_map elem */,
                  float *maps /* [P][3][H][W] */, double *sums /* [0]=ssim, [1]=l1 */, int want_maps) {
    __shared__ float s1[TS][TS + 1];
    __shared__ float s2[TS][TS + 1];
    __shared__ float tmp5[5][TS][TB + 1];
    __shared__ float red[2][32];
    const int tid = threadIdx.y * TB + threadIdx.x;
    const int plane = blockIdx.z;
    const int y0 = blockIdx.y * TB - HALO, x0 = blockIdx.x * TB - HALO;
    const float *ip = img + plane * img_plane_stride;
    const GT *gp = gt + (int64_t)plane * H * W;
    for (int e = tid; e < TS * TS; e += TB * TB) {
        int ly = e / TS, lx = e - ly * TS;
        int y = y0 + ly, x = x0 + lx;
        float a = 0.f, b = 0.f;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            a = ip[(int64_t)y * img_row_stride + (int64_t)x * img_pix_stride];
            b = load_gt<GT>(gp + (int64_t)y * W + x);
        }
        s1[ly][lx] = a;
        s2[ly][lx] = b;
    }
    // all five moments in ONE separable pass: the horizontal sweep forms x1^2, x2^2, x1 x2 on the fly (same
    // operations, same order per map as five separate convolutions, so results are unchanged), two barriers in all
    __syncthreads();
    for (int e = tid; e < TS * TB; e += TB * TB) {
        const int y = e / TB, x = e - y * TB;
        float h1 = 0.f, h2 = 0.f, h11 = 0.f, h22 = 0.f, h12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float a = s1[y][x + k], b = s2[y][x + k], w = c_gauss[k];
            h1 += w * a; h2 += w * b; h11 += w * (a * a); h22 += w * (b * b); h12 += w * (a * b);
        }
        tmp5[0][y][x] = h1; tmp5[1][y][x] = h2; tmp5[2][y][x] = h11; tmp5[3][y][x] = h22; tmp5[4][y][x] = h12;
    }
    __syncthreads();
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
    {
        const int ty = threadIdx.y, tx = threadIdx.x;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float w = c_gauss[k];
            mu1 += w * tmp5[0][ty + k][tx]; mu2 += w * tmp5[1][ty + k][tx]; e11 += w * tmp5[2][ty + k][tx];
            e22 += w * tmp5[3][ty + k][tx]; e12 += w * tmp5[4][ty + k][tx];
        }
    }

    const int py = blockIdx.y * TB + threadIdx.y, px = blockIdx.x * TB + threadIdx.x;
    const bool inside = py < H && px < W;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float sig1 = e11 - mu1 * mu1, sig2 = e22 - mu2 * mu2, sig12 = e12 - mu1 * mu2;
    const float Cc = 2.0f * mu1 * mu2 + C1, D = 2.0f * sig12 + C2;
    const float A = mu1 * mu1 + mu2 * mu2 + C1, B = sig1 + sig2 + C2;
    float ssim_v = 0.f, l1_v = 0.f;
    if (inside) {
        ssim_v = (Cc * D) / (A * B);
        l1_v = fabsf(s1[threadIdx.y + HALO][threadIdx.x + HALO] - s2[threadIdx.y + HALO][threadIdx.x + HALO]);
        if (want_maps) {
            const int64_t HW = (int64_t)H * W;
            float *m = maps + (int64_t)plane * 3 * HW + (int64_t)py * W + px;
            m[0] = dmap * ((mu2 * 2.0f * D) / (A * B) - (mu2 * 2.0f * Cc) / (A * B) - (mu1 * 2.0f * Cc * D) / (A * A * B) +
                           (mu1 * 2.0f * Cc * D) / (A * B * B));
            m[HW] = dmap * ((-Cc * D) / (A * B * B));
            m[2 * HW] = dmap * ((2.0f * Cc) / (A * B));
        }
    }
    // block reduction of the two sums
    float a = warp_sum(ssim_v), b = warp_sum(l1_v);
    if ((tid & 31) == 0) { red[0][tid >> 5] = a; red[1][tid >> 5] = b; }
    __syncthreads();
    if (tid < 32) {
        a = warp_sum(red[0][tid]);
        b = warp_sum(red[1][tid]);
        if (tid == 0) { atomicAdd(&sums[0], (double)a); atomicAdd(&sums[1], (double)b); }
    }
}

template <typename GT>
__global__ void __launch_bounds__(TB *TB)
ssim_pass2_kernel(int H, int W, const float *img, int64_t img_plane_stride, int64_t img_row_stride,
                  int img_pix_stride, const GT *gt, const float *maps, float l1w, float *v_img) {
    __shared__ float sm[TS][TS + 1];
    __shared__ float tmp3[3][TS][TB + 1];
    const int tid = threadIdx.y * TB + threadIdx.x;
    const int plane = blockIdx.z;
    const int y0 = blockIdx.y * TB - HALO, x0 = blockIdx.x * TB - HALO;
    const int64_t HW = (int64_t)H * W;
    float acc[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {  // horizontal sweeps share one staging tile; one vertical sweep for all three
        const float *mp = maps + ((int64_t)plane * 3 + k) * HW;
        __syncthreads();
        for (int e = tid; e < TS * TS; e += TB * TB) {
            int ly = e / TS, lx = e - ly * TS;
            int y = y0 + ly, x = x0 + lx;
            sm[ly][lx] = (y >= 0 && y < H && x >= 0 && x < W) ? mp[(int64_t)y * W + x] : 0.f;
        }
        __syncthreads();
        for (int e = tid; e < TS * TB; e += TB * TB) {
            const int y = e / TB, x = e - y * TB;
            tmp3[k][y][x] = conv_h(sm, y, x);
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float v = 0.f;
#pragma unroll
        for (int t = 0; t < 11; ++t) v += c_gauss[t] * tmp3[k][threadIdx.y + t][threadIdx.x];
        acc[k] = v;
    }
    const int py = blockIdx.y * TB + threadIdx.y, px = blockIdx.x * TB + threadIdx.x;
    if (py >= H || px >= W) return;
    const int64_t off = plane * img_plane_stride + (int64_t)py * img_row_stride + (int64_t)px * img_pix_stride;
    const float p1 = img[off];
    const float p2 = load_gt<GT>(gt + (int64_t)plane * HW + (int64_t)py * W + px);
    float v = acc[0] + p1 * 2.0f * acc[1] + p2 * acc[2];
    const float d = p1 - p2;
    v += l1w * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
    v_img[off] = v;
}

__global__ void loss_finalize_kernel(const double *sums, double n, float lambda, float *out) {
    double ssim = sums[0] / n, l1 = sums[1] / n;
    out[0] = (float)((1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - ssim));
    out[1] = (float)l1;
    out[2] = (float)ssim;
}
__global__ void ssim_finalize_kernel(const double *sums, double n, float *out) { out[0] = (float)(sums[0] / n); }

inline size_t maps_bytes(int64_t planes, int H, int W) { return (size_t)planes * 3 * H * W * sizeof(float); }

}  // namespace

CLMGS_EXPORT size_t clmgs_loss_scratch_bytes(int height, int width) {
    return 256 + maps_bytes(3, height, width);
}

CLMGS_EXPORT int clmgs_loss_fwd_bwd(int height, int width, const float *image, int hwc, const uint8_t *gt_u8,
                                    float lambda_dssim, float *loss_out, float *v_image, void *scratch,
                                    size_t scratch_bytes, void *stream) {
    CLMGS_REQUIRE(height > 0 && width > 0, "loss: empty image");
    if (scratch_bytes < clmgs_loss_scratch_bytes(height, width)) {
        clmgs_set_error("loss: scratch too small (%zu < %zu)", scratch_bytes, clmgs_loss_scratch_bytes(height, width));
        return CLMGS_ERR_SCRATCH;
    }
    cudaStream_t st = as_stream(stream);
    double *sums = (double *)scratch;
    float *maps = (float *)((char *)scratch + 256);
    CLMGS_CUDA(cudaMemsetAsync(sums, 0, 2 * sizeof(double), st));
    const double n = 3.0 * height * width;
    const int64_t plane_stride = hwc ? 1 : (int64_t)height * width;
    const int64_t row_stride = hwc ? (int64_t)width * 3 : width;
    const int pix_stride = hwc ? 3 : 1;
    dim3 grid((width + TB - 1) / TB, (height + TB - 1) / TB, 3), block(TB, TB, 1);
    const float dmap = (float)(-(double)lambda_dssim / n);
    ssim_pass1_kernel<uint8_t><<<grid, block, 0, st>>>(height, width, image, plane_stride, row_stride, pix_stride,
                                                       gt_u8, dmap, maps, sums, v_image != nullptr);
    CLMGS_LAUNCH_OK();
    if (v_image) {
        ssim_pass2_kernel<uint8_t><<<grid, block, 0, st>>>(height, width, image, plane_stride, row_stride,
                                                           pix_stride, gt_u8, maps,
                                                           (float)((1.0 - (double)lambda_dssim) / n), v_image);
        CLMGS_LAUNCH_OK();
    }
    loss_finalize_kernel<<<1, 1, 0, st>>>(sums, n, lambda_dssim, loss_out);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_fused_ssim(int B, int CH, int height, int width, const float *img1, const float *img2,
                                  float *ssim_out, float *v_img1, void *scratch, size_t scratch_bytes,
                                  void *stream) {
    CLMGS_REQUIRE(B > 0 && CH > 0 && height > 0 && width > 0, "fused_ssim: empty input");
    const int64_t planes = (int64_t)B * CH;
    CLMGS_REQUIRE(planes <= 65535, "fused_ssim: too many planes");
    const size_t need = 256 + (v_img1 ? maps_bytes(planes, height, width) : 0);
    if (scratch_bytes < need) {
        clmgs_set_error("fused_ssim: scratch too small (%zu < %zu)", scratch_bytes, need);
        return CLMGS_ERR_SCRATCH;
    }
    cudaStream_t st = as_stream(stream);
    double *sums = (double *)scratch;
    float *maps = (float *)((char *)scratch + 256);
    CLMGS_CUDA(cudaMemsetAsync(sums, 0, 2 * sizeof(double), st));
    const double n = (double)planes * height * width;
    dim3 grid((width + TB - 1) / TB, (height + TB - 1) / TB, (unsigned)planes), block(TB, TB, 1);
    ssim_pass1_kernel<float><<<grid, block, 0, st>>>(height, width, img1, (int64_t)height * width, width, 1, img2,
                                                     (float)(1.0 / n), maps, sums, v_img1 != nullptr);
    CLMGS_LAUNCH_OK();
    if (v_img1) {
        ssim_pass2_kernel<float><<<grid, block, 0, st>>>(height, width, img1, (int64_t)height * width, width, 1,
                                                         img2, maps, 0.f, v_img1);
        CLMGS_LAUNCH_OK();
    }
    ssim_finalize_kernel<<<1, 1, 0, st>>>(sums, n, ssim_out);
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
