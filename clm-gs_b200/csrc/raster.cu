// S5 alpha-composited rasterisation, forward and backward, 3 colour channels,
// 16x16-pixel tiles (one CTA of 256 threads per tile, one pixel per thread).
//
// Semantics follow rasterize_to_pixels_{fwd,bwd}_kernel<3>
// (CSRC/rasterize_to_pixels_fwd.cu:14-160, CSRC/rasterize_to_pixels_bwd.cu:14-249):
// front-to-back compositing, alpha = min(0.999, o*exp(-sigma)), skip when
// sigma < 0 or alpha < 1/255, exclusive termination at T*(1-alpha) <= 1e-4.
//
// Differences in how it is computed (B200-first):
//  * a batch of 256 Gaussians is staged in shared memory as two float4 + one
//    float per Gaussian INCLUDING its colour, so the inner loop touches no
//    global memory (the reference fetches colours from global per hit);
//  * while a batch is staged, every Gaussian is tested against the tile exactly (minimum of its quadratic
//    form over the tile's pixel-centre rectangle vs ln(255 * opacity)): the reference's tile lists come from
//    the square of the 3-sigma radius, so for anisotropic or faint Gaussians most (tile, Gaussian) pairs can
//    never reach alpha >= 1/255 at any pixel of the tile.  Those are dropped during staging (stable
//    ballot compaction in shared memory) and the survivors carry the image rows they can touch, which lets
//    each warp (2 rows) skip the rest with one compare.  The test is conservative (tolerances cover the
//    rounding of the per-pixel evaluation), so images, last_ids and gradients are unchanged;
//  * backward: branch-free inner loop, the 9 per-Gaussian partial gradients of a warp are reduced with
//    a multi-value butterfly (14 shuffles instead of 45) and the 8 finished
//    sums are added by 8 different lanes in parallel.
#include "common.cuh"
#include <math_constants.h>
#include <stdlib.h>

constexpr int kTile = 16;
constexpr int kBlock = kTile * kTile;
// resident CTAs per SM the forward kernel is compiled for: 5 = 48 registers (4 would give 54 registers and a
// slightly shorter loop; tools/raster_variants.sh builds and times both)
#ifndef CLMGS_RASTER_FWD_MIN_BLOCKS
#define CLMGS_RASTER_FWD_MIN_BLOCKS 5
#endif

struct RasterArgs {
    int C;
    int64_t N;
    int64_t n_isects;
    int width, height, tile_width, tile_height;
};

__device__ __forceinline__ void tile_range(const RasterArgs &a, const int32_t *tile_offsets, int cam, int tile_id,
                                           int &start, int &end) {
    int64_t t = (int64_t)cam * a.tile_width * a.tile_height + tile_id;
    start = tile_offsets[t];
    end = (t == (int64_t)a.C * a.tile_width * a.tile_height - 1) ? (int)a.n_isects : tile_offsets[t + 1];
}


// Can a Gaussian (mean, conic a/b/c, opacity) reach alpha = min(0.999, o exp(-sigma)) >= 1/255 at ANY pixel centre
// of the rectangle [xmin,xmax] x [ymin,ymax]?  sigma(d) = 0.5 (a dx^2 + c dy^2) + b dx dy must get below
// tau = ln(255 o).  Conservative: slack covers the rounding of the per-pixel evaluation, degenerate conics are kept.
// ylo / yhi bound the rows where sigma <= tau is possible at all (|dy| <= sqrt(2 tau a / det)).
__device__ __forceinline__ bool tile_touch(float mx, float my, float ca, float cb, float cc, float opac, float xmin,
                                           float xmax, float ymin, float ymax, float &ylo, float &yhi) {
    ylo = -CUDART_INF_F; yhi = CUDART_INF_F;
    const float o255 = opac * 255.0f;
    if (o255 < 0.999f) return false;  // alpha <= opacity < 1/255 everywhere (NaN opacities fall through)
    // det = a c - b^2 cancels badly for elongated Gaussians: compensated product difference (error ~1 ulp of det)
    const float bb = cb * cb;
    const float det = fmaf(ca, cc, -bb) - fmaf(cb, cb, -bb);
    if (!(ca > 0.f && cc > 0.f && det > 0.f) || !(o255 < CUDART_INF_F)) return true;
    const float tau = __logf(o255) * 1.001f + 0.05f;
    const float ey = sqrtf(2.f * tau * ca / det) * 1.001f + 0.01f;
    const float ex = sqrtf(2.f * tau * cc / det) * 1.001f + 0.01f;
    if (!(ey < CUDART_INF_F) || !(ex < CUDART_INF_F)) return true;
    ylo = my - ey; yhi = my + ey;
    if (mx + ex < xmin || mx - ex > xmax || yhi < ymin || ylo > ymax) return false;
    const float dx0 = xmin - mx, dx1 = xmax - mx, dy0 = ymin - my, dy1 = ymax - my;
    if (dx0 <= 0.f && dx1 >= 0.f && dy0 <= 0.f && dy1 >= 0.f) return true;  // the mean is inside
    float best = CUDART_INF_F;  // min over the 4 edges of (sigma - rounding allowance)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        const float X = e ? dx1 : dx0;
        const float y = fminf(fmaxf(-cb * X / cc, dy0), dy1);
        const float q = 0.5f * (ca * X * X + cc * y * y), m = cb * X * y;
        best = fminf(best, (q + m) - 1e-5f * (q + fabsf(m)));
        const float Y = e ? dy1 : dy0;
        const float x = fminf(fmaxf(-cb * Y / ca, dx0), dx1);
        const float q2 = 0.5f * (ca * x * x + cc * Y * Y), m2 = cb * x * Y;
        best = fminf(best, (q2 + m2) - 1e-5f * (q2 + fabsf(m2)));
    }
    return !(best > tau);
}

// Stable compaction of the staged batch: thread tr holds one candidate; survivors are written to consecutive
// slots of the shared arrays in list order.  Returns the number of survivors (same value in every thread).
__device__ __forceinline__ int stage_compact(bool keep, int tr, int *s_wcnt, int &slot) {
    const int lane = tr & 31, warp = tr >> 5;
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kBlock / 32; ++w) {
        const int c = s_wcnt[w];
        if (w < warp) before += c;
        total += c;
    }
    slot = before + __popc(bal & ((1u << lane) - 1u));
    return total;
}

// COUNT: also accumulate the work counters the bench's FLOP roofline needs into counters[0..2]:
//   [0] P_f = sum over pixels of the list entries the REFERENCE kernel examines (tile start through the entry
//             that terminates the pixel, or the whole list) -- the algorithmic figure, independent of the pruning
//   [1] entries this kernel actually evaluated per pixel (after tile-touch pruning and row skipping)
//   [2] P_b = sum over pixels of (last blended entry - tile start + 1): what the reference backward walks
template <bool HAS_BG, bool COUNT>
__global__ void __launch_bounds__(kBlock, CLMGS_RASTER_FWD_MIN_BLOCKS)
raster_fwd_kernel(RasterArgs a, const float2 *__restrict__ means2d, const float *__restrict__ conics,
                  const float *__restrict__ colors, const float *__restrict__ opacities,
                  const float *__restrict__ backgrounds, const int32_t *__restrict__ tile_offsets,
                  const int32_t *__restrict__ flatten_ids, float *__restrict__ render_colors,
                  float *__restrict__ render_alphas, int32_t *__restrict__ last_ids,
                  unsigned long long *__restrict__ counters) {
    // one staged Gaussian = 3 consecutive float4 (one pointer walks the batch):
    //   [0] x, y, conic.x, conic.y   [1] conic.z, opacity, r, g   [2] b, first row, last row (pixel-centre y), position in the batch
    __shared__ float4 s_g[3 * kBlock];
    __shared__ int s_wcnt[kBlock / 32];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * a.tile_width + blockIdx.x;
    const int i = blockIdx.y * kTile + threadIdx.y;
    const int j = blockIdx.x * kTile + threadIdx.x;
    const int tr = threadIdx.y * kTile + threadIdx.x;
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const bool inside = i < a.height && j < a.width;
    bool done = !inside;
    int start, end;
    tile_range(a, tile_offsets, cam, tile_id, start, end);
    const int num_batches = (end - start + kBlock - 1) / kBlock;
    // pixel centres of this tile / of this warp's two rows
    const float txmin = (float)(blockIdx.x * kTile) + 0.5f, txmax = (float)min(blockIdx.x * kTile + kTile - 1, a.width - 1) + 0.5f;
    const float tymin = (float)(blockIdx.y * kTile) + 0.5f, tymax = (float)min(blockIdx.y * kTile + kTile - 1, a.height - 1) + 0.5f;
    const float wy0 = (float)(blockIdx.y * kTile + 2 * (tr >> 5)) + 0.5f, wy1 = wy0 + 1.0f;

    float T = 1.0f, r = 0.f, g = 0.f, b = 0.f;
    int cur_idx = 0;
    int n_examined = inside ? end - start : 0, n_eval = 0;  // COUNT only
    bool blended = false;
    for (int bt = 0; bt < num_batches; ++bt) {
        if (__syncthreads_count(done) >= kBlock) break;
        const int batch_start = start + kBlock * bt;
        const int idx = batch_start + tr;
        bool keep = false;
        float4 A, Bv, Cv;
        if (idx < end) {
            const int gid = flatten_ids[idx];
            const float2 xy = means2d[gid];
            const float *cn = conics + 3 * (int64_t)gid;
            const float *cl = colors + 3 * (int64_t)gid;
            A = make_float4(xy.x, xy.y, cn[0], cn[1]);
            Bv = make_float4(cn[2], opacities[gid], cl[0], cl[1]);
            float ylo, yhi;
            keep = tile_touch(A.x, A.y, A.z, A.w, Bv.x, Bv.y, txmin, txmax, tymin, tymax, ylo, yhi);
            Cv = make_float4(cl[2], ylo, yhi, __int_as_float(tr));
        }
        int slot;
        const int batch_size = stage_compact(keep, tr, s_wcnt, slot);
        if (keep) { float4 *d = s_g + 3 * slot; d[0] = A; d[1] = Bv; d[2] = Cv; }
        __syncthreads();
        if (done) continue;
        const float4 *pg = s_g;
        for (int t = batch_size; t > 0; --t, pg += 3) {
            const float4 Cc = pg[2];
            if (wy1 < Cc.y || wy0 > Cc.z) continue;  // cannot reach 1/255 on this warp's rows (warp-uniform)
            const float4 A = pg[0];
            const float4 Bv = pg[1];
            const float dx = A.x - px, dy = A.y - py;
            const float sigma = 0.5f * (A.z * dx * dx + Bv.x * dy * dy) + A.w * dx * dy;
            const float alpha = fminf(0.999f, Bv.y * __expf(-sigma));
            if (COUNT) ++n_eval;
            if (sigma < 0.f || alpha < 1.f / 255.f) continue;
            const float next_T = T * (1.0f - alpha);
            if (next_T <= 1e-4f) {
                done = true;
                if (COUNT) n_examined = batch_start + __float_as_int(Cc.w) - start + 1;
                break;
            }
            const float vis = alpha * T;
            r += Bv.z * vis; g += Bv.w * vis; b += Cc.x * vis;
            cur_idx = batch_start + __float_as_int(Cc.w);
            if (COUNT) blended = true;
            T = next_T;
        }
    }
    if (inside) {
        const int64_t pix = ((int64_t)cam * a.height + i) * a.width + j;
        render_alphas[pix] = 1.0f - T;
        if (HAS_BG) {
            const float *bg = backgrounds + 3 * cam;
            r += T * bg[0]; g += T * bg[1]; b += T * bg[2];
        }
        render_colors[3 * pix] = r; render_colors[3 * pix + 1] = g; render_colors[3 * pix + 2] = b;
        last_ids[pix] = cur_idx;
    }
    if (COUNT) {
        int n_back = blended ? cur_idx - start + 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            n_examined += __shfl_xor_sync(0xffffffffu, n_examined, o);
            n_eval += __shfl_xor_sync(0xffffffffu, n_eval, o);
            n_back += __shfl_xor_sync(0xffffffffu, n_back, o);
        }
        // 64 partial sums per counter, picked by tile: thousands of warps adding to three words would serialise
        if ((tr & 31) == 0) {
            unsigned long long *c = counters + 3 * ((tile_id + (tr >> 5)) & 63);
            atomicAdd(c + 0, (unsigned long long)n_examined);
            atomicAdd(c + 1, (unsigned long long)n_eval);
            atomicAdd(c + 2, (unsigned long long)n_back);
        }
    }
}

static int rasterize_fwd_impl(int C, int64_t N, int64_t n_isects, const float *means2d, const float *conics,
                              const float *colors, const float *opacities, const float *backgrounds,
                              int width, int height, int tile_size, int tile_width, int tile_height,
                              const int32_t *tile_offsets, const int32_t *flatten_ids, float *render_colors,
                              float *render_alphas, int32_t *last_ids, unsigned long long *counters, void *stream) {
    CLMGS_REQUIRE(tile_size == kTile, "rasterize: tile_size must be 16 (got %d)", tile_size);
    CLMGS_REQUIRE(tile_width * kTile >= width && tile_height * kTile >= height,
                  "rasterize: tile grid %dx%d does not cover %dx%d", tile_width, tile_height, width, height);
    CLMGS_REQUIRE(n_isects < ((int64_t)1 << 31), "rasterize: n_isects too large");
    if (C == 0 || width == 0 || height == 0) return CLMGS_OK;
    RasterArgs a{C, N, n_isects, width, height, tile_width, tile_height};
    dim3 grid(tile_width, tile_height, C), block(kTile, kTile, 1);
    cudaStream_t st = as_stream(stream);
#define LAUNCH_FWD(BG, CNT)                                                                                      \
    raster_fwd_kernel<BG, CNT><<<grid, block, 0, st>>>(a, (const float2 *)means2d, conics, colors, opacities,      \
                                                       backgrounds, tile_offsets, flatten_ids, render_colors,      \
                                                       render_alphas, last_ids, counters)
    if (counters) { if (backgrounds) LAUNCH_FWD(true, true); else LAUNCH_FWD(false, true); }
    else { if (backgrounds) LAUNCH_FWD(true, false); else LAUNCH_FWD(false, false); }
#undef LAUNCH_FWD
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_rasterize_fwd(int C, int64_t N, int64_t n_isects, const float *means2d, const float *conics,
                                     const float *colors, const float *opacities, const float *backgrounds,
                                     int width, int height, int tile_size, int tile_width, int tile_height,
                                     const int32_t *tile_offsets, const int32_t *flatten_ids, float *render_colors,
                                     float *render_alphas, int32_t *last_ids, void *stream) {
    return rasterize_fwd_impl(C, N, n_isects, means2d, conics, colors, opacities, backgrounds, width, height, tile_size,
                              tile_width, tile_height, tile_offsets, flatten_ids, render_colors, render_alphas, last_ids,
                              nullptr, stream);
}

// Same outputs, plus work_counters[64][3] (device, uint64, ACCUMULATED into: the caller zeroes them; the three totals
// are the column sums) -- see the COUNT note on the kernel.  Measurement only.
CLMGS_EXPORT int clmgs_rasterize_fwd_counted(int C, int64_t N, int64_t n_isects, const float *means2d,
                                             const float *conics, const float *colors, const float *opacities,
                                             const float *backgrounds, int width, int height, int tile_size,
                                             int tile_width, int tile_height, const int32_t *tile_offsets,
                                             const int32_t *flatten_ids, float *render_colors, float *render_alphas,
                                             int32_t *last_ids, int64_t *work_counters, void *stream) {
    CLMGS_REQUIRE(work_counters != nullptr, "rasterize_fwd_counted: work_counters is null");
    return rasterize_fwd_impl(C, N, n_isects, means2d, conics, colors, opacities, backgrounds, width, height, tile_size,
                              tile_width, tile_height, tile_offsets, flatten_ids, render_colors, render_alphas, last_ids,
                              (unsigned long long *)work_counters, stream);
}

// ---------------------------------------------------------------------------
// Backward
// ---------------------------------------------------------------------------
// Butterfly reduction of 8 values across a warp in 9 shuffles.  On return,
// lane L with (L & 3) == 0 holds the full warp sum of value index
// ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1) in v[0].
__device__ __forceinline__ float butterfly8(float v[8], int lane) {
    const bool up16 = lane & 16;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float send = up16 ? v[k] : v[k + 4];
        float keep = up16 ? v[k + 4] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    const bool up8 = lane & 8;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        float send = up8 ? v[k] : v[k + 2];
        float keep = up8 ? v[k + 2] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    const bool up4 = lane & 4;
    {
        float send = up4 ? v[0] : v[1];
        float keep = up4 ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    return v[0];
}

// bin_final is 0 both for "entry 0 was blended" and for "nothing blended"; either way entries
// above it are dead, so the value itself is the bound (clamped below the range it means "none").
__device__ __forceinline__ int inside_any(int warp_bin_final, int start) {
    return warp_bin_final >= start ? warp_bin_final : start - 1;
}

// TILE_ACC: the per-Gaussian sums of a tile's 8 warps meet in shared memory (shared atomics, one slot of 9-11 floats
// per staged Gaussian) and leave as ONE set of global atomics per (tile, Gaussian) after the batch, instead of one
// set per (warp, Gaussian) as in the reference (rasterize_to_pixels_bwd.cu:215-246).
template <bool HAS_BG, bool ABSGRAD, bool TILE_ACC>
__global__ void __launch_bounds__(kBlock, 4)
raster_bwd_kernel(RasterArgs a, const float2 *__restrict__ means2d, const float *__restrict__ conics,
                  const float *__restrict__ colors, const float *__restrict__ opacities,
                  const float *__restrict__ backgrounds, const int32_t *__restrict__ tile_offsets,
                  const int32_t *__restrict__ flatten_ids, const float *__restrict__ render_alphas,
                  const int32_t *__restrict__ last_ids, const float *__restrict__ v_render_colors,
                  const float *__restrict__ v_render_alphas, float *__restrict__ v_means2d_abs,
                  float *__restrict__ v_means2d, float *__restrict__ v_conics, float *__restrict__ v_colors,
                  float *__restrict__ v_opacities) {
    // one staged Gaussian = 4 consecutive float4 (one pointer walks the batch):
    //   [0] x, y, conic.x, conic.y   [1] conic.z, opacity, r, g   [2] b, first row, last row (pixel-centre y), position in the batch
    //   [3] .x = Gaussian id (int bits)
    __shared__ float4 s_g[4 * kBlock];
    __shared__ int s_wcnt[kBlock / 32];
    constexpr int kAcc = ABSGRAD ? 11 : 9;  // rgb, conic xyz, mean xy, opacity (, |mean| xy)
    __shared__ float s_acc[TILE_ACC ? kBlock * kAcc : 1];
    __shared__ int s_hit[TILE_ACC ? kBlock : 1];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * a.tile_width + blockIdx.x;
    const int i = blockIdx.y * kTile + threadIdx.y;
    const int j = blockIdx.x * kTile + threadIdx.x;
    const int tr = threadIdx.y * kTile + threadIdx.x;
    const int lane = tr & 31;
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const bool inside = i < a.height && j < a.width;
    int start, end;
    tile_range(a, tile_offsets, cam, tile_id, start, end);
    const int num_batches = (end - start + kBlock - 1) / kBlock;

    const int64_t pix = inside ? ((int64_t)cam * a.height + i) * a.width + j : 0;
    const float T_final = inside ? 1.0f - render_alphas[pix] : 1.0f;
    float T = T_final;
    float buf_v = 0.f;  // sum over the entries behind the current one of (colour . v_render) * alpha * T
    const int bin_final = inside ? last_ids[pix] : 0;
    float vr = 0.f, vg = 0.f, vb = 0.f, va = 0.f;
    if (inside) {
        vr = v_render_colors[3 * pix]; vg = v_render_colors[3 * pix + 1]; vb = v_render_colors[3 * pix + 2];
        if (v_render_alphas) va = v_render_alphas[pix];
    }
    float bg_dot = 0.f;
    if (HAS_BG) {
        const float *bg = backgrounds + 3 * cam;
        bg_dot = bg[0] * vr + bg[1] * vg + bg[2] * vb;
    }
    const float tf_va = T_final * (va - bg_dot);
    int warp_bin_final = bin_final;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_bin_final = max(warp_bin_final, __shfl_xor_sync(0xffffffffu, warp_bin_final, o));
    // Entries behind the last one any pixel of this tile blended contribute nothing: with early
    // termination that is most of a long tile list, so do not even stage those batches (the
    // reference loads every batch).  tile_last = max over the CTA of bin_final.
    __shared__ int s_tile_last;
    if (tr == 0) s_tile_last = -1;
    __syncthreads();
    if (lane == 0) atomicMax(&s_tile_last, inside_any(warp_bin_final, start));
    __syncthreads();
    const int tile_last = s_tile_last;
    if (tile_last < start) return;  // nothing was blended anywhere in this tile
    const int first_batch = (end - 1 - tile_last) / kBlock;
    const float txmin = (float)(blockIdx.x * kTile) + 0.5f, txmax = (float)min(blockIdx.x * kTile + kTile - 1, a.width - 1) + 0.5f;
    const float tymin = (float)(blockIdx.y * kTile) + 0.5f, tymax = (float)min(blockIdx.y * kTile + kTile - 1, a.height - 1) + 0.5f;
    const float wy0 = (float)(blockIdx.y * kTile + 2 * (tr >> 5)) + 0.5f, wy1 = wy0 + 1.0f;

    // Where this lane's warp sum goes (per-warp atomics, !TILE_ACC): after the butterfly, lanes 0,4,..,28 hold the
    // eight sums (rgb, conic xyz, mean xy), lane 1 the opacity sum (, lanes 2 / 3 the |mean| sums); element of
    // Gaussian g = lane_dst[g * lane_stride].  Fixed per lane, so the inner loop needs no address selection.
    float *lane_dst = nullptr;
    unsigned lane_stride = 0;
    if (!TILE_ACC) {
        const int k = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
        if ((lane & 3) == 0) {
            if (k < 3) { lane_dst = v_colors + k; lane_stride = 3; }
            else if (k < 6) { lane_dst = v_conics + (k - 3); lane_stride = 3; }
            else { lane_dst = v_means2d + (k - 6); lane_stride = 2; }
        } else if (lane == 1) {
            lane_dst = v_opacities; lane_stride = 1;
        } else if (ABSGRAD && (lane == 2 || lane == 3)) {
            lane_dst = v_means2d_abs + (lane - 2); lane_stride = 2;
        }
    }

    for (int bt = first_batch; bt < num_batches; ++bt) {
        __syncthreads();
        const int batch_end = end - 1 - kBlock * bt;  // absolute index of the back-most entry of this batch
        const int idx = batch_end - tr;
        bool keep = false;
        float4 sA, sB, sC;
        int gid_s = 0;
        if (idx >= start) {
            gid_s = flatten_ids[idx];
            const float2 xy = means2d[gid_s];
            const float *cn = conics + 3 * (int64_t)gid_s;
            const float *cl = colors + 3 * (int64_t)gid_s;
            sA = make_float4(xy.x, xy.y, cn[0], cn[1]);
            sB = make_float4(cn[2], opacities[gid_s], cl[0], cl[1]);
            float ylo, yhi;
            keep = tile_touch(sA.x, sA.y, sA.z, sA.w, sB.x, sB.y, txmin, txmax, tymin, tymax, ylo, yhi);
            sC = make_float4(cl[2], ylo, yhi, __int_as_float(tr));
        }
        int slot;
        const int batch_size = stage_compact(keep, tr, s_wcnt, slot);
        if (keep) { float4 *d = s_g + 4 * slot; d[0] = sA; d[1] = sB; d[2] = sC; d[3].x = __int_as_float(gid_s); }
        if (TILE_ACC) {
#pragma unroll
            for (int q = 0; q < kAcc; ++q) s_acc[q * kBlock + tr] = 0.f;
            s_hit[tr] = 0;
        }
        __syncthreads();
        const int t_lo = batch_end - warp_bin_final;  // entries with a smaller position are behind every pixel's last
        const float4 *pg = s_g;
        for (int tc = 0; tc < batch_size; ++tc, pg += 4) {
            const float4 Cc = pg[2];
            const int t = __float_as_int(Cc.w);  // position in the (uncompacted) batch
            if (t < t_lo || wy1 < Cc.y || wy0 > Cc.z) continue;  // warp-uniform
            // Branch-free from here: every lane evaluates the Gaussian; a lane that does not blend it (outside the
            // image, behind its last entry, sigma < 0, alpha < 1/255) carries alpha = vis = 0, which makes each of its
            // terms below exactly zero and leaves its running state untouched.
            const float4 A = pg[0], Bv = pg[1];
            const float opac = Bv.y;
            const float dx = A.x - px, dy = A.y - py;
            const float sigma = 0.5f * (A.z * dx * dx + Bv.x * dy * dy) + A.w * dx * dy;
            float vis = __expf(-sigma);
            float alpha = fminf(0.999f, opac * vis);
            const bool valid = inside && (batch_end - t <= bin_final) && !(sigma < 0.f || alpha < 1.f / 255.f);
            if (!__any_sync(0xffffffffu, valid)) continue;
            if (!valid) { alpha = 0.f; vis = 0.f; }
            const float ra = 1.0f / (1.0f - alpha);
            if (valid) T *= ra;
            const float fac = alpha * T;
            // sum_c (c_c T - buf_c ra) v_c + T_final ra (v_a - bg . v)  with  buf . v  kept as ONE running scalar
            const float cv = Bv.z * vr + Bv.w * vg + Cc.x * vb;
            const float v_alpha = T * cv + ra * (tf_va - buf_v);
            buf_v += cv * fac;
            const float v_opa_px = (opac * vis <= 0.999f) ? vis * v_alpha : 0.f;  // alpha clamped: no gradient through it
            const float v_sigma = -opac * v_opa_px;
            float v[8];  // rgb, then conic xyz + mean xy (TILE_ACC: their raw moments, combined once per tile at the flush)
            v[0] = fac * vr; v[1] = fac * vg; v[2] = fac * vb;
            const float sx = v_sigma * dx, sy = v_sigma * dy;
            float v_ax = 0.f, v_ay = 0.f;
            if (TILE_ACC) {
                v[3] = sx * dx; v[4] = sx * dy; v[5] = sy * dy; v[6] = sx; v[7] = sy;
                if (ABSGRAD) { v_ax = fabsf(A.z * sx + A.w * sy); v_ay = fabsf(A.w * sx + Bv.x * sy); }
            } else {
                v[3] = 0.5f * sx * dx; v[4] = sx * dy; v[5] = 0.5f * sy * dy;
                v[6] = A.z * sx + A.w * sy; v[7] = A.w * sx + Bv.x * sy;
                if (ABSGRAD) { v_ax = fabsf(v[6]); v_ay = fabsf(v[7]); }
            }
            const float red = butterfly8(v, lane);
            const float v_opa = warp_sum(v_opa_px);
            if (ABSGRAD) { v_ax = warp_sum(v_ax); v_ay = warp_sum(v_ay); }
            if (TILE_ACC) {
                // value q of staged Gaussian tc lives at s_acc[q * kBlock + tc].  Lanes 0,4,..,28 hold the eight
                // butterfly sums, lane 1 the opacity sum (, lanes 2 / 3 the |mean| sums): ONE shared atomic for all of
                // them (a float atomic on shared memory is a compare-and-swap loop; one loop, not one per lane group)
                int q = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
                float val = red;
                bool part = (lane & 3) == 0;
                if (lane == 1) { q = 8; val = v_opa; part = true; }
                if (ABSGRAD && lane == 2) { q = 9; val = v_ax; part = true; }
                if (ABSGRAD && lane == 3) { q = 10; val = v_ay; part = true; }
                if (part) atomicAdd(&s_acc[q * kBlock + tc], val);
                if (lane == 1) s_hit[tc] = 1;
                continue;
            }
            // one reduction instruction for the whole warp: every participating lane adds its sum to its own array
            float val = red;
            if (lane == 1) val = v_opa;
            if (ABSGRAD && lane == 2) val = v_ax;
            if (ABSGRAD && lane == 3) val = v_ay;
            if (lane_dst) atomicAdd(lane_dst + (unsigned)__float_as_int(pg[3].x) * lane_stride, val);  // 32-bit index: N < 2^32 / 3 (checked)
        }
        if (TILE_ACC) {
            __syncthreads();  // every warp's contribution to this batch is in s_acc
            if (tr < batch_size && s_hit[tr]) {
                const int64_t gid = __float_as_int(s_g[4 * tr + 3].x);
                atomicAdd(v_colors + 3 * gid, s_acc[0 * kBlock + tr]);
                atomicAdd(v_colors + 3 * gid + 1, s_acc[1 * kBlock + tr]);
                atomicAdd(v_colors + 3 * gid + 2, s_acc[2 * kBlock + tr]);
                // moments of v_sigma over the tile -> conic and mean gradients (conic a, b, c of staged Gaussian tr)
                const float ca = s_g[4 * tr].z, cb = s_g[4 * tr].w, cc = s_g[4 * tr + 1].x;
                const float m_x = s_acc[6 * kBlock + tr], m_y = s_acc[7 * kBlock + tr];
                atomicAdd(v_conics + 3 * gid, 0.5f * s_acc[3 * kBlock + tr]);
                atomicAdd(v_conics + 3 * gid + 1, s_acc[4 * kBlock + tr]);
                atomicAdd(v_conics + 3 * gid + 2, 0.5f * s_acc[5 * kBlock + tr]);
                atomicAdd(v_means2d + 2 * gid, ca * m_x + cb * m_y);
                atomicAdd(v_means2d + 2 * gid + 1, cb * m_x + cc * m_y);
                atomicAdd(v_opacities + gid, s_acc[8 * kBlock + tr]);
                if (ABSGRAD) {
                    atomicAdd(v_means2d_abs + 2 * gid, s_acc[9 * kBlock + tr]);
                    atomicAdd(v_means2d_abs + 2 * gid + 1, s_acc[10 * kBlock + tr]);
                }
            }
        }
    }
}

CLMGS_EXPORT int clmgs_rasterize_bwd(int C, int64_t N, int64_t n_isects, const float *means2d, const float *conics,
                                     const float *colors, const float *opacities, const float *backgrounds,
                                     int width, int height, int tile_size, int tile_width, int tile_height,
                                     const int32_t *tile_offsets, const int32_t *flatten_ids,
                                     const float *render_alphas, const int32_t *last_ids,
                                     const float *v_render_colors, const float *v_render_alphas,
                                     float *v_means2d_abs, float *v_means2d, float *v_conics, float *v_colors,
                                     float *v_opacities, void *stream) {
    CLMGS_REQUIRE(tile_size == kTile, "rasterize: tile_size must be 16 (got %d)", tile_size);
    CLMGS_REQUIRE(tile_width * kTile >= width && tile_height * kTile >= height,
                  "rasterize: tile grid %dx%d does not cover %dx%d", tile_width, tile_height, width, height);
    CLMGS_REQUIRE((int64_t)C * N < ((int64_t)1 << 32) / 3, "rasterize_bwd: C * N too large for 32-bit element indices");
    if (C == 0 || width == 0 || height == 0 || n_isects == 0) return CLMGS_OK;
    RasterArgs a{C, N, n_isects, width, height, tile_width, tile_height};
    dim3 grid(tile_width, tile_height, C), block(kTile, kTile, 1);
    cudaStream_t st = as_stream(stream);
    // CLMGS_RASTER_BWD_TILE_ACC=1 selects the per-tile accumulation in shared memory (kept for A/B measurements:
    // on B200 it is 10 % slower than the per-warp global reductions)
    static const bool tile_acc = [] { const char *e = getenv("CLMGS_RASTER_BWD_TILE_ACC"); return e && e[0] == '1'; }();
#define LAUNCH(BG, AB, TA)                                                                                          \
    raster_bwd_kernel<BG, AB, TA><<<grid, block, 0, st>>>(a, (const float2 *)means2d, conics, colors, opacities,     \
                                                          backgrounds, tile_offsets, flatten_ids, render_alphas,     \
                                                          last_ids, v_render_colors, v_render_alphas, v_means2d_abs, \
                                                          v_means2d, v_conics, v_colors, v_opacities)
#define LAUNCH2(BG, AB) do { if (tile_acc) LAUNCH(BG, AB, true); else LAUNCH(BG, AB, false); } while (0)
    if (backgrounds) { if (v_means2d_abs) LAUNCH2(true, true); else LAUNCH2(true, false); }
    else { if (v_means2d_abs) LAUNCH2(false, true); else LAUNCH2(false, false); }
#undef LAUNCH2
#undef LAUNCH
    CLMGS_LAUNCH_OK();
    return CLMGS_OK;
}
