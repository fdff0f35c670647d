/*
 * clmgs.h -- C ABI of libclmgs.so, the B200 (sm_100a) implementation of the
 * CLM-GS per-camera-batch training hot path.
 *
 * This is the drop-in boundary: one entry point per native function the
 * reference binds for this path through pybind11 (gsplat.csrc:
 * submodules/gsplat/gsplat/cuda/csrc/ext.cpp:4-32; clm_kernels._C:
 * submodules/clm_kernels/ext.cpp:7-23), plus the cpu-adam / fast-tsp calls of
 * strategies/clm_offload/engine.py.  Paths below are relative to the reference
 * root; CSRC = submodules/gsplat/gsplat/cuda/csrc, CK = submodules/clm_kernels.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross the boundary
 *   - `d_` / unprefixed pointers are device memory, `h_` pointers are PINNED
 *     host memory mapped into the device address space (UVA zero-copy), `host_`
 *     pointers are ordinary host memory read by the CPU
 *   - outputs are caller-allocated; functions that need temporary global memory
 *     take (scratch, scratch_bytes) and have a *_scratch_bytes() query
 *   - `stream` is a cudaStream_t passed as void*; every launch is asynchronous
 *     on that stream and no entry point synchronises unless it says so
 *   - return value: 0 = ok, negative = error; clmgs_last_error() gives the
 *     message (thread-local).  No C++ exception crosses the boundary.
 *   - all floating point is fp32; ids are int64 where the reference uses int64
 *     (filters, camera_ids/gaussian_ids) and int32 where it uses int32.
 */
#ifndef CLMGS_H
#define CLMGS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CLMGS_OK 0
#define CLMGS_ERR_INVALID (-1)
#define CLMGS_ERR_CUDA (-2)
#define CLMGS_ERR_SCRATCH (-3)

const char *clmgs_last_error(void);
/* ABI version of this header; bumped on any signature change. */
int clmgs_abi_version(void);

/* ---------------------------------------------------------------------------
 * S1  Frustum cull + compaction.
 * Replaces fully_fused_projection_packed_fwd_tensor
 * (CSRC/fully_fused_projection_packed_fwd.cu:184-274, two projection passes +
 * cumsum + .item()) as used by calculate_filters (strategies/base_engine.py:18-76).
 *
 * clmgs_cull_count: every Gaussian is projected ONCE for all B cameras
 * (B <= 64); writes the visibility word of each Gaussian (bit c = camera c;
 * word_bytes = 1,2,4,8 chosen by the caller with 8*word_bytes >= B), the
 * per-camera totals counts[B] (int64) and the per-block partial counts in
 * scratch.  If `activated` is 0 the inputs are the RAW model tensors
 * (log-scale, unnormalised quaternion) and exp()/normalize() are fused in
 * (strategies/base_gaussian_model.py:42-50); if 1 they are already activated,
 * as the reference's python passes them.
 *
 * clmgs_cull_emit: from bitmap + scratch writes, for camera c, its ascending
 * Gaussian ids at gaussian_ids[offset_c ...] with offset_c = sum(counts[<c]),
 * i.e. exactly the flat (camera_ids, gaussian_ids) of the reference.  Optional
 * outputs (may be NULL) reproduce the rest of the packed tuple: camera_ids,
 * radii, means2d, depths, conics.
 * ------------------------------------------------------------------------- */
size_t clmgs_cull_scratch_bytes(int64_t N, int B);
int clmgs_cull_count(const float *means, const float *scales, const float *quats, int activated,
                     int64_t N, const float *viewmats, const float *Ks, int B, int width,
                     int height, float eps2d, float near_plane, float far_plane,
                     float radius_clip, void *bitmap, int word_bytes, int64_t *counts,
                     void *scratch, size_t scratch_bytes, void *stream);
int clmgs_cull_emit(const float *means, const float *scales, const float *quats, int activated,
                    int64_t N, const float *viewmats, const float *Ks, int B, int width,
                    int height, float eps2d, float near_plane, float far_plane,
                    float radius_clip, const void *bitmap, int word_bytes, const void *scratch,
                    int64_t *gaussian_ids, int64_t *camera_ids, int32_t *radii, float *means2d,
                    float *depths, float *conics, void *stream);

/* ---------------------------------------------------------------------------
 * A4  Filter gather + activations (strategies/clm_offload/engine.py:650-666,
 * strategies/base_gaussian_model.py:42-50) and its backward + scatter-add into
 * the dense gradient tensors (engine.py:725-742).
 * ------------------------------------------------------------------------- */
int clmgs_gather_activate(const float *xyz, const float *log_scale, const float *raw_quat,
                          const float *opacity_logit, const int64_t *ids, int64_t n,
                          float *means_out, float *scales_out, float *quats_out,
                          float *opacities_out, void *stream);
/* v_* are gradients w.r.t. the ACTIVATED filtered values; g_* are the dense
 * [N,*] accumulators (+=).  Re-reads the raw parameters to differentiate
 * exp / normalize / sigmoid. */
int clmgs_activate_bwd_scatter(const float *log_scale, const float *raw_quat,
                               const float *opacity_logit, const int64_t *ids, int64_t n,
                               const float *v_means, const float *v_scales, const float *v_quats,
                               const float *v_opacities, float *g_xyz, float *g_log_scale,
                               float *g_raw_quat, float *g_opacity_logit, void *stream);

/* A17  Densification statistics of one camera ("next" row of SURVEY.md 8f;
 * update_densification_stats_offload_accum_grads, densification.py:59-102 ->
 * gsplat_add_densification_stats_exact_filter, strategies/clm_offload/gaussian_model.py:833-851):
 * max_radii2D[id] = max(., radii), xyz_gradient_accum[id] += |v_means2d * (W/2, H/2)|, denom[id] += 1
 * for every filtered Gaussian (ids unique). v_means2d [n,2], radii [n] i32; the three [N] float
 * tables are updated in place. */
int clmgs_densification_stats(const int64_t *ids, int64_t n, const float *v_means2d,
                              const int32_t *radii, int width, int height, float *max_radii2d,
                              float *xyz_gradient_accum, float *denom, void *stream);

/* ---------------------------------------------------------------------------
 * S3  Dense projection.  fully_fused_projection_fwd_tensor / _bwd_tensor
 * (CSRC/fully_fused_projection_fwd.cu:132-181, CSRC/fully_fused_projection_bwd.cu:188-283).
 * Shapes: means [N,3], quats [N,4] wxyz, scales [N,3], viewmats [C,4,4]
 * row-major world->camera, Ks [C,3,3]; outputs radii [C,N] i32, means2d
 * [C,N,2], depths [C,N], conics [C,N,3], compensations [C,N] or NULL.
 * Unlike the reference (torch::empty) culled rows are written as zeros.
 * bwd accumulates (+=) into v_means/v_quats/v_scales, which the caller zeroes;
 * v_depths may be NULL (== zeros, as in the CLM loss).
 * ------------------------------------------------------------------------- */
int clmgs_projection_fwd(int C, int64_t N, const float *means, const float *quats,
                         const float *scales, const float *viewmats, const float *Ks, int width,
                         int height, float eps2d, float near_plane, float far_plane,
                         float radius_clip, int32_t *radii, float *means2d, float *depths,
                         float *conics, float *compensations, void *stream);
int clmgs_projection_bwd(int C, int64_t N, const float *means, const float *quats,
                         const float *scales, const float *viewmats, const float *Ks, int width,
                         int height, float eps2d, const int32_t *radii, const float *conics,
                         const float *v_means2d, const float *v_depths, const float *v_conics,
                         float *v_means, float *v_quats, float *v_scales, void *stream);

/* ---------------------------------------------------------------------------
 * Spherical harmonics.  compute_sh_fwd_tensor (CSRC/compute_sh_fwd.cu:31-58)
 * and compute_sh_bwd_inplace_tensor (CK/compute_sh_bwd.cu:309-344).
 * dirs [N,3], coeffs [N,K,3], masks [N] (uint8/bool) or NULL, colors [N,3].
 * bwd: v_coeffs [N,K,3]; accumulate != 0 gives the clm `+=` behaviour,
 * 0 gives gsplat's overwrite (bases above the degree are written as 0).
 * v_dirs [N,3] may be NULL.
 * ------------------------------------------------------------------------- */
int clmgs_sh_fwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                 const uint8_t *masks, float *colors, void *stream);
int clmgs_sh_bwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                 const uint8_t *masks, const float *v_colors, float *v_coeffs, int accumulate,
                 float *v_dirs, void *stream);

/* Engine-fused colour stage (engine.py:67-78 forward, :709-716 backward):
 * dirs = means - campos is formed on the fly, the forward applies
 * clamp_min(c + 0.5, 0), the backward re-derives the clamp mask (c + 0.5 >= 0,
 * torch.clamp_min's rule) instead of keeping an autograd graph, accumulates
 * (+=) into the retained SH gradient rows and adds v_dirs into v_means (+=). */
int clmgs_sh_colors_fwd(int64_t N, int K, int degree, const float *means, const float *campos,
                        const float *coeffs, float *colors, void *stream);
int clmgs_sh_colors_bwd(int64_t N, int K, int degree, const float *means, const float *campos,
                        const float *coeffs, const float *v_colors, float *v_coeffs,
                        float *v_means, void *stream);

/* ---------------------------------------------------------------------------
 * S4  Tile intersection, sort, offsets.  isect_tiles_tensor
 * (CSRC/isect_tiles.cu:100-232) and isect_offset_encode_tensor (:280-301).
 * Dense layout only ([C,N]); packed inputs are handled by passing camera_ids
 * (int64 [nnz]) with C = n_cameras and N = nnz.
 *
 * clmgs_isect_count: tiles_per_gauss [C*N] i32, cum_tiles [C*N] i64 (inclusive
 * prefix sum, == torch.cumsum in the reference), n_isects = cum_tiles[last]
 * (device scalar, also returned in *d_n_isects if non-NULL).
 * clmgs_isect_emit_sort: writes isect_ids [n_isects] i64 and flatten_ids
 * [n_isects] i32; if `sort`, they come out in the order a stable LSD radix
 * sort on the low 32+tile_n_bits+cam_n_bits key bits gives
 * (cub::DeviceRadixSort semantics, CSRC/isect_tiles.cu:196-203), computed as
 * depth sort of the Gaussians + stable (cam|tile) partition of the
 * intersections (same permutation, a quarter of the traffic).  scratch:
 * clmgs_isect_sort_scratch_bytes(total = C*N or nnz, n_isects).
 * ------------------------------------------------------------------------- */
size_t clmgs_isect_count_scratch_bytes(int64_t total);
int clmgs_isect_count(int64_t total, const float *means2d, const int32_t *radii, int tile_size,
                      int tile_width, int tile_height, int32_t *tiles_per_gauss,
                      int64_t *cum_tiles, void *scratch, size_t scratch_bytes, void *stream);
/* No counterpart in the reference (its torch extensions always allocate through
 * torch).  Same result as clmgs_isect_count for hosts that cannot give it
 * scratch memory: an embedding without a device allocator, a process whose HBM
 * was fully handed out at start-up, or a loop where cudaMalloc/cudaFree (which
 * may synchronise the device) is off limits.  The temporary lives in a 128 KiB
 * array inside the library image (held on every GPU that loads the library
 * until exit), so nothing is allocated by this call or by anything it calls.
 * total must be <= clmgs_isect_count_static_max() (16 Mi inputs); more is
 * rejected with an error and nothing is launched.  The first call on each
 * device resolves the array's device address -- make it before any stream
 * capture begins.  The array is shared by all calls on a device and keeps the
 * last call's contents (each call overwrites what it reads, nothing needs
 * clearing); calls that overlap on the GPU race on it and must be ordered. */
int64_t clmgs_isect_count_static_max(void);
int clmgs_isect_count_static(int64_t total, const float *means2d, const int32_t *radii,
                             int tile_size, int tile_width, int tile_height,
                             int32_t *tiles_per_gauss, int64_t *cum_tiles, void *stream);
size_t clmgs_isect_sort_scratch_bytes(int64_t total, int64_t n_isects);
int clmgs_isect_emit_sort(int C, int64_t N, const int64_t *camera_ids, const float *means2d,
                          const int32_t *radii, const float *depths, const int64_t *cum_tiles,
                          int tile_size, int tile_width, int tile_height, int64_t n_isects,
                          int sort, int64_t *isect_ids, int32_t *flatten_ids, void *scratch,
                          size_t scratch_bytes, void *stream);
int clmgs_isect_offset_encode(int64_t n_isects, const int64_t *isect_ids, int C, int tile_width,
                              int tile_height, int32_t *offsets, void *stream);

/* ---------------------------------------------------------------------------
 * S5  Rasterise.  rasterize_to_pixels_fwd_tensor (CSRC/rasterize_to_pixels_fwd.cu:236-267)
 * and rasterize_to_pixels_bwd_tensor (CSRC/rasterize_to_pixels_bwd.cu:347-381),
 * 3 colour channels (CLM uses CDIM=3, engine.py:106-117), tile_size 16.
 * means2d [C,N,2], conics [C,N,3], colors [C,N,3], opacities [C,N],
 * backgrounds [C,3] or NULL, tile_offsets [C,th,tw] i32, flatten_ids i32.
 * Outputs: render_colors [C,H,W,3], render_alphas [C,H,W,1], last_ids [C,H,W].
 * bwd accumulates (+=) into v_means2d, v_conics, v_colors, v_opacities (and
 * v_means2d_abs if non-NULL), which the caller zeroes.  v_render_alphas may be
 * NULL (== zeros, as in the CLM loss).
 * ------------------------------------------------------------------------- */
int clmgs_rasterize_fwd(int C, int64_t N, int64_t n_isects, const float *means2d,
                        const float *conics, const float *colors, const float *opacities,
                        const float *backgrounds, int width, int height, int tile_size,
                        int tile_width, int tile_height, const int32_t *tile_offsets,
                        const int32_t *flatten_ids, float *render_colors, float *render_alphas,
                        int32_t *last_ids, void *stream);
/* Measurement variant of clmgs_rasterize_fwd (no counterpart in the reference): same outputs, and
 * work_counters[64][3] (device int64, accumulated into -- zero them first; 64 partial sums per counter so that
 * the adds do not serialise: sum over the first index) receive
 *   [0] P_f: list entries the reference forward examines per pixel, summed (27 P_f FLOP),
 *   [1] entries this kernel evaluated per pixel after its tile-touch pruning, summed,
 *   [2] P_b: last blended entry - tile start + 1 per pixel, summed (75 P_b FLOP for the backward). */
int clmgs_rasterize_fwd_counted(int C, int64_t N, int64_t n_isects, const float *means2d,
                                const float *conics, const float *colors, const float *opacities,
                                const float *backgrounds, int width, int height, int tile_size,
                                int tile_width, int tile_height, const int32_t *tile_offsets,
                                const int32_t *flatten_ids, float *render_colors, float *render_alphas,
                                int32_t *last_ids, int64_t *work_counters, void *stream);

int clmgs_rasterize_bwd(int C, int64_t N, int64_t n_isects, const float *means2d,
                        const float *conics, const float *colors, const float *opacities,
                        const float *backgrounds, int width, int height, int tile_size,
                        int tile_width, int tile_height, const int32_t *tile_offsets,
                        const int32_t *flatten_ids, const float *render_alphas,
                        const int32_t *last_ids, const float *v_render_colors,
                        const float *v_render_alphas, float *v_means2d_abs, float *v_means2d,
                        float *v_conics, float *v_colors, float *v_opacities, void *stream);

/* ---------------------------------------------------------------------------
 * A9  Loss.  fusedssim + fusedssim_backward (CK/ssim.cu:368-444) fused with
 * loss_combined (strategies/base_engine.py:79-94, a torch.compile/Triton
 * kernel in the reference):  loss = (1-l)*mean|I-GT| + l*(1-mean ssim_map).
 * image is [H,W,3] (the rasteriser's layout, hwc=1) or [3,H,W] (hwc=0); gt_u8
 * is uint8 [3,H,W] (train.py:311).  Writes loss_out[0..2] = {loss, l1, ssim}
 * (device) and v_image = d loss / d image in the image's own layout.
 * scratch: clmgs_loss_scratch_bytes(H,W).
 * ------------------------------------------------------------------------- */
size_t clmgs_loss_scratch_bytes(int height, int width);
int clmgs_loss_fwd_bwd(int height, int width, const float *image, int hwc, const uint8_t *gt_u8,
                       float lambda_dssim, float *loss_out, float *v_image, void *scratch,
                       size_t scratch_bytes, void *stream);
/* Plain fused SSIM map mean (clm_kernels.fused_ssim forward, CK/clm_kernels/__init__.py:54-61)
 * on float [B,CH,H,W] inputs, for drop-in use; ssim_out is one float; v_img1
 * (optional) receives d mean(ssim_map) / d img1 (fusedssim_backward with
 * dL_dmap = 1/numel).  scratch: clmgs_loss_scratch_bytes(B*CH*H/3 + 1, W). */
int clmgs_fused_ssim(int B, int CH, int height, int width, const float *img1, const float *img2,
                     float *ssim_out, float *v_img1, void *scratch, size_t scratch_bytes,
                     void *stream);

/* ---------------------------------------------------------------------------
 * S2 / S6  Offloaded SH rows.  48 floats = 192 B per row.
 * send_shs2gpu_stream (CK/clm_kernels.cu:26-48), send_shs2gpu_stream_retention
 * (:92-126), send_shs2cpu_grad_buffer_stream (:150-175),
 * send_shs2cpu_grad_buffer_stream_retention (:230-268), set_signal (:280-289).
 * The grid/block arguments of the reference are dropped: the library sizes its
 * own launches.
 * ------------------------------------------------------------------------- */
int clmgs_rows_gather_i64(float *d_rows, const float *h_rows, const int64_t *ids, int64_t n,
                          void *stream);
int clmgs_rows_gather_retention(float *d_rows, const float *h_rows, const float *d_prev_rows,
                                const int32_t *host_ids, const int32_t *prev_src_rows,
                                const int32_t *dst_rows_from_host,
                                const int32_t *dst_rows_from_prev, int64_t n_host,
                                int64_t n_prev, void *stream);
int clmgs_rows_scatter_grad_i64(const float *d_grad, float *h_grad, const int64_t *ids,
                                int64_t n, int accum, void *stream);
int clmgs_rows_scatter_grad_retention(const float *d_grad, float *h_grad, float *d_next_grad,
                                      const int32_t *host_ids, const int32_t *next_dst_rows,
                                      const int32_t *src_rows_to_host,
                                      const int32_t *src_rows_to_next, int64_t n_host,
                                      int64_t n_next, int accum, void *stream);
int clmgs_set_signal(int32_t *h_flags, int index, int value, void *stream);
/* *h_dst = *d_src by a one-thread kernel (h_dst pinned + mapped).  For data-dependent sizes the host needs (the
 * tile total the reference reads with .item(), isect_tiles.cu:165-166): unlike a cudaMemcpyAsync it does not queue
 * behind bulk DMA transfers on the device-to-host copy engine.  Record an event after it and read *h_dst once the
 * event has fired. */
int clmgs_store_i64_to_host(int64_t *h_dst, const int64_t *d_src, void *stream);
/* Batch-resident SH rows ("slots"; new design replacing the per-camera re-packing of retained rows,
 * CK/clm_kernels.cu:71-126,206-268 + engine.py:519-641,767-803).  Every Gaussian the batch touches
 * owns one row of a device table for the whole batch: rows are loaded when their first camera comes
 * up and flushed after their last, so each row crosses the host link once each way per batch, with
 * plain stores (no read-modify-write over PCIe) and no device-to-device copies between cameras.
 *   clmgs_extract_first_use : key 0 = untouched, f+1 = first seen by the camera at position f of the
 *                             ORDERED bitmap (counterpart of clmgs_extract_ffs, which gives the last use)
 *   clmgs_slot_map          : slot_of[ids[p]] = p for the touched Gaussians grouped by first use
 *   clmgs_rows_gather_i32   : clmgs_rows_gather_i64 with int32 ids (a first-use group -> its contiguous slots)
 *   clmgs_rows_gather_to_slots : d_rows[slot_of[ids[r]]] = h_rows[ids[r]] (slots in an order other than the
 *                             first-use grouping: the multi-GPU engine keeps them in ascending-id = owner-major order)
 *   clmgs_rows_flush_slots  : h_grad[ids[r]] = d_grad_rows[slot_of[ids[r]]] (a last-use group, stored once)
 *   clmgs_sh_colors_*_slots : the engine-fused SH colour stage reading / accumulating rows through
 *                             slot_of[ids[g]] (K = 16, degree <= 3). */
int clmgs_extract_first_use(const void *bitmap, int word_bytes, int bsz, uint8_t *out, int64_t N,
                            void *stream);
int clmgs_slot_map(const int32_t *ids, int64_t n, int32_t *slot_of, void *stream);
int clmgs_rows_gather_i32(float *d_rows, const float *h_rows, const int32_t *ids, int64_t n,
                          void *stream);
int clmgs_rows_gather_to_slots(float *d_rows, const float *h_rows, const int32_t *ids, int64_t n,
                               const int32_t *slot_of, void *stream);
int clmgs_rows_flush_slots(const float *d_grad_rows, const int32_t *slot_of, const int32_t *ids,
                           int64_t n, float *h_grad, void *stream);
int clmgs_sh_colors_fwd_slots(int64_t N, int degree, const float *means, const float *campos,
                              const float *d_rows, const int64_t *ids, const int32_t *slot_of,
                              float *colors, void *stream);
int clmgs_sh_colors_bwd_slots(int64_t N, int degree, const float *means, const float *campos,
                              const float *d_rows, const float *v_colors, float *d_grad_rows,
                              const int64_t *ids, const int32_t *slot_of, float *v_means,
                              void *stream);
/* Multi-GPU owner-side sum (new design; the reference is single-GPU):
 * d_dst_rows[dst_index[r]] += d_src_rows[r] for 48-float rows in device memory;
 * indices may repeat (several ranks touched the same row). */
int clmgs_rows_accumulate(float *d_dst_rows, const float *d_src_rows, const int64_t *dst_index,
                          int64_t n, void *stream);

/* ---------------------------------------------------------------------------
 * A2  Camera ordering / retention bookkeeping bitmaps.
 * scatter_to_bit (CK/clm_kernels.cu:362-421), extract_ffs (:291-360),
 * compute_cnt_h (:423-627).  word_bytes in {1,2,4,8}.
 * clmgs_compute_cnt_d writes cnt_d[k] = |F_k & F_{k+1}| for k < bsz-1 directly
 * (int32 [bsz-1], zeroed inside) instead of the reference's (bsz-1, 16384)
 * per-thread partials + torch.sum.
 * clmgs_bitmap_permute re-encodes a cull bitmap (bit c = input camera c) into
 * the ordered convention of engine.py:151-153,199-201: camera order[i] -> bit
 * bsz-1-i, replacing zero_() + bsz scatter_to_bit launches.
 * ------------------------------------------------------------------------- */
int clmgs_scatter_to_bit(void *bitmap, int word_bytes, const int64_t *ids, int64_t n, int bit,
                         void *stream);
int clmgs_extract_ffs(const void *bitmap, int word_bytes, uint8_t *out, int64_t N, void *stream);
int clmgs_compute_cnt_d(const void *bitmap, int word_bytes, int bsz, int64_t N, int32_t *cnt_d,
                        void *stream);
int clmgs_bitmap_permute(const void *bitmap_in, void *bitmap_out, int word_bytes, int bsz,
                         const int32_t *order /* host, [bsz] */, int64_t N, void *stream);
/* Sampled pairwise XOR distance matrix of engine.py:159-175: dist[i*bsz+j] =
 * #samples where cameras i and j differ.  Here bit c of a word = camera c.
 * dist must hold 2*bsz*bsz int32: the first bsz*bsz are the matrix, the rest
 * is working space (co-visibility counts).  Zeroed inside. */
int clmgs_bitmap_distance(const void *bitmap, int word_bytes, int bsz, const int64_t *sample_ids,
                          int64_t n_samples, int32_t *dist, void *stream);
/* Groups Gaussian ids by ffs value (the torch.sort + unique_consecutive +
 * split of engine.py:203-213): out_ids [N] int32 grouped by key 0..bsz, with
 * group_counts [bsz+1] int64; ids ascending inside each group (stable).
 * scratch: clmgs_group_by_ffs_scratch_bytes(N). */
size_t clmgs_group_by_ffs_scratch_bytes(int64_t N);
int clmgs_group_by_ffs(const uint8_t *ffs, int64_t N, int bsz, int32_t *out_ids,
                       int64_t *group_counts, void *scratch, size_t scratch_bytes, void *stream);
/* Retention without index lists (new design; replaces the nonzero_static /
 * gather / scatter_ glue of engine.py:519-641 and :767-803 plus the two
 * transfer launches).  `bitmap` is the ORDERED bitmap (camera at position k of
 * the processing order <-> bit bsz-1-k, engine.py:151-153); ids_this/ids_next
 * are the ascending filters of positions k and k+1.
 *
 * clmgs_rows_prefetch_next: for every row r' of the next camera, copies its
 * 192-B SH row either device->device from the current camera's buffer (the
 * Gaussian is also visible now: rank found by binary search in ids_this) or
 * from the pinned host table with 16-byte zero-copy loads.  src_map[r'] gets
 * the source row in the current buffer, or -1 for rows that came from host.
 *
 * clmgs_rows_grad_handoff: after the backward of camera k: next_grad[r'] =
 * grad_this[src_map[r']] (retained rows) or 0 (rows new to the GPU); rows of
 * camera k that camera k+1 does not see go to the pinned gradient table
 * (accum 0: store; 1: += over the link, the reference's read-modify-write;
 * 2: += only for rows already flushed once in this batch -- known from the
 * bitmap -- and a plain store otherwise, which requires the table rows to be
 * zero at batch start, as cpu-adam leaves them).  n_next == 0 flushes every
 * row (last camera, engine.py:812-822). */
int clmgs_rows_prefetch_next(float *d_next_rows, const float *h_rows, const float *d_this_rows,
                             const int64_t *ids_this, int64_t n_this, const int64_t *ids_next,
                             int64_t n_next, const void *bitmap, int word_bytes, int bit_this,
                             int32_t *src_map, void *stream);
int clmgs_rows_grad_handoff(const float *d_grad_this, float *d_next_grad, float *h_grad,
                            const int64_t *ids_this, int64_t n_this, int64_t n_next,
                            const int32_t *src_map, const void *bitmap, int word_bytes,
                            int bit_this, int bit_next, int accum, void *stream);

/* ---------------------------------------------------------------------------
 * A16  GPU Adam.  selective_adam_update (CK/adam.cu:43-78; masked, no bias
 * correction) and the dense torch.optim.Adam(fused=True) rule with bias
 * correction (optimizer.py:127-129).  grad_scale multiplies the gradient
 * first (the `param.grad /= bsz` of engine.py:870-872, fused).
 * ------------------------------------------------------------------------- */
int clmgs_selective_adam(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                         const uint8_t *mask, float lr, float b1, float b2, float eps,
                         float grad_scale, int64_t N, int M, void *stream);
int clmgs_adam_dense(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, float lr,
                     float b1, float b2, float eps, float grad_scale, int64_t step, int64_t n,
                     void *stream);

/* ---------------------------------------------------------------------------
 * A15  Host Adam for the offloaded rows (HOST code, runs on the CPU by design:
 * the reference's cpu-adam, submodule TarzanZhao/cpu-adam @ f169c36, source
 * not in the mount; call sites optimizer.py:130-142, engine.py:301-335).
 * Blocking; releases nothing itself -- call it from a thread that does not
 * hold the GIL (ctypes does that).  Work is given as n_segments segments
 * processed in order: segment s = host_rows[seg_begin[s] .. seg_end[s]) of the
 * pinned [N,cols] tensors; before touching it every worker waits (spin with
 * pause) until h_signal[seg_signal[s]] != 0, or not at all if seg_signal[s] < 0.
 * The reference's batched_sparse_step(batch_size, lists, signals) is the case
 * "segment m = rows whose last user is camera m-1, waiting on flag m-1; segment
 * 0 = untouched rows, no wait, omitted when sparse_adam"; a caller may split a
 * segment further and signal the pieces separately (the engine does this for
 * the last camera so that the update overlaps the final gradient flush).
 * Gradients are multiplied by `scale`, and zeroed in place afterwards
 * (the reference's version=3).  n_threads <= 0 picks hardware concurrency.
 * seg_zero_grad (optional): non-zero marks a segment whose gradient rows are known to be
 * zero (rows no camera touched, in dense mode): the same update is applied without reading
 * or re-zeroing the gradient table (a quarter less DRAM traffic for those rows).
 * host_progress (optional): set to s+1 once segments 0..s are complete, so another
 * host thread can follow the update (used to start loading a batch's rows while the
 * previous batch's update is still running).
 * ------------------------------------------------------------------------- */
int clmgs_cpu_adam_batched_sparse_step(float *host_param, float *host_grad, float *host_exp_avg,
                                       float *host_exp_avg_sq, int cols, const float *col_lr,
                                       const int32_t *host_rows, const int64_t *seg_begin,
                                       const int64_t *seg_end, const int32_t *seg_signal,
                                       const int32_t *seg_zero_grad, int n_segments,
                                       const volatile int32_t *h_signal, float b1,
                                       float b2, float eps, int64_t step, float scale,
                                       int n_threads, volatile int64_t *host_progress);
/* Same update with the gradients in a PACKED pinned staging buffer instead of the [N, cols] table (new; the
 * reference accumulates into the table over PCIe, CK/clm_kernels.cu:128-148, and cpu-adam zeroes it afterwards):
 * the gradient of list entry r is row (r - packed_base) of h_grad_packed.  The engine's flush fills that buffer
 * with contiguous DMA copies (copy engine, no SMs, no read-modify-write), the host reads every row once and does
 * not zero it again; the gradient table stays untouched (all zero).  Segments flagged in seg_zero_grad may lie
 * before packed_base; all others must not. */
int clmgs_cpu_adam_batched_sparse_step_packed(float *host_param, float *h_grad_packed, int64_t packed_base,
                                              float *host_exp_avg, float *host_exp_avg_sq, int cols,
                                              const float *col_lr, const int32_t *host_rows,
                                              const int64_t *seg_begin, const int64_t *seg_end,
                                              const int32_t *seg_signal, const int32_t *seg_zero_grad,
                                              int n_segments, const volatile int32_t *h_signal, float b1,
                                              float b2, float eps, int64_t step, float scale, int n_threads,
                                              volatile int64_t *host_progress);

/* Statistics of this process's last clmgs_cpu_adam_batched_sparse_step call (measurement; no counterpart
 * in the reference apart from --log_cpu_adam_trailing_overhead, engine.py:888-917): host_out4 =
 * {wall ms of the call, ms of it spent waiting for gradient flags (slowest thread), rows updated,
 * of which rows that took the zero-gradient shortcut}. */
int clmgs_cpu_adam_last_stats(double *host_out4);

/* What the rasteriser consumes, in one call: clmgs_isect_emit_sort (sorted) + clmgs_isect_offset_encode
 * without materialising the 64-bit isect_ids -- the last radix pass writes flatten_ids [n_isects] at the
 * sorted positions and marks the first position of every (camera, tile) key; tile_offsets [C, th, tw] is
 * identical to isect_offset_encode's (CSRC/isect_tiles.cu:234-301).  Scratch as clmgs_isect_emit_sort. */
int clmgs_isect_sorted_offsets(int C, int64_t N, const int64_t *camera_ids, const float *means2d,
                               const int32_t *radii, const float *depths, const int64_t *cum_tiles,
                               int tile_size, int tile_width, int tile_height, int64_t n_isects,
                               int32_t *flatten_ids, int32_t *tile_offsets, void *scratch,
                               size_t scratch_bytes, void *stream);

/* Stable LSD radix sort of (uint32 key, int32 value) pairs on key bits [0, end_bit), 8-bit digits,
 * hand-written (csrc/radix.cu) -- the sort under clmgs_isect_emit_sort, which replaces the reference's
 * cub::DeviceRadixSort::SortPairs call (CSRC/isect_tiles.cu:186-229).  Ping-pongs between the two buffer
 * pairs; *result_in_alt = 1 if the sorted pairs ended up in (keys_alt, vals_alt).  Scratch holds the per-tile
 * digit histograms. */
size_t clmgs_radix_sort_pairs_u32_scratch_bytes(int64_t n);
int clmgs_radix_sort_pairs_u32(int64_t n, uint32_t *keys, uint32_t *keys_alt, int32_t *vals,
                               int32_t *vals_alt, int end_bit, void *scratch, size_t scratch_bytes,
                               int32_t *result_in_alt, void *stream);

/* ---------------------------------------------------------------------------
 * Densify / prune / opacity reset (the non-iteration step
 * that runs every densification_interval iterations, densification.py:5-56).
 * Replaces GaussianModelCLMOffload.densify_and_prune (strategies/
 * base_gaussian_model.py:364-388), densify_and_clone / densify_and_split
 * (strategies/clm_offload/gaussian_model.py:801-831, 741-799), prune_points
 * (:578-611, _prune_unified_adam :522-576), cat_tensors_to_unified_adam
 * (:613-700) and reset_opacity (:488-520).
 * ------------------------------------------------------------------------- */

/* One pass over the N current rows -> class byte per row: bit0 clone (|grad| >=
 * grad_threshold and max(exp(log_scale)) <= dense_threshold), bit1 split (grad >=
 * grad_threshold and max scale > dense_threshold), bit2 "this row and its clone are
 * pruned" (sigmoid(opacity) < min_opacity, or use_world_size and max scale >
 * world_size_threshold), bit3 "its split children are pruned" (same test on the
 * children's scale exp(log(s / 1.6))).  grad = grad_accum / denom with NaN -> 0. */
int clmgs_densify_classify(int64_t N, const float *grad_accum, const float *denom, const float *log_scale,
                           const float *opacity_logit, float grad_threshold, float dense_threshold,
                           float min_opacity, float world_size_threshold, int use_world_size,
                           uint8_t *cls, void *stream);

/* prune_points(mask) alone: class byte = "pruned" where prune_mask[i] != 0. */
int clmgs_prune_mask_to_class(int64_t N, const uint8_t *prune_mask, uint8_t *cls, void *stream);

/* Class bytes -> the final row order.  totals[4] (device) = {nA kept rows, nB kept
 * clones, nS split parents, nC split parents whose children are kept}.
 * src[nA + nB + 2 nC] (caller sizes it N + N + 2 N at most, or reads totals first):
 * source row of every output row, in the reference's order (kept originals, clones,
 * children copy 0, children copy 1; each ascending).  noise_row[nC]: rank of each
 * kept split parent among ALL split parents (the reference draws one normal sample
 * per split parent and copy, GM:761-763: child (copy w, kept parent r) uses
 * noise[w * nS + noise_row[r]]). */
size_t clmgs_densify_plan_scratch_bytes(int64_t N);
int clmgs_densify_plan(int64_t N, const uint8_t *cls, int32_t *src, int32_t *noise_row, int64_t *totals,
                       void *scratch, size_t scratch_bytes, void *stream);

/* Builds the new GPU-resident tensors and their Adam moments out of place.
 * in12 / out12: HOST arrays of 12 device pointers {xyz[.,3], opacity[.,1],
 * log_scale[.,3], quat[.,4], then exp_avg of the four, then exp_avg_sq of the four}.
 * Kept rows carry their moments, appended rows start at zero.  Children: xyz =
 * R(quat) (exp(log_scale) * noise) + xyz, log_scale = log(exp(log_scale) / 1.6),
 * opacity and quat copied (GM:761-776).  noise [2 nS, 3] standard normal. */
int clmgs_densify_apply(int64_t nA, int64_t nB, int64_t nC, int64_t nS, const int32_t *src,
                        const int32_t *noise_row, const float *noise, const float *const *in12,
                        float *const *out12, void *stream);

/* Pinned tables [capacity, cols] (SH coefficients + both host Adam moments) compacted in
 * place on the host threads: rows host_src[0..nA) kept (strictly ascending), then n_tail
 * appended rows copying the coefficients of host_src[nA..nA+n_tail) with zero moments.
 * HOST function (no stream); the caller has copied src to the host. */
int clmgs_host_rows_densify(float *host_param, float *host_exp_avg, float *host_exp_avg_sq, int cols,
                            int64_t n_old, int64_t capacity, const int32_t *host_src, int64_t nA,
                            int64_t n_tail, int n_threads);

/* reset_opacity (GM:488-520): opacity_logit <- logit(min(sigmoid(opacity_logit), max_opacity)),
 * Adam moments of the opacity <- 0. */
int clmgs_reset_opacity(int64_t N, float *opacity_logit, float *exp_avg, float *exp_avg_sq,
                        float max_opacity, void *stream);

/* Lazy dense Adam -- a scheduling change on top of the reference's dense host Adam, which updates every row of the
 * table at every step (optimizer.py:130-142 + engine.py:320-327): zero-gradient steps of rows no camera touched are postponed and
 * replayed -- bit-identically, same per-step bias corrections -- when the row is next needed.  last_step[row] = the
 * step through which the row is up to date.  Brings host_rows[0..n_rows) (all rows 0..n_rows if NULL) to
 * through_step, then marks them mark_step.  HOST function. */
int clmgs_cpu_adam_lazy_catch_up(float *host_param, float *host_exp_avg, float *host_exp_avg_sq, int cols,
                                 const float *col_lr, const int32_t *host_rows, int64_t n_rows,
                                 int32_t *last_step, int64_t through_step, int64_t mark_step, float b1,
                                 float b2, float eps, int n_threads);

/* fast_tsp.find_tour(dist, 0.001) (engine.py:179; submodule TarzanZhao/fast-tsp
 * @ 069eb3d, source not in the mount): deterministic greedy + 2-opt/Or-opt
 * open... closed tour over n <= 64 cameras, int32 dist [n*n]; writes tour[n]. */
int clmgs_tsp_tour(const int32_t *host_dist, int n, int32_t *host_tour);

/* ---------------------------------------------------------------------------
 * Measurement helpers (no counterpart in the reference; SURVEY.md 8d asks for the
 * roofline denominators to be measured on the box): an FP32 FMA chain kernel
 * (one launch = grid_ctas x 1024 threads x iters x 128 FMAs), a plain
 * cudaMemcpyAsync (pinned host <-> device = the host-link peak) and a host STREAM
 * triad a = b + s c over n floats on n_threads threads (blocking).  bench.py and
 * tools/peaks.py time them; nothing on the training path calls them. */
int clmgs_measure_fma(int grid_ctas, int iters, float *sink, void *stream);
int clmgs_measure_memcpy(void *dst, const void *src, size_t bytes, void *stream);
int clmgs_measure_host_triad(float *host_a, const float *host_b, const float *host_c, int64_t n, float s,
                             int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* CLMGS_H */
