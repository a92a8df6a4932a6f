/*
 * isogs_rast.h -- C ABI of the B200-native (sm_100a) differentiable Gaussian
 * rasterizer and the SLAM step kernels around it.
 *
 * Drop-in boundary for the hot path of mzli112358/IsoGS-SLAM-cpp.  Every entry
 * point takes raw DEVICE pointers, sizes and a cudaStream_t (passed as void*);
 * no torch types, no allocation, no exceptions.  Return value: 0 on success,
 * negative ISOGS_E_* on error (text via isogs_last_error()).  All arrays are
 * contiguous row-major fp32 unless stated.  The library is re-entrant per
 * (device, stream) and keeps no global device state.
 *
 * Each entry point names the reference interface it replaces (file:line under
 * /root/reference).
 */
#ifndef ISOGS_RAST_H
#define ISOGS_RAST_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ISOGS_REC 12   /* floats per projected record and per gradient record */
#define ISOGS_TILE 16

#define ISOGS_OK 0
#define ISOGS_E_ARG (-1)      /* bad argument */
#define ISOGS_E_CUDA (-2)     /* CUDA runtime error */
#define ISOGS_E_CAPACITY (-3) /* intersection buffers too small: *n_isects holds the need */

/* flags (match the oracle's) */
#define ISOGS_F_LOG_SCALES 1     /* scales are log-scales (python_ordinary/utils/slam_helpers.py:136) */
#define ISOGS_F_LOGIT_OPAC 2     /* opacities are logits  (slam_helpers.py:135) */
#define ISOGS_F_NO_OPAC_EXTENT 4 /* fixed 3.33 sigma extent (gsplat _torch_impl.py:356) */

/* GaussianRasterizationSettings reduced to what the kernels need
 * (python_ordinary/utils/recon_helpers.py:4-27; gsplat Ks/viewmats of Ops.h:42-64). */
typedef struct isogs_camera {
    int32_t width, height;
    float fx, fy, cx, cy;
    float R[9]; /* row-major world->camera rotation of the view matrix */
    float t[3];
    float near_plane, far_plane, eps2d;
    float bg[3];
} isogs_camera;

/* per-frame pose applied before the view matrix == transform_to_frame
 * (python_ordinary/utils/slam_helpers.py:252-304).  HOST struct. */
typedef struct isogs_pose {
    float q[4]; /* w,x,y,z un-normalised */
    float t[3];
} isogs_pose;

/* DEVICE-resident step state of the fused Adam: advanced by isogs_step_begin, read by isogs_adam_fused, so a
 * captured CUDA graph of the step stays replayable (torch.optim.Adam keeps `step` on the host and the
 * reference re-creates the optimizer per frame, python_ordinary/scripts/splatam.py:402-408,1237,1406: zero this
 * struct to do the same).  32 bytes, 8-byte aligned. */
typedef struct isogs_adam_state {
    int64_t step;
    double inv_bc1;      /* 1 / (1 - beta1^step) */
    double inv_bc2_sqrt; /* 1 / sqrt(1 - beta2^step) */
    double reserved;
} isogs_adam_state;

/* DEVICE-resident "current frame" slot read by the fused steps: ground-truth planes, the frame's pose (the
 * keyframe's fixed pose for mapping; ignored by tracking, whose pose is the optimised variable) and
 * 1 / #(gt_depth > 0).  Lets ONE captured graph serve every keyframe of the window: the caller copies 48 bytes
 * into the slot (stream-ordered) before each launch.  gt_rgb [3,H,W] and gt_depth [H,W] must be 16-byte aligned. */
typedef struct isogs_frame {
    const float* gt_rgb;
    const float* gt_depth;
    float pose[7]; /* q (w,x,y,z un-normalised), t */
    float inv_n_depth;
} isogs_frame;

const char* isogs_last_error(void);
int isogs_version(void);

/* ---- rasterizer forward, stage by stage ------------------------------------------------ */

/* Bytes of scratch the scan / sort stages need for N Gaussians and `cap` intersections. */
int isogs_temp_bytes(int64_t N, int64_t cap, size_t* scan_bytes, size_t* sort_bytes);

/* preprocess: replaces diff_gaussian_rasterization's preprocessCUDA ==
 * gsplat projection_ewa_3dgs_fused_fwd (Ops.h:42-64; ProjectionEWA3DGSFused.cu:16-212)
 * + intersect_tile pass 1 (IntersectTile.cu:66-84) + cumsum (Intersect.cpp:79).
 * scale_dim 1 = isotropic [N,1]; 3 = [N,3] with quats [N,4].
 * Alignment: rec, quats (and v_rec, v_quats in the backward) are read / written as float4 rows and must be
 * 16-byte aligned, radii 8-byte aligned (ISOGS_E_ARG otherwise); every other array is 4-byte aligned.
 * pose (HOST, nullable) or pose_dev (DEVICE float[7] q,t; nullable) is applied first.
 * ch 4: record colours {c0,c1,c2,z}; ch 5: {c0,c1,c2,z,z*z}.
 * Outputs: rec [N,12], radii [N,2] i32, tiles [N] i32, cum [N] i32 (inclusive scan of tiles).
 * If n_isects_host (PINNED or pageable HOST int64*) is non-null the call copies cum[N-1]
 * there and synchronises the stream (the one sync of a forward, == Intersect.cpp:80).
 * tiles and cum may both be null (then n_isects_host must be null too): the sync-free pipeline
 * (isogs_tile_bin_async) counts intersections per tile itself and needs neither.
 * v_rec_zero (nullable, [N,12], 16-byte aligned): rows are cleared by the same kernel -- the gradient record
 * that isogs_blend_bwd accumulates into, so the step needs no separate memset launch.
 * tile_rects (nullable, [N,4] u32, 16-byte aligned): {x0 | x1 << 16, y0 | y1 << 16, depth bits, 0} -- the tile rect of
 * intersect_tile (IntersectTile.cu:66-77; empty for a culled splat) and the key's depth bits, so that the binning
 * kernels stream 16 bytes per Gaussian.  tile_counts (nullable, DEVICE int32[n_tiles], ZEROED by the caller): the
 * per-tile intersection counts are accumulated by the same kernel (CTA-private shared-memory bins, one global
 * atomic per touched tile) -- hand both to isogs_tile_bin_async with counts_ready = 1. */
int isogs_preprocess_fwd(const isogs_camera* cam, const isogs_pose* pose, const float* pose_dev,
                         int64_t N, const float* means, const float* scales, int scale_dim,
                         const float* quats, const float* opacities, const float* colors,
                         int ch, int flags,
                         float* rec, int32_t* radii, int32_t* tiles, int32_t* cum,
                         void* scan_temp, size_t scan_temp_bytes,
                         int64_t* n_isects_host, float* v_rec_zero, uint32_t* tile_rects, int32_t* tile_counts,
                         void* stream);

/* duplicateWithKeys + radix sort + identifyTileRanges: replaces gsplat intersect_tile pass 2,
 * radix_sort_double_buffer and intersect_offset (Ops.h:186-204; IntersectTile.cu:87-113,
 * 296-339, 209-257).  keys_* [M] u64, vals_* [M] i32, tile_ranges [n_tiles+1] i32. */
int isogs_tile_bin(const isogs_camera* cam, int64_t N, int64_t M,
                   const float* rec, const int32_t* radii, const int32_t* cum,
                   uint64_t* keys_unsorted, int32_t* vals_unsorted,
                   uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges,
                   void* sort_temp, size_t sort_temp_bytes, void* stream);

/* Same arguments and bit-identical outputs, computed by the LITERAL pipeline north_star (2) names: emit the (tile | depth)
 * keys (intersect_tile pass 2, IntersectTile.cu:87-113), cub::DeviceRadixSort::SortPairs over the used key bits
 * (:296-339), identifyTileRanges (:209-257).  isogs_tile_bin reaches the same result by a counting sort by tile plus a
 * per-tile depth sort, in about half the time; this entry is the cross-check and the reference's own kernel sequence.
 * cum must be given. */
int isogs_tile_bin_cub(const isogs_camera* cam, int64_t N, int64_t M,
                       const float* rec, const int32_t* radii, const int32_t* cum,
                       uint64_t* keys_unsorted, int32_t* vals_unsorted,
                       uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges,
                       void* sort_temp, size_t sort_temp_bytes, void* stream);

/* Sync-free variant of isogs_tile_bin for the fused step: removes the blocking
 * `n_isects = cum[-1].item()` of gsplat Intersect.cpp:80.  The intersection count is NOT known on
 * the host; `capacity` is the size of bucket / keys_sorted / vals_sorted.  totals (DEVICE int32[2])
 * receives {n_isects, overflow}: when n_isects > capacity the tile lists are truncated to the
 * capacity (no out-of-bounds access) and overflow is set to 1.  The overflow word is STICKY: the library never
 * clears it (while it is set totals[0] keeps the largest count seen); hand `totals + 1` to isogs_adam_fused /
 * isogs_pose_keep_best as skip_flag so the affected steps update nothing, and clear it on the host after
 * growing the buffers.  bucket [capacity] u64 is scratch; keys_sorted may be null.  Output order is
 * bit-identical to isogs_tile_bin.  tile_rects (nullable): the rows isogs_preprocess_fwd wrote (else the rects are
 * re-derived from rec / radii); counts_ready = 1: the first n_tiles ints of sort_temp already hold the per-tile
 * counts (isogs_preprocess_fwd's tile_counts pointed there), so the count pass is skipped. */
int isogs_tile_bin_async(const isogs_camera* cam, int64_t N, int64_t capacity,
                         const float* rec, const int32_t* radii,
                         uint64_t* bucket, uint64_t* keys_sorted /*nullable*/, int32_t* vals_sorted,
                         int32_t* tile_ranges, int32_t* totals,
                         void* sort_temp, size_t sort_temp_bytes,
                         const uint32_t* tile_rects /*nullable*/, int counts_ready, void* stream);

/* renderCUDA forward: replaces rasterize_to_pixels_3dgs_fwd (Ops.h:223-238;
 * RasterizeToPixels3DGSFwd.cu:17-188).  render [ch,H,W] planar (the Inria API's CHW),
 * alpha [H,W] = 1 - T_final (the silhouette), last [H,W] = sorted index of the last contributor. */
int isogs_blend_fwd(const isogs_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                    const int32_t* tile_ranges, float* render, float* alpha, int32_t* last,
                    void* stream);

/* ---- rasterizer backward --------------------------------------------------------------- */

/* renderCUDA backward: replaces rasterize_to_pixels_3dgs_bwd (Ops.h:239-264;
 * RasterizeToPixels3DGSBwd.cu:15-278).  v_rec [N,12] must be ZEROED by the caller; receives
 * v_x,v_y,v_conic_a,v_conic_b,v_conic_c,v_opacity,v_c0..v_c4,0 by atomic accumulation. */
int isogs_blend_bwd(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                    const int32_t* vals_sorted, const int32_t* tile_ranges,
                    const float* alpha, const int32_t* last,
                    const float* v_render, const float* v_alpha /*nullable*/,
                    float* v_rec, void* stream);

/* Tracking variant of isogs_blend_bwd (same arguments): only the pose gets a gradient in the tracking loop
 * (python_ordinary/scripts/splatam.py:1247-1290, transform_to_frame(gaussians_grad=False)), and the pose sees
 * the record through {x, y, conic, z, z*z} alone, so only v_rec[0..4] and v_rec[9..10] are accumulated; the
 * colour / opacity slots keep the caller's zeros.  Feed the result to isogs_preprocess_bwd with v_pose. */
int isogs_blend_bwd_geom(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                         const int32_t* vals_sorted, const int32_t* tile_ranges,
                         const float* alpha, const int32_t* last,
                         const float* v_render, const float* v_alpha /*nullable*/,
                         float* v_rec, void* stream);

/* preprocess backward: replaces projection_ewa_3dgs_fused_bwd (Ops.h:65-88;
 * ProjectionEWA3DGSFused.cu:294-531) and, when a pose is given, the autograd backward of
 * transform_to_frame (slam_helpers.py:252-304) including the N->7 pose reduction.
 * Any output pointer may be null (gradient not needed).  v_pose: DEVICE float[7]
 * (d/dq un-normalised, d/dt), must be zeroed by the caller; pose_acc: DEVICE float[16] scratch,
 * zeroed by the caller.  v_means2d [N,3] (nullable) receives the Inria-style screen-space
 * gradient side channel (splatam.py:528-530) = v_xy * (W/2, H/2). */
int isogs_preprocess_bwd(const isogs_camera* cam, const isogs_pose* pose, const float* pose_dev,
                         int64_t N, const float* means, const float* scales, int scale_dim,
                         const float* quats, const float* opacities, int ch, int flags,
                         const int32_t* radii, const float* v_rec,
                         float* v_means, float* v_scales, float* v_quats, float* v_opac,
                         float* v_colors, float* v_means2d,
                         float* v_pose, float* pose_acc, void* stream);

/* ---- the step around it ---------------------------------------------------------------- */

/* fused masked L1 colour/depth loss on the 5-channel render: replaces get_loss's mask + L1
 * section (python_ordinary/scripts/splatam.py:534-570,747-749).  mode 0 tracking (sums over
 * gt_d>0 & !nan & sil>thres), mode 1 mapping (means).  Writes v_render [5,H,W] and
 * losses (DEVICE float[4], zeroed by the caller) = {total, colour, depth, n_valid_depth}.
 * For mode 1 n_valid (number of gt_depth>0 pixels) must be supplied by the caller as
 * inv_n_depth = 1/n_valid (it is a per-frame constant).
 * outlier_thresh (DEVICE float*, nullable; mode 0 only): ignore_outlier_depth_loss of splatam.py:551-554 --
 * pixels with |gt_d - d| >= *outlier_thresh leave the depth AND colour masks; fill it with
 * isogs_depth_outlier_threshold. */
int isogs_loss_l1_fused(const isogs_camera* cam, int mode, const float* render, const float* alpha,
                        const float* gt_rgb, const float* gt_depth, float w_im, float w_depth,
                        float sil_thres, int use_sil, float inv_n_depth, const float* outlier_thresh,
                        float* v_render, float* losses, void* stream);

/* *thresh_out = factor * median(|gt_d - d| * (gt_d > 0)) over all pixels (torch.median = lower median;
 * splatam.py:552-553 uses factor 10).  render_depth: DEVICE [H,W] plane (render + 3*H*W); scratch: DEVICE
 * float[2*H*W]; sort_temp: isogs_temp_bytes(0, H*W, ...) sort bytes suffice.  No host sync. */
int isogs_depth_outlier_threshold(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                                  float factor, float* scratch, void* sort_temp, size_t sort_temp_bytes,
                                  float* thresh_out, void* stream);

/* Allocation-free variant of isogs_depth_outlier_threshold.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  The scratch (two H*W float planes + the sort temp) is a file-scope device array inside the
 * library -- ISOGS_NOALLOC_BYTES on every GPU that loads it, until exit -- so the caller passes inputs and the
 * output only.  Images of more than ISOGS_NOALLOC_MAX_PIXELS pixels are rejected with ISOGS_E_ARG.  The call
 * allocates nothing and does not synchronise.  All calls share the one array: calls that overlap on the GPU
 * (different streams or threads) race; serialise them.  Same result, bit for bit, as the entry above. */
#define ISOGS_NOALLOC_MAX_PIXELS (1 << 21)                                   /* >= 1752 x 1168 */
#define ISOGS_NOALLOC_BYTES (8ull * ISOGS_NOALLOC_MAX_PIXELS + (4ull << 20)) /* 20 MiB */
int isogs_depth_outlier_threshold_noalloc(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                                          float factor, float* thresh_out, void* stream);

/* fused SSIM term of the mapping colour loss 0.8*L1 + 0.2*(1 - SSIM): replaces calc_ssim / _ssim
 * (python_ordinary/utils/slam_external.py:54-97; call site splatam.py:568) and its autograd backward
 * (5 depthwise 11x11 conv2d, zero padding, sigma 1.5, c1 = 1e-4, c2 = 9e-4, mean over [3,H,W]).
 * render / gt_rgb: DEVICE [>=3,H,W] planes (the first three are used).  dmaps: DEVICE scratch [9,H,W].
 * ssim_out: DEVICE float[1], += mean SSIM (caller zeroes).  losses (nullable): losses[0] += weight*(1-SSIM).
 * v_render (nullable): planes 0..2 += d(weight*(1-SSIM))/d(render) -- call AFTER isogs_loss_l1_fused, which
 * overwrites v_render (pass it w_im = 0.8 * w and this call weight = 0.2 * w for the reference's mix). */
int isogs_ssim_loss_fused(const isogs_camera* cam, const float* render, const float* gt_rgb, float weight,
                          float* dmaps, float* v_render, float* losses, float* ssim_out, void* stream);

/* fused multi-tensor Adam with torch.optim.Adam semantics: replaces launch_adam_update
 * (cpp/cuda/kernels/optimizer.cu:15-98), gsplat adam (Ops.h:173-183) and the Python
 * torch.optim.Adam of splatam.py:402-408.  One launch covers n_groups tensors. */
typedef struct isogs_adam_group {
    float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
    int64_t numel; float lr;
} isogs_adam_group;
/* `step` (>= 1) is torch's host-side step count; if `state` (DEVICE, nullable) is given the bias corrections are
 * read from it instead (see isogs_step_begin) and `step` is ignored.  skip_flag (DEVICE int32*, nullable): the
 * launch is a no-op when *skip_flag != 0 (the sticky overflow word of isogs_tile_bin_async). */
int isogs_adam_fused(const isogs_adam_group* groups /*HOST*/, int n_groups, float beta1, float beta2,
                     float eps, int step, const isogs_adam_state* state, const int32_t* skip_flag, void* stream);

/* First launch of a fused step: small[0..n_small) = 0 (losses, pose gradient and accumulators), and, if state
 * is non-null, state->step += 1 with the bias corrections of that step (double precision, like torch). */
int isogs_step_begin(float* small, int n_small, isogs_adam_state* state, float beta1, float beta2, void* stream);

/* ---- gradient-driven densification (off in every reference config, part of the operator surface) ---- */

/* accumulate_mean2d_gradient + max_2D_radius (python_ordinary/utils/slam_external.py:100-104,
 * splatam.py:751-753): for every Gaussian with radius > 0: grad_accum += |(scale_x v_x, scale_y v_y)|,
 * denom += 1, max_radius = max(max_radius, radius).  v_xy: DEVICE rows of `v_stride` floats whose first two
 * are dL/d(mean2D) in pixels (v_rec of isogs_blend_bwd: stride ISOGS_REC); scale (1,1) keeps pixel units. */
int isogs_densify_stats(int64_t N, const int32_t* radii, const float* v_xy, int v_stride, float scale_x,
                        float scale_y, float* grad_accum, float* denom, float* max_radius, void* stream);

/* densify (slam_external.py:191-243; GaussianModel::densify, cpp/src/core/gaussian_model.cpp:157-289;
 * clone_gaussian_kernel / split_gaussian_kernel, cpp/cuda/kernels/densify.cu:10-92) in two calls so the host
 * can size the pool in between:
 *   select: grads = accum/denom (nan -> 0); clone = grads >= grad_thresh && max exp(s) <= small_scale;
 *           split = grads >= grad_thresh && max exp(s) > small_scale  (small_scale = 0.01 * scene_radius).
 *           scratch: DEVICE int32[4N] (flags + inclusive scans, consumed by emit);
 *           counts: DEVICE int32[3] = {N + n_clone + n_split_into * n_split, n_clone, n_split}.
 *   emit:   appends the clones in index order, then n_split_into blocks of the split rows (the reference's
 *           `.repeat(n, 1)` order) with mean + R(q) (exp(s) * z), log_scale = log(exp(s) / (0.8 n)); z ~ N(0,1)
 *           from `noise` (DEVICE [n_split_into * n_split, 3], the reference's torch.normal draw divided by std)
 *           or, if null, from a counter-based generator keyed by (seed, parent, copy, axis).
 *           keep: DEVICE uint8[>= counts[0]]: 0 for split parents, 1 for everything else (feed isogs_compact).
 *           The caller zeroes the Adam moments of the appended rows. */
int isogs_densify_select(int64_t N, const float* grad_accum, const float* denom, const float* log_scales,
                         int scale_dim, float grad_thresh, float small_scale, int n_split_into,
                         int32_t* scratch, void* scan_temp, size_t scan_temp_bytes, int32_t* counts, void* stream);
int isogs_densify_emit(int64_t N, int64_t capacity, const int32_t* scratch, float* means, float* rgb,
                       float* quats /*nullable*/, float* logit_opac, float* log_scales, int scale_dim,
                       int n_split_into, const float* noise /*nullable*/, uint64_t seed, uint8_t* keep, void* stream);

/* dst[i] = dst[i] && src[i] (combine the split-parent mask with isogs_prune_mask before one compaction). */
int isogs_mask_and(int64_t N, uint8_t* dst, const uint8_t* src, void* stream);

/* keyframe selection by re-projection overlap: replaces keyframe_selection_overlap + get_pointcloud
 * (python_ordinary/utils/keyframe_selection.py:9-95; cpp/src/utils/keyframe_selection.cpp) up to the final
 * host-side ranking.  Samples n_samples valid-depth pixels of gt_depth (sample_idx: DEVICE int32[n] flat pixel
 * indices = the reference's torch.randint draw, or null: drawn on the device from `seed`), back-projects them
 * with the inverse of cur_w2c (HOST float[12], row-major 3x4 [R|t]), removes points that coincide with another
 * sample or with the origin after |round(.,4)| (the reference's unique/isin filter), and writes for each
 * keyframe (kf_w2c: DEVICE [K,12]) the fraction projecting inside the image minus `edge` pixels with z > 0.
 * scratch_i: DEVICE int32[H*W + n]; scratch_f: DEVICE float[6n]; percent_inside: DEVICE float[K].  No sync.
 * The caller ranks (descending), keeps > 0 and draws k of them (keyframe_selection.py:86-93). */
int isogs_keyframe_overlap(const isogs_camera* cam, const float* gt_depth, const float* cur_w2c,
                           const int32_t* sample_idx, int n_samples, uint64_t seed, const float* kf_w2c,
                           int n_keyframes, int edge, int32_t* scratch_i, float* scratch_f, void* scan_temp,
                           size_t scan_temp_bytes, float* percent_inside, void* stream);

/* ---- IsoGS geometric losses: flat + iso-surface density, three-axis scales only ---- */

/* Exact 16-nearest-neighbour search among the N Gaussian centres for Q query centres: replaces the chunked
 * torch.cdist + topk of compute_iso_surface_loss_sampled (python_ordinary/scripts/splatam.py:203-208) and the
 * spatial hash of cpp/cuda/kernels/spatial_hash.cu.  A uniform grid (at most res+1 cells per axis over the
 * bounding box of the centres) is rebuilt on the device on every call; no host sync.
 * query_idx: DEVICE int32[Q] (nullable).  If null and sampled_idx (DEVICE int32[Q]) is given, Q centres are
 * drawn from `seed` (with replacement; the reference uses randperm) and written there; if both are null the
 * queries are centres 0..Q-1 (Q must be N: the reference's N < 8192 case, :654-656).
 * grid / sorted: DEVICE scratch of isogs_knn_workspace bytes (sorted 16-byte aligned); scan_temp must hold a
 * CUB scan of (res+1)^3 + 1 ints (isogs_temp_bytes with that N).  nbr: DEVICE int32[Q,16], nearest first,
 * -1 padded when N < 16. */
int isogs_knn_workspace(int64_t N, int res, size_t* grid_bytes, size_t* sorted_bytes);
int isogs_knn16(int64_t N, const float* means, int res, int Q, const int32_t* query_idx, uint64_t seed,
                int32_t* sampled_idx, int32_t* grid, float* sorted, void* scan_temp, size_t scan_temp_bytes,
                int32_t* nbr, void* stream);

/* flat = mean_i min_k clamp(exp(log_scales_ik), 1e-5)  (compute_flat_loss, splatam.py:42-54,580-583) and
 * iso = mean_q (sum_{j in nbr_q} sigmoid(o_j) exp(-1/2 d^T R_j (S_j^2 + 1e-8)^-1 R_j^T d) - target)^2
 * (splatam.py:612-668), d = mu_q - mu_j.  losses (nullable): losses[0] += w_flat*flat + w_iso*iso
 * (reference weights 50 / 2, :733-739).  out: DEVICE float[3], += {flat, iso, mean density} (caller zeroes).
 * g_*: gradients of the weighted sum, ACCUMULATED into [N,3] / [N,3] / [N,4] / [N] (all four or none);
 * call after isogs_preprocess_bwd, which overwrites the gradient buffers. */
int isogs_iso_flat_loss(int64_t N, const float* means, const float* log_scales, const float* quats,
                        const float* logit_opac, int Q, const int32_t* query_idx, const int32_t* nbr,
                        float w_iso, float w_flat, float target, float* g_means, float* g_log_scales,
                        float* g_quats, float* g_logit_opac, float* losses, float* out, void* stream);

/* best-candidate selection of the tracking loop (python_ordinary/scripts/splatam.py:1287-1290),
 * on the device so the loop needs no host sync: if losses[0] (evaluated BEFORE this iteration's
 * Adam step) < best[7], then best[0..6] = pose (the values AFTER the step) and best[7] = losses[0].
 * best: DEVICE float[8] = {q[4], t[3], min_loss}; initialise best[7] to 1e20 (splatam.py:1242). */
int isogs_pose_keep_best(const float* losses, const float* pose, float* best, const int32_t* skip_flag /*nullable*/,
                         void* stream);

/* silhouette-driven prune compaction: replaces prune_gaussians + remove_points
 * (python_ordinary/utils/slam_external.py:139-188) and GaussianModel::prune
 * (cpp/src/core/gaussian_model.cpp:291-331).  keep[i] = !(sigmoid(logit_o) < opac_thres ||
 * max exp(log_s) > max_scale).  Order-preserving gather of n_rows row-major tensors.
 * new_count: DEVICE int32. */
typedef struct isogs_rows { const float* in; float* out; int32_t width; } isogs_rows;
int isogs_prune_mask(int64_t N, const float* logit_opac, const float* log_scales, int scale_dim,
                     float opac_thres, float max_scale, uint8_t* keep, void* stream);
int isogs_compact(int64_t N, const uint8_t* keep, const isogs_rows* rows /*HOST*/, int n_rows,
                  int32_t* scan /*[N] scratch*/, void* scan_temp, size_t scan_temp_bytes,
                  int32_t* new_count, void* stream);

/* silhouette-driven densification: replaces add_new_gaussians + get_pointcloud +
 * initialize_new_params (splatam.py:799-841,298-348,763-796) and Mapper::addNewGaussians
 * (cpp/src/slam/mapper.cpp:175-206).  Selects pixels with (alpha < sil_thres ||
 * (render_depth > gt_d && |gt_d - render_depth| > depth_err_thres)) && gt_d > 0,
 * back-projects them with the inverse of (pose, view) and appends, in pixel order, at
 * offset *count (DEVICE int32, updated) up to capacity.  Appended rows:
 * means[3], rgb[3], quat=[1,0,0,0], logit_opac=0, log_scale=log(z/((fx+fy)/2)) x scale_dim
 * (+ noise_std * N(0,1) per axis from a counter-based generator keyed by (seed, pixel, axis)
 * when scale_dim == 3: the IsoGS symmetry-breaking jitter of splatam.py:771-777). */
int isogs_backproject_append(const isogs_camera* cam, const isogs_pose* pose,
                             const float* gt_rgb, const float* gt_depth,
                             const float* render_depth, const float* alpha,
                             float sil_thres, float depth_err_thres,
                             float* means, float* rgb, float* quats, float* logit_opac,
                             float* log_scales, int scale_dim, float noise_std, uint64_t seed,
                             int64_t capacity,
                             int32_t* scan /*[2*H*W] scratch*/, void* scan_temp, size_t scan_temp_bytes,
                             int32_t* count, void* stream);

/* ---- the fused steps: ONE host call enqueues a whole iteration ------------------------- */

/* Parameters, gradients and Adam moments of the map in the reference's order (splatam.py:351-399):
 * 0 means3D [N,3], 1 rgb_colors [N,3], 2 unnorm_rotations [N,4], 3 logit_opacities [N,1], 4 log_scales [N,sd].
 * g / m / v entries may be null where no gradient is wanted (all of them for tracking; index 2 when sd == 1). */
typedef struct isogs_gaussians {
    int64_t N;
    int32_t scale_dim, reserved;
    float* p[5];
    float* g[5];
    float* m[5];
    float* v[5];
} isogs_gaussians;

/* Caller-allocated DEVICE scratch of a step (nothing is allocated or freed by the library). */
typedef struct isogs_workspace {
    float* rec;            /* [n_cap,12] */
    float* v_rec;          /* [n_cap,12] */
    int32_t* radii;        /* [n_cap,2] */
    uint64_t* bucket;      /* [isect_cap] */
    int32_t* vals;         /* [isect_cap] */
    int32_t* ranges;       /* [n_tiles+1] */
    int32_t* totals;       /* [2] {n_isects, sticky overflow} */
    void* sort_temp;
    size_t sort_temp_bytes;
    int64_t isect_cap;
    float* render;         /* [5,H,W] */
    float* v_render;       /* [5,H,W] */
    float* alpha;          /* [H,W] */
    int32_t* last;         /* [H,W] */
    float* small;          /* [32]: losses[4] | v_pose[7] | pose_acc[16] | ssim | outlier thr | geo[3] */
    isogs_adam_state* adam_state;
    float* dmaps;          /* [9,H,W] SSIM derivative maps (nullable unless ISOGS_STEP_SSIM) */
    float* median_scratch; /* [2,H,W]  (nullable unless ISOGS_STEP_OUTLIER) */
    void* median_temp;
    size_t median_temp_bytes;
    uint32_t* tile_rects;  /* [n_cap,4] u32 tile rect + depth bits per Gaussian (nullable: slower binning) */
} isogs_workspace;

#define ISOGS_STEP_SSIM 1     /* colour term 0.8*L1 + 0.2*(1-SSIM) (splatam.py:568) instead of pure L1 */
#define ISOGS_STEP_OUTLIER 2  /* ignore_outlier_depth_loss (splatam.py:545-554); mapping: mean over the masked pixels */
#define ISOGS_STEP_NO_SIL 4   /* tracking: use_sil_for_loss = False */
#define ISOGS_PHASE_RENDER 1  /* begin + forward + loss + backward (gradients land in gaussians->g) */
#define ISOGS_PHASE_ADAM 2    /* fused Adam on every group with a gradient */
#define ISOGS_STEP_EVENTS 9   /* begin, preprocess_fwd, tile_bin, blend_fwd, loss, ssim, blend_bwd, preprocess_bwd, adam */

typedef struct isogs_step_params {
    int32_t flags, phases;
    float w_im, w_depth, sil_thres;
    float lr[5];            /* mapping: per parameter group; tracking: lr[0] = rotation, lr[1] = translation */
    float beta1, beta2, eps;
    float lr_pose[2];       /* isogs_map_step_ba: rotation, translation */
} isogs_step_params;

/* One mapping iteration on one keyframe == get_loss(mapping=True) + loss.backward() + optimizer.step()
 * (python_ordinary/scripts/splatam.py:1412-1476; transform_to_frame + both renders + masked L1 (+ SSIM) +
 * autograd + torch.optim.Adam; Mapper::optimize, cpp/src/slam/mapper.cpp:34-151).  frame: DEVICE slot (pose,
 * ground truth, 1/n_valid_depth).  No host sync, no allocation: capturable in a CUDA graph.  Split the phases to
 * put a gradient exchange (or the iso / flat terms) between backward and Adam.  events (nullable): HOST array of
 * ISOGS_STEP_EVENTS cudaEvent_t recorded after the stages for per-stage timing. */
int isogs_map_step(const isogs_camera* cam, const isogs_frame* frame /*DEVICE*/, const isogs_workspace* ws,
                   const isogs_gaussians* gs, const isogs_step_params* prm, void* const* events, void* stream);

/* get_loss(mapping=True, do_ba=True) (splatam.py:505-510): the same iteration with the keyframe's pose a parameter
 * too -- the transform is differentiated with respect to both, the pose takes its own Adam step (prm->lr_pose,
 * pose_state).  pose / pose_m / pose_v: DEVICE float[7]; frame->pose is ignored.  Both phases in one call. */
int isogs_map_step_ba(const isogs_camera* cam, const isogs_frame* frame /*DEVICE*/, const isogs_workspace* ws,
                      const isogs_gaussians* gs, float* pose, float* pose_m, float* pose_v,
                      isogs_adam_state* pose_state, const isogs_step_params* prm, void* const* events, void* stream);

/* One tracking iteration == get_loss(tracking=True) + backward + Adam on the pose + best-candidate rule
 * (splatam.py:1247-1290; Tracker::optimizePose, cpp/src/slam/tracker.cpp:15-115).  pose / pose_m / pose_v:
 * DEVICE float[7] (cam_unnorm_rots | cam_trans of the frame and their Adam moments), best: DEVICE float[8],
 * pose_state: the pose optimizer's step state.  frame->pose is ignored. */
int isogs_track_step(const isogs_camera* cam, const isogs_frame* frame /*DEVICE*/, const isogs_workspace* ws,
                     const isogs_gaussians* gs, float* pose, float* pose_m, float* pose_v, float* best,
                     isogs_adam_state* pose_state, const isogs_step_params* prm, void* const* events, void* stream);

/* uint8 HWC colour + uint16 depth (the reference's source formats, python_ordinary/datasets/gradslam_datasets/
 * basedataset.py:299-340: color / 255 -> [3,H,W], depth / png_depth_scale) unpacked into the fp32 planes the
 * step reads: 5 bytes per pixel cross PCIe instead of 16. */
int isogs_unpack_rgbd(const isogs_camera* cam, const uint8_t* rgb_hwc, const uint16_t* depth_u16,
                      float inv_depth_scale, float* gt_rgb, float* gt_depth, void* stream);

/* CUDA-graph capture of anything enqueued on `stream` between begin and end (thread-local capture mode);
 * launch replays it.  graph_exec is an opaque handle. */
int isogs_graph_begin(void* stream);
int isogs_graph_end(void* stream, void** graph_exec);
int isogs_graph_launch(void* graph_exec, void* stream);
int isogs_graph_destroy(void* graph_exec);
/* Kernel nodes / other nodes (memsets, copies) of a captured graph, counted from the graph at isogs_graph_end. */
int isogs_graph_info(void* graph_exec, int32_t* n_kernel_nodes, int32_t* n_other_nodes);

/* keyframe-sharded mapping, SURVEY 8(e): the average of the flat gradient buffer over the ranks as ONE kernel over
 * NVLink peer memory (two-shot: every rank reduces the 1/world slice it owns from all ranks in rank order and stores the
 * result into every rank's buffer), instead of an NCCL all-reduce.  The reference has no distributed code; the
 * sequential analogue is one keyframe per iteration (python_ordinary/scripts/splatam.py:1415-1426).
 * isogs_peer_alloc: cudaMalloc'd, zeroed, IPC-exportable memory (handle: 64 bytes, HOST); isogs_peer_open maps a
 * peer's allocation into this process (peer access is enabled lazily).  Layout chosen by the caller: the float
 * gradient buffer plus a block of 64 uint32 flags (zero at start), both inside such allocations.
 * isogs_allreduce_avg_p2p: bufs / flags are HOST arrays of `world` DEVICE pointers (index = rank; own entry = own
 * memory); numel a multiple of 4, buffers 16-byte aligned; epoch must increase by one per call, identically on
 * every rank; n_ctas <= 0 picks the grid.  In-place: afterwards every buffer holds the same bits.
 * isogs_peer_wait (same stream, before the first consumer of the reduced gradients) waits until every rank's slice
 * has landed locally.  status: DEVICE int32, set non-zero if a spin timed out after 2 s (results then undefined). */
int isogs_peer_alloc(size_t bytes, void** ptr, unsigned char* handle);
int isogs_peer_open(const unsigned char* handle, void** ptr);
int isogs_peer_close(void* ptr);
int isogs_peer_free(void* ptr);
int isogs_allreduce_avg_p2p(int world, int rank, float* const* bufs /*HOST*/, uint32_t* const* flags /*HOST*/,
                            int64_t numel, uint32_t epoch, int n_ctas, int32_t* status, void* stream);
int isogs_peer_wait(int world, const uint32_t* flags_local, uint32_t epoch, int32_t* status, void* stream);

/* The exchange FUSED with the optimizer step: the owner of a slice averages it over the ranks (as above), applies Adam to
 * it in registers -- its slice of the moments is the only up-to-date copy, i.e. the optimizer state is sharded over the
 * ranks -- and stores the updated PARAMETERS into every rank's parameter tensors; isogs_adam_fused is not called at
 * all.  Replaces, per rank, the all-reduce above + torch.optim.Adam.step over all N Gaussians (splatam.py:1475) by one
 * kernel whose Adam work is N / world.  groups (HOST): position of each parameter group in the flat gradient buffer
 * (`offset`, `numel` in floats; offset a multiple of 4), the group's parameter tensor on EVERY rank (peer-mapped, own
 * entry = local, 16-byte aligned, at least 3 floats of slack behind numel), the LOCAL moments and the learning rate.
 * state: DEVICE isogs_adam_state advanced by isogs_step_begin this iteration.  Flags / epoch / status / isogs_peer_wait
 * as for isogs_allreduce_avg_p2p (wait before the next reader of the parameters). */
#define ISOGS_PEER_MAX_WORLD 16
typedef struct isogs_peer_adam_group {
    int64_t offset, numel;
    float* param[ISOGS_PEER_MAX_WORLD];
    float* exp_avg;
    float* exp_avg_sq;
    float lr;
} isogs_peer_adam_group;
int isogs_reduce_adam_p2p(int world, int rank, float* const* grad_bufs /*HOST*/, uint32_t* const* flags /*HOST*/,
                          int64_t numel, const isogs_peer_adam_group* groups /*HOST*/, int n_groups, float beta1,
                          float beta2, float eps, const isogs_adam_state* state, uint32_t epoch, int n_ctas,
                          int32_t* status, void* stream);

/* The same fused exchange through the NVSwitch's in-fabric reduction (NVLS): mc_grad is the MULTICAST address of the
 * flat gradient buffer (every rank's buffer bound to one multicast object), mc_params[g] the multicast address of
 * parameter group g.  The slice owner reads its slice with ONE multimem.ld_reduce.add.v4.f32 per 16 bytes (the switch
 * sums the ranks' copies: 1/world of the buffer crosses the owner's link instead of (world-1)/world) and broadcasts
 * the updated parameters with ONE multimem.st (instead of world peer stores).  The summation order inside the switch
 * is the hardware's, so the result agrees with isogs_reduce_adam_p2p to rounding, not bit for bit; replicas stay
 * bit-identical by construction.  No reference counterpart (the reference has no distributed code). */
int isogs_reduce_adam_nvls(int world, int rank, float* const* grad_bufs /*HOST*/, uint32_t* const* flags /*HOST*/,
                           int64_t numel, const isogs_peer_adam_group* groups /*HOST*/, int n_groups, float beta1,
                           float beta2, float eps, const isogs_adam_state* state, uint32_t epoch, int n_ctas,
                           int32_t* status, const float* mc_grad, float* const* mc_params /*HOST*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif
