/*
 * isogs_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain C++17/OpenMP restatement, in IEEE fp32 with no fast-math and no FMA
 * contraction (-ffp-contract=off; fmaf only where written), of the one hot
 * path this repository accelerates: project -> tile keys -> sort -> tile
 * ranges -> alpha-blend forward -> blend backward -> projection backward, plus
 * the masked-L1 loss, Adam and compaction around it.
 *
 * It follows the arithmetic of the only rasterizer source present in the
 * reference tree (vendored gsplat 1.5.3):
 *   projection fwd  cpp/third_party/gsplat/gsplat/cuda/csrc/ProjectionEWA3DGSFused.cu:56-208
 *   projection bwd  .../ProjectionEWA3DGSFused.cu:346-530 + include/Utils.cuh:18-80,142-261,373-388,498-616
 *   keys / ranges   .../IntersectTile.cu:55-113,227-256
 *   blend fwd       .../RasterizeToPixels3DGSFwd.cu:62-63,85-92,140-187
 *   blend bwd       .../RasterizeToPixels3DGSBwd.cu:107-112,156-243
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (the CUDA
 * library behind include/isogs_rast.h) never links or calls it.
 *
 * Parity status: projection / keys / ranges are pinned against the
 * reference's own CPU-runnable transcription (gsplat/cuda/_torch_impl.py,
 * see oracle/pin_against_gsplat.py and tests/golden/).  The blend has no
 * CPU-runnable reference implementation in the tree: this restatement of it is
 * verified by finite differences, and the CUDA kernels it checks are
 * additionally compared on the GPU with the reference's own CUDA kernels
 * (oracle/build_ref.sh -> oracle/_ref/gsplat_ref.so, tests/test_gpu_vs_gsplat.py).
 */
#pragma once
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_REC 12 /* floats per projected record / per gradient record */

typedef struct orc_camera {
    int32_t width, height;
    float fx, fy, cx, cy;
    float R[9]; /* row-major world->camera rotation (the rasterizer's view matrix) */
    float t[3];
    float near_plane, far_plane, eps2d;
    float bg[3];
} orc_camera;

/* optional per-frame pose applied BEFORE the view matrix (SplaTAM's
 * transform_to_frame, python_ordinary/utils/slam_helpers.py:252-304):
 * p_frame = Rot(normalize(q)) p + t */
typedef struct orc_pose {
    float q[4]; /* w,x,y,z un-normalised */
    float t[3];
} orc_pose;

/* flags */
#define ORC_F_LOG_SCALES 1   /* scales are log-scales: s = exp(.)      (slam_helpers.py:136) */
#define ORC_F_LOGIT_OPAC 2   /* opacities are logits: o = sigmoid(.)   (slam_helpers.py:135) */
#define ORC_F_NO_OPAC_EXTENT 4 /* extent = 3.33 (gsplat _torch_impl.py:356), for pinning */

/* deterministic elementary functions shared (by specification, not by code)
 * with the CUDA kernels -- see DESIGN.md "bit-exact keys". */
float orc_logf(float x);
float orc_expf(float x);

/* preprocess (north_star (1)).  scale_dim 1 = isotropic [N,1], 3 = [N,3]+quats.
 * ch = 4: record colours {c0,c1,c2,z}; ch = 5: {c0,c1,c2,z,z*z}.
 * rec [N,12] = x,y,0.5*conic_a,conic_b,0.5*conic_c,L,opacity,c0..c4
 * radii [N,2]; tiles [N] = number of 16x16 tiles touched. */
void orc_preprocess_fwd(const orc_camera* cam, const orc_pose* pose /*nullable*/, int64_t N,
                        const float* means, const float* scales, int scale_dim,
                        const float* quats /*nullable when scale_dim==1*/,
                        const float* opacities, const float* colors /*[N,3]*/,
                        int ch, int flags,
                        float* rec, int32_t* radii, int32_t* tiles);

/* duplicateWithKeys + stable radix sort + identifyTileRanges (north_star (2)).
 * Returns M; keys/vals may be null to only count.  tile_ranges has n_tiles+1 entries. */
int64_t orc_tile_bin(const orc_camera* cam, int64_t N, const float* rec, const int32_t* radii,
                     const int32_t* tiles, int64_t cap,
                     uint64_t* keys_unsorted /*nullable*/, int32_t* vals_unsorted /*nullable*/,
                     uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges);

/* renderCUDA forward (north_star (3)).  render [ch,H,W] planar, alpha [H,W] = 1-T, last [H,W]. */
void orc_blend_fwd(const orc_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                   const int32_t* tile_ranges, float* render, float* alpha, int32_t* last);

/* renderCUDA backward (north_star (4)).  v_rec [N,12] = v_x,v_y,v_a,v_b,v_c,v_opac,v_c0..v_c4,0 */
void orc_blend_bwd(const orc_camera* cam, int ch, int64_t N, const float* rec,
                   const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                   const int32_t* last, const float* v_render /*[ch,H,W]*/,
                   const float* v_alpha /*[H,W] nullable*/, float* v_rec);

/* preprocess backward: chains v_rec into the inputs of orc_preprocess_fwd.
 * v_means [N,3] (gradient w.r.t. the `means` argument, i.e. through the pose when
 * one is given), v_scales [N,scale_dim], v_quats [N,4] (nullable), v_opac [N],
 * v_colors [N,3], v_pose[7] = d/dq(4) (un-normalised), d/dt(3) (nullable). */
void orc_preprocess_bwd(const orc_camera* cam, const orc_pose* pose, int64_t N,
                        const float* means, const float* scales, int scale_dim,
                        const float* quats, const float* opacities, const float* colors,
                        int ch, int flags, const int32_t* radii, const float* rec,
                        const float* v_rec,
                        float* v_means, float* v_scales, float* v_quats, float* v_opac,
                        float* v_colors, double* v_pose);

/* masked L1 colour/depth loss on the 5-channel render (splatam.py:534-570,747-749).
 * mode 0 = tracking (sum over mask = gt_d>0 & !nan & sil>sil_thres), 1 = mapping (mean
 * over gt_d>0 for depth, plain mean for colour).  Writes v_render [5,H,W] and
 * losses[3] = {weighted total, colour term, depth term}. */
float orc_depth_outlier_threshold(int64_t P, const float* render_depth, const float* gt_depth, float factor);

void orc_loss_l1(const orc_camera* cam, int mode, const float* render /*[5,H,W]*/,
                 const float* alpha, const float* gt_rgb /*[3,H,W]*/, const float* gt_depth /*[H,W]*/,
                 float w_im, float w_depth, float sil_thres, int use_sil,
                 float outlier_thresh /* < 0: off; splatam.py:551-554 */,
                 float* v_render, double* losses);

/* torch.optim.Adam semantics (bias-corrected, step counted from 1; splatam.py:402-408). */
/* SSIM term (python_ordinary/utils/slam_external.py:54-97): returns mean SSIM over [3,H,W]; if v_x is
 * non-null, v_x[3,H,W] += coef * d(mean SSIM)/dx  (x = first image).  Double accumulation. */
double orc_ssim(int width, int height, const float* x /*[3,H,W]*/, const float* y /*[3,H,W]*/, float coef,
                float* v_x /*nullable*/);

/* IsoGS flat + iso-surface density losses (splatam.py:42-54,162-262,572-675); see the .cpp for the math. */
double orc_iso_flat_loss(int64_t N, const float* means, const float* log_scales, const float* quats,
                         const float* logit_opac, int Q, const int32_t* query_idx /*nullable: 0..Q-1*/, float w_iso,
                         float w_flat, float target, float* g_means, float* g_ls, float* g_quats, float* g_logit,
                         double* out /*[3]*/, int32_t* nbr_out /*[Q,16] nullable*/);

void orc_adam(int64_t n, float* p, const float* g, float* m, float* v, float lr, float b1,
              float b2, float eps, int step);

/* order-preserving compaction of row-major [N,width] float rows by keep mask
 * (slam_external.py:139-160).  Returns kept count. */
int64_t orc_compact_rows(int64_t N, int width, const uint8_t* keep, const float* in, float* out);

int orc_num_threads(void);
/* launchers such as torchrun export OMP_NUM_THREADS=1: the timed CPU baseline asks for the host's cores explicitly */
void orc_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
