// step.cu -- north_star (5): fused masked L1 colour/depth loss, fused multi-tensor Adam, and the
// silhouette-driven densify / prune compaction.  All HBM-streaming kernels.
// References: python_ordinary/scripts/splatam.py:534-570,747-749 (loss), :402-408 (Adam),
// :799-841,298-348,763-796 (densify); python_ordinary/utils/slam_external.py:139-188 (prune);
// cpp/cuda/kernels/optimizer.cu:15-98; cpp/src/core/gaussian_model.cpp:291-331.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>
#include <mutex>

#include "common.cuh"

namespace isogs {

// ------------------------------------------------------------------------------------------
// loss
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float sgnf(float d) { return (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f); }

// V pixels per thread per trip: V = 4 moves every plane as float4 (P % 4 == 0 and 16-byte aligned planes)
template <int V> struct Px { float v[V]; };
template <int V> __device__ __forceinline__ Px<V> ldv(const float* p) {
    Px<V> r;
    if constexpr (V == 4) { const float4 t = *reinterpret_cast<const float4*>(p); r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w; }
    else r.v[0] = *p;
    return r;
}
template <int V> __device__ __forceinline__ void stv(float* p, const Px<V>& r) {
    if constexpr (V == 4) *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
    else *p = r.v[0];
}

template <int V>
__global__ void __launch_bounds__(256)
loss_l1_kernel(int64_t P, int mode, const float* __restrict__ render, const float* __restrict__ alpha,
               const float* __restrict__ gt_rgb, const float* __restrict__ gt_depth, float w_im, float w_depth,
               float sil_thres, int use_sil, float inv_nd, float inv_nc, const float* __restrict__ outlier_thresh,
               float* __restrict__ v_render, float* __restrict__ losses, const isogs_frame* __restrict__ fr,
               const float* __restrict__ inv_n_dev) {
    if (fr != nullptr) {   // current frame through the device-side frame slot (CUDA-graph replay)
        gt_rgb = fr->gt_rgb;
        gt_depth = fr->gt_depth;
        if (mode == 1) inv_nd = fr->inv_n_depth;
        if (V == 4 && ((((uintptr_t)gt_rgb | (uintptr_t)gt_depth) & 15) != 0)) { // documented requirement
            if (blockIdx.x == 0 && threadIdx.x == 0) losses[0] = __int_as_float(0x7fc00000);
            return;
        }
    }
    if (inv_n_dev != nullptr) inv_nd = *inv_n_dev;   // mapping + outlier mask: 1 / n(mask), computed on the device
    float sum_d = 0.f, sum_c = 0.f, cnt = 0.f;
    const bool outl = outlier_thresh != nullptr;
    const float othr = outl ? *outlier_thresh : 0.f;
    const bool need_alpha = (mode == 0) && use_sil;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V; i < P; i += (int64_t)gridDim.x * blockDim.x * V) {
        const Px<V> d = ldv<V>(render + 3 * P + i), d2 = ldv<V>(render + 4 * P + i), gd = ldv<V>(gt_depth + i);
        Px<V> al;
        if (need_alpha) al = ldv<V>(alpha + i);
        Px<V> c[3], g[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { c[k] = ldv<V>(render + k * P + i); g[k] = ldv<V>(gt_rgb + k * P + i); }
        Px<V> od, oc[3], zero;
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const float unc = d2.v[j] - d.v[j] * d.v[j];
            const bool nan_ok = !(isnan(d.v[j]) || isnan(unc));
            bool m = (gd.v[j] > 0.f) && nan_ok;
            bool mc = true;
            const float dd = d.v[j] - gd.v[j];
            if (outl) m = m && (fabsf(dd) < othr);               // ignore_outlier_depth_loss, splatam.py:551-554
            if (mode == 0) {
                if (use_sil) m = m && (al.v[j] > sil_thres);
                mc = (use_sil || outl) ? m : true;               // :563-568
            }
            if (gd.v[j] > 0.f) cnt += 1.f;
            if (m) sum_d += fabsf(dd);
            od.v[j] = m ? (w_depth * inv_nd) * sgnf(dd) : 0.f;
            zero.v[j] = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float dc = c[k].v[j] - g[k].v[j];
                if (mc) sum_c += fabsf(dc);
                oc[k].v[j] = mc ? (w_im * inv_nc) * sgnf(dc) : 0.f;
            }
        }
        stv<V>(v_render + 3 * P + i, od);
        stv<V>(v_render + 4 * P + i, zero);
#pragma unroll
        for (int k = 0; k < 3; ++k) stv<V>(v_render + k * P + i, oc[k]);
    }
    __shared__ float sm[3][8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        sum_d += __shfl_xor_sync(0xffffffffu, sum_d, off);
        sum_c += __shfl_xor_sync(0xffffffffu, sum_c, off);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    }
    if (lane == 0) { sm[0][wid] = sum_d; sm[1][wid] = sum_c; sm[2][wid] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float sd = 0.f, sc = 0.f, n = 0.f;
        for (int w = 0; w < 8; ++w) { sd += sm[0][w]; sc += sm[1][w]; n += sm[2][w]; }
        const float ld = sd * inv_nd, lc = sc * inv_nc;
        atomicAdd(losses + 0, w_im * lc + w_depth * ld);
        atomicAdd(losses + 1, lc);
        atomicAdd(losses + 2, ld);
        atomicAdd(losses + 3, n);
    }
}

// ------------------------------------------------------------------------------------------
// Adam (torch.optim.Adam semantics: bias-corrected, step counted from 1)
// ------------------------------------------------------------------------------------------
constexpr int MAX_GROUPS = 8;
struct AdamArgs {
    float* p[MAX_GROUPS]; const float* g[MAX_GROUPS]; float* m[MAX_GROUPS]; float* v[MAX_GROUPS];
    int64_t numel[MAX_GROUPS];
    int64_t block_start[MAX_GROUPS + 1]; // in units of blocks
    float step_size[MAX_GROUPS];     // host-step mode: lr / (1 - b1^t); device-step mode: lr
    int n;
    float b1, b2, one_minus_b1, one_minus_b2, inv_bc2_sqrt, eps;
    const isogs_adam_state* state;   // nullable: bias corrections written by step_begin_kernel (graph-replayable)
    const int32_t* skip;             // nullable: *skip != 0 turns the launch into a no-op (overflowed forward)
};

constexpr int ADAM_BLOCK = 256, ADAM_VEC = 4;

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamArgs& a, float step_size,
                                         float inv_bc2_sqrt) {
    adam_update(p, g, m, v, a.one_minus_b1, a.b2, a.one_minus_b2, inv_bc2_sqrt, a.eps, step_size);
}

__global__ void __launch_bounds__(ADAM_BLOCK) adam_kernel(AdamArgs a) {
    if (a.skip != nullptr && *a.skip != 0) return;
    int grp = 0;
#pragma unroll
    for (int k = 1; k < MAX_GROUPS; ++k)
        if (k < a.n && (int64_t)blockIdx.x >= a.block_start[k]) grp = k;
    const int64_t local_block = (int64_t)blockIdx.x - a.block_start[grp];
    const int64_t base = (local_block * ADAM_BLOCK + threadIdx.x) * ADAM_VEC;
    const int64_t n = a.numel[grp];
    if (base >= n) return;
    float* p = a.p[grp]; const float* g = a.g[grp]; float* m = a.m[grp]; float* v = a.v[grp];
    float ss = a.step_size[grp], ib2 = a.inv_bc2_sqrt;
    if (a.state != nullptr) {   // the same double -> float roundings as the host-step path
        ss = (float)((double)ss * a.state->inv_bc1);
        ib2 = (float)a.state->inv_bc2_sqrt;
    }
    const bool vec_ok = (base + ADAM_VEC <= n) &&
                        ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
    if (vec_ok) {
        float4 P4 = *reinterpret_cast<float4*>(p + base);
        const float4 G4 = *reinterpret_cast<const float4*>(g + base);
        float4 M4 = *reinterpret_cast<float4*>(m + base);
        float4 V4 = *reinterpret_cast<float4*>(v + base);
        adam_one(P4.x, G4.x, M4.x, V4.x, a, ss, ib2);
        adam_one(P4.y, G4.y, M4.y, V4.y, a, ss, ib2);
        adam_one(P4.z, G4.z, M4.z, V4.z, a, ss, ib2);
        adam_one(P4.w, G4.w, M4.w, V4.w, a, ss, ib2);
        *reinterpret_cast<float4*>(p + base) = P4;
        *reinterpret_cast<float4*>(m + base) = M4;
        *reinterpret_cast<float4*>(v + base) = V4;
    } else {
        for (int64_t i = base; i < n && i < base + ADAM_VEC; ++i) adam_one(p[i], g[i], m[i], v[i], a, ss, ib2);
    }
}

// ------------------------------------------------------------------------------------------
// prune mask + order-preserving compaction
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
prune_mask_kernel(int64_t N, const float* __restrict__ logit_opac, const float* __restrict__ log_scales, int sd,
                  float opac_thres, float max_scale, uint8_t* __restrict__ keep) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    bool rm = det_sigmoidf(logit_opac[i]) < opac_thres; // slam_external.py:175
    float smax = det_expf(log_scales[sd * i]);
    for (int k = 1; k < sd; ++k) smax = fmaxf(smax, det_expf(log_scales[sd * i + k]));
    rm = rm || (smax > max_scale); // :177-179
    keep[i] = rm ? 0 : 1;
}

constexpr int MAX_ROWS = 24;
struct RowsArgs { const float* in[MAX_ROWS]; float* out[MAX_ROWS]; int width[MAX_ROWS]; int n; };

__global__ void __launch_bounds__(256)
compact_kernel(int64_t N, const uint8_t* __restrict__ keep, const int32_t* __restrict__ scan, RowsArgs rows,
               int32_t* __restrict__ new_count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const bool k = keep[i] != 0;
    const int32_t dst = scan[i] - (k ? 1 : 0); // inclusive scan -> exclusive position
    if (i == N - 1) *new_count = scan[i];
    if (!k) return;
    for (int r = 0; r < rows.n; ++r) {
        const int w = rows.width[r];
        const float* src = rows.in[r] + i * w;
        float* d = rows.out[r] + (int64_t)dst * w;
        for (int c = 0; c < w; ++c) d[c] = src[c];
    }
}

__global__ void __launch_bounds__(256) u8_to_i32_kernel(int64_t N, const uint8_t* __restrict__ in, int32_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) out[i] = in[i] ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// silhouette-driven densification
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
densify_mask_kernel(int64_t P, const float* __restrict__ gt_depth, const float* __restrict__ render_depth,
                    const float* __restrict__ alpha, float sil_thres, float depth_err_thres, int32_t* __restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float gd = gt_depth[i], rd = render_depth[i];
    const float err = fabsf(gd - rd) * ((gd > 0.f) ? 1.f : 0.f); // splatam.py:811
    const bool np = (alpha[i] < sil_thres) || ((rd > gd) && (err > depth_err_thres)); // :807,812,814
    flag[i] = (np && gd > 0.f) ? 1 : 0;                                                // :826-827
}

__device__ __forceinline__ uint32_t hash_u32(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return (uint32_t)x;
}
__device__ __forceinline__ float counter_normal(uint64_t seed, uint64_t idx, int axis) {
    const uint64_t c = seed * 0x9E3779B97F4A7C15ull + idx * 6ull + (uint64_t)axis * 2ull;
    const float u1 = ((float)(hash_u32(c) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float u2 = ((float)(hash_u32(c + 1) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

__global__ void __launch_bounds__(256)
densify_append_kernel(DevCam cam, DevPose pose, const float* __restrict__ gt_rgb, const float* __restrict__ gt_depth,
                      const int32_t* __restrict__ flag, const int32_t* __restrict__ scan, float* __restrict__ means,
                      float* __restrict__ rgb, float* __restrict__ quats, float* __restrict__ logit_opac,
                      float* __restrict__ log_scales, int sd, float noise_std, uint64_t seed, int64_t capacity,
                      const int32_t* __restrict__ count_in) {
    const int64_t P = (int64_t)cam.width * cam.height;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int32_t base = *count_in;
    if (!flag[i]) return;
    const int64_t dst = (int64_t)base + scan[i] - 1;
    if (dst >= capacity) return;
    const int px = (int)(i % cam.width), py = (int)(i / cam.width);
    const float z = gt_depth[i];
    // get_pointcloud, splatam.py:306-322 (integer pixel coordinates)
    float pc[3] = {((float)px - cam.cx) / cam.fx * z, ((float)py - cam.cy) / cam.fy * z, z};
    // inverse of view then of pose: p_f = Rv^T (p_c - tv), p_w = Rp^T (p_f - tp)
    float d[3] = {pc[0] - cam.t[0], pc[1] - cam.t[1], pc[2] - cam.t[2]};
    float pf[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) pf[k] = cam.R[k] * d[0] + cam.R[3 + k] * d[1] + cam.R[6 + k] * d[2];
    float pw[3] = {pf[0], pf[1], pf[2]};
    if (pose.mode != 0) {
        float q[4], t[3], Rp[9];
        for (int k = 0; k < 4; ++k) q[k] = (pose.mode == 2) ? pose.dev[k] : pose.q[k];
        for (int k = 0; k < 3; ++k) t[k] = (pose.mode == 2) ? pose.dev[4 + k] : pose.t[k];
        quat_to_R(q, Rp, nullptr, nullptr);
        float e[3] = {pf[0] - t[0], pf[1] - t[1], pf[2] - t[2]};
#pragma unroll
        for (int k = 0; k < 3; ++k) pw[k] = Rp[k] * e[0] + Rp[3 + k] * e[1] + Rp[6 + k] * e[2];
    }
    means[3 * dst] = pw[0]; means[3 * dst + 1] = pw[1]; means[3 * dst + 2] = pw[2];
    rgb[3 * dst] = gt_rgb[i]; rgb[3 * dst + 1] = gt_rgb[P + i]; rgb[3 * dst + 2] = gt_rgb[2 * P + i];
    if (quats) { quats[4 * dst] = 1.f; quats[4 * dst + 1] = 0.f; quats[4 * dst + 2] = 0.f; quats[4 * dst + 3] = 0.f; }
    logit_opac[dst] = 0.f;
    const float ls = det_logf(z / ((cam.fx + cam.fy) / 2.0f)); // splatam.py:330-331,769
    for (int k = 0; k < sd; ++k)
        log_scales[sd * dst + k] = (sd == 3 && noise_std > 0.f) ? ls + noise_std * counter_normal(seed, (uint64_t)i, k) : ls;
}

// runs after the append kernel (stream order), so no thread still reads the old count
__global__ void densify_count_kernel(int64_t P, const int32_t* __restrict__ scan, int64_t capacity, int32_t* count) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int64_t total = (int64_t)(*count) + scan[P - 1];
        *count = (int32_t)(total < capacity ? total : capacity);
    }
}

// keep the best tracking candidate (splatam.py:1287-1290): the pose AFTER this iteration's Adam step
// is saved when the loss evaluated BEFORE the step is the smallest seen so far
__global__ void pose_keep_best_kernel(const float* __restrict__ losses, const float* __restrict__ pose,
                                      float* __restrict__ best, const int32_t* __restrict__ skip) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (skip != nullptr && *skip != 0) return;
    const float loss = losses[0];
    if (loss < best[7]) {
        best[7] = loss;
        for (int k = 0; k < 7; ++k) best[k] = pose[k];
    }
}

// Start of a fused step: clears the per-step scalars (losses, pose gradient, pose accumulators ...) and advances the
// DEVICE-side Adam step counter with its bias corrections, so that the whole step can be replayed from a CUDA graph
// (the host-side `step` argument of isogs_adam_fused would be frozen into the captured launch).
__global__ void step_begin_kernel(float* __restrict__ small, int n_small, isogs_adam_state* __restrict__ st, double b1,
                                  double b2) {
    for (int i = threadIdx.x; i < n_small; i += blockDim.x) small[i] = 0.f;
    if (threadIdx.x == 0 && st != nullptr) {
        const int64_t t = st->step + 1;
        st->step = t;
        st->inv_bc1 = 1.0 / (1.0 - pow(b1, (double)t));
        st->inv_bc2_sqrt = 1.0 / sqrt(1.0 - pow(b2, (double)t));
    }
}

} // namespace isogs
using namespace isogs;

extern "C" int isogs_step_begin(float* small, int n_small, isogs_adam_state* state, float beta1, float beta2,
                                void* stream_) {
    if (n_small < 0 || (n_small > 0 && !small)) { set_error("isogs_step_begin: bad argument"); return ISOGS_E_ARG; }
    step_begin_kernel<<<1, 64, 0, (cudaStream_t)stream_>>>(small, n_small, state, (double)beta1, (double)beta2);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_pose_keep_best(const float* losses, const float* pose, float* best, const int32_t* skip_flag,
                                    void* stream_) {
    if (!losses || !pose || !best) { set_error("isogs_pose_keep_best: null argument"); return ISOGS_E_ARG; }
    pose_keep_best_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(losses, pose, best, skip_flag);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

namespace isogs {
int launch_loss_l1(const isogs_camera* cam, int mode, const float* render, const float* alpha, const float* gt_rgb,
                   const float* gt_depth, const isogs_frame* frame_dev, float w_im, float w_depth, float sil_thres,
                   int use_sil, float inv_n_depth, const float* outlier_thresh, float* v_render, float* losses,
                   cudaStream_t stream, const float* inv_n_dev = nullptr);
}

extern "C" int isogs_loss_l1_fused(const isogs_camera* cam, int mode, const float* render, const float* alpha,
                                   const float* gt_rgb, const float* gt_depth, float w_im, float w_depth,
                                   float sil_thres, int use_sil, float inv_n_depth, const float* outlier_thresh,
                                   float* v_render, float* losses, void* stream_) {
    return launch_loss_l1(cam, mode, render, alpha, gt_rgb, gt_depth, nullptr, w_im, w_depth, sil_thres, use_sil,
                          inv_n_depth, outlier_thresh, v_render, losses, (cudaStream_t)stream_);
}

int isogs::launch_loss_l1(const isogs_camera* cam, int mode, const float* render, const float* alpha, const float* gt_rgb,
                          const float* gt_depth, const isogs_frame* frame_dev, float w_im, float w_depth,
                          float sil_thres, int use_sil, float inv_n_depth, const float* outlier_thresh,
                          float* v_render, float* losses, cudaStream_t stream, const float* inv_n_dev) {
    if (!cam || (mode != 0 && mode != 1) || (!frame_dev && (!gt_rgb || !gt_depth))) {
        set_error("isogs_loss_l1_fused: bad argument");
        return ISOGS_E_ARG;
    }
    if (mode == 1 && outlier_thresh && !inv_n_dev) {   // the mean's normaliser is 1 / n(mask): data dependent
        set_error("isogs_loss_l1_fused: mapping with the outlier mask goes through isogs_map_step (ISOGS_STEP_OUTLIER)");
        return ISOGS_E_ARG;
    }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P == 0) return ISOGS_OK;
    const float inv_nd = (mode == 1) ? inv_n_depth : 1.0f;
    const float inv_nc = (mode == 1) ? 1.0f / (float)(3 * P) : 1.0f;
    const bool vec = (P % 4 == 0) && ((((uintptr_t)render | (uintptr_t)alpha | (uintptr_t)gt_rgb | (uintptr_t)gt_depth |
                                        (uintptr_t)v_render) & 15) == 0);
    if (vec) {
        const int blocks = (int)std::min<int64_t>((P / 4 + 255) / 256, 148 * 8);
        loss_l1_kernel<4><<<blocks, 256, 0, stream>>>(P, mode, render, alpha, gt_rgb, gt_depth, w_im, w_depth, sil_thres,
                                                      use_sil, inv_nd, inv_nc, outlier_thresh, v_render, losses, frame_dev,
                                                      inv_n_dev);
    } else {
        const int blocks = (int)std::min<int64_t>((P + 255) / 256, 148 * 8);
        loss_l1_kernel<1><<<blocks, 256, 0, stream>>>(P, mode, render, alpha, gt_rgb, gt_depth, w_im, w_depth, sil_thres,
                                                      use_sil, inv_nd, inv_nc, outlier_thresh, v_render, losses, frame_dev,
                                                      inv_n_dev);
    }
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

// depth_error = |gt - d| * (gt > 0); threshold = factor * lower median (torch.median), splatam.py:552-553
__global__ void __launch_bounds__(256)
depth_error_kernel(int64_t P, const float* __restrict__ render_depth, const float* __restrict__ gt_depth,
                   float* __restrict__ err, const isogs_frame* __restrict__ fr) {
    if (fr != nullptr) gt_depth = fr->gt_depth;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) { const float gd = gt_depth[i]; err[i] = fabsf(gd - render_depth[i]) * ((gd > 0.f) ? 1.f : 0.f); }
}
__global__ void median_pick_kernel(int64_t P, const float* __restrict__ sorted, float factor, float* out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = factor * sorted[(P - 1) / 2];
}
// mapping mode: also 1 / n(mask) with mask = (gt > 0) & (|gt - d| < thr) & finite (splatam.py:545-549, 556-559)
__global__ void __launch_bounds__(1024)
median_pick_count_kernel(int64_t P, const float* __restrict__ sorted, float factor, const float* __restrict__ render_depth,
                         const float* __restrict__ gt_depth, const isogs_frame* __restrict__ fr, float* out,
                         float* inv_n_out) {
    if (fr != nullptr) gt_depth = fr->gt_depth;
    const float thr = factor * sorted[(P - 1) / 2];
    int n = 0;
    for (int64_t i = threadIdx.x; i < P; i += blockDim.x) {
        const float gd = gt_depth[i], d = render_depth[i], d2 = render_depth[P + i];
        const float unc = d2 - d * d;
        n += (gd > 0.f) && (fabsf(d - gd) < thr) && !(isnan(d) || isnan(unc));
    }
    __shared__ int sm[32];
    for (int off = 16; off > 0; off >>= 1) n += __shfl_xor_sync(0xffffffffu, n, off);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 32; ++w) t += sm[w];
        *out = thr;
        *inv_n_out = t > 0 ? 1.0f / (float)t : 0.0f;
    }
}

namespace isogs {
int launch_outlier_threshold(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                             const isogs_frame* frame_dev, float factor, float* scratch, void* sort_temp,
                             size_t sort_temp_bytes, float* thresh_out, cudaStream_t stream, float* inv_n_out = nullptr);
}

extern "C" int isogs_depth_outlier_threshold(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                                             float factor, float* scratch, void* sort_temp, size_t sort_temp_bytes,
                                             float* thresh_out, void* stream_) {
    return launch_outlier_threshold(cam, render_depth, gt_depth, nullptr, factor, scratch, sort_temp, sort_temp_bytes,
                                    thresh_out, (cudaStream_t)stream_);
}

int isogs::launch_outlier_threshold(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                                    const isogs_frame* frame_dev, float factor, float* scratch, void* sort_temp,
                                    size_t sort_temp_bytes, float* thresh_out, cudaStream_t stream, float* inv_n_out) {
    if (!cam || !render_depth || (!gt_depth && !frame_dev) || !scratch || !thresh_out) {
        set_error("isogs_depth_outlier_threshold: bad argument");
        return ISOGS_E_ARG;
    }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P <= 0 || P >= (1ll << 31)) { set_error("isogs_depth_outlier_threshold: bad image size"); return ISOGS_E_ARG; }
    depth_error_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(P, render_depth, gt_depth, scratch, frame_dev);
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, need, scratch, scratch + P, (int)P, 0, 32, stream));
    if (need > sort_temp_bytes) { set_error("sort temp too small (%zu needed)", need); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceRadixSort::SortKeys(sort_temp, sort_temp_bytes, scratch, scratch + P, (int)P, 0, 32, stream));
    if (inv_n_out != nullptr)   // render_depth is plane 3 of the 5-channel render: plane 4 (depth^2) follows it
        median_pick_count_kernel<<<1, 1024, 0, stream>>>(P, scratch + P, factor, render_depth, gt_depth, frame_dev,
                                                         thresh_out, inv_n_out);
    else
        median_pick_kernel<<<1, 32, 0, stream>>>(P, scratch + P, factor, thresh_out);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

// ---- allocation-free variant of the outlier threshold ------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The depth-error
// planes and the radix-sort temp of isogs_depth_outlier_threshold live in ONE file-scope device array that every
// GPU loading the library holds until exit (ISOGS_NOALLOC_BYTES = 20 MiB).  Nothing in it has to start at zero:
// depth_error_kernel writes all H*W errors, the sort writes all H*W outputs and CUB initialises its own temp.
// Every call shares the array, so calls that overlap on the GPU (two streams, two threads) race.
constexpr int64_t NOALLOC_MAX_PIXELS = ISOGS_NOALLOC_MAX_PIXELS;
constexpr size_t NOALLOC_SORT_TEMP = (size_t)4 << 20;
constexpr size_t NOALLOC_BYTES = sizeof(float) * 2 * (size_t)NOALLOC_MAX_PIXELS + NOALLOC_SORT_TEMP;
static_assert(NOALLOC_BYTES == ISOGS_NOALLOC_BYTES, "header and kernel file disagree on the scratch size");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

// In host code g_noalloc_scratch is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, taken once per device with that device current.
static int noalloc_scratch_ptr(unsigned char** out) {
    constexpr int MAX_DEV = 64;
    static unsigned char* ptrs[MAX_DEV] = {};
    static std::mutex mu;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) { set_error("isogs_depth_outlier_threshold_noalloc: device %d out of range", dev); return ISOGS_E_ARG; }
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        ISOGS_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return ISOGS_OK;
}

extern "C" int isogs_depth_outlier_threshold_noalloc(const isogs_camera* cam, const float* render_depth,
                                                     const float* gt_depth, float factor, float* thresh_out,
                                                     void* stream_) {
    if (!cam) { set_error("isogs_depth_outlier_threshold_noalloc: null camera"); return ISOGS_E_ARG; }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P > NOALLOC_MAX_PIXELS) {
        set_error("isogs_depth_outlier_threshold_noalloc: %lld pixels exceed the built-in scratch (max %lld); use "
                  "isogs_depth_outlier_threshold with caller scratch", (long long)P, (long long)NOALLOC_MAX_PIXELS);
        return ISOGS_E_ARG;
    }
    unsigned char* base = nullptr;
    const int rc = noalloc_scratch_ptr(&base);
    if (rc != ISOGS_OK) return rc;
    float* scratch = reinterpret_cast<float*>(base);
    void* sort_temp = base + sizeof(float) * 2 * (size_t)NOALLOC_MAX_PIXELS;   // 256-byte aligned: 8 * 2^21 bytes in
    return launch_outlier_threshold(cam, render_depth, gt_depth, nullptr, factor, scratch, sort_temp, NOALLOC_SORT_TEMP,
                                    thresh_out, (cudaStream_t)stream_);
}

// uint8 HWC colour + uint16 depth -> fp32 planes (basedataset.py:299-340: color / 255, depth / png_depth_scale)
__global__ void __launch_bounds__(256)
unpack_rgbd_kernel(int64_t P, const uint8_t* __restrict__ rgb, const uint16_t* __restrict__ depth, float inv_scale,
                   float* __restrict__ gt_rgb, float* __restrict__ gt_depth) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    gt_rgb[i] = (float)rgb[3 * i] / 255.0f;
    gt_rgb[P + i] = (float)rgb[3 * i + 1] / 255.0f;
    gt_rgb[2 * P + i] = (float)rgb[3 * i + 2] / 255.0f;
    gt_depth[i] = (float)depth[i] * inv_scale;
}

extern "C" int isogs_unpack_rgbd(const isogs_camera* cam, const uint8_t* rgb_hwc, const uint16_t* depth_u16,
                                 float inv_depth_scale, float* gt_rgb, float* gt_depth, void* stream_) {
    if (!cam || !rgb_hwc || !depth_u16 || !gt_rgb || !gt_depth) { set_error("isogs_unpack_rgbd: null argument"); return ISOGS_E_ARG; }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P == 0) return ISOGS_OK;
    unpack_rgbd_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(P, rgb_hwc, depth_u16, inv_depth_scale,
                                                                                      gt_rgb, gt_depth);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_adam_fused(const isogs_adam_group* groups, int n_groups, float beta1, float beta2, float eps,
                                int step, const isogs_adam_state* state, const int32_t* skip_flag, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (state != nullptr) step = 1;   // ignored: the bias corrections come from the device
    if (!groups || n_groups < 0 || n_groups > MAX_GROUPS || step < 1) {
        set_error("isogs_adam_fused: bad argument (n_groups=%d, max %d, step=%d)", n_groups, MAX_GROUPS, step);
        return ISOGS_E_ARG;
    }
    AdamArgs a;
    a.n = 0;
    const double bc1 = 1.0 - std::pow((double)beta1, step);
    const double bc2 = 1.0 - std::pow((double)beta2, step);
    a.b1 = beta1; a.b2 = beta2;
    a.one_minus_b1 = 1.0f - beta1; a.one_minus_b2 = 1.0f - beta2;
    a.inv_bc2_sqrt = (float)(1.0 / std::sqrt(bc2));
    a.eps = eps;
    a.state = state;
    a.skip = skip_flag;
    int64_t blocks = 0;
    for (int k = 0; k < n_groups; ++k) {
        const isogs_adam_group& g = groups[k];
        if (g.grad == nullptr || g.numel <= 0) continue; // torch skips params whose .grad is None
        const int j = a.n++;
        a.p[j] = g.param; a.g[j] = g.grad; a.m[j] = g.exp_avg; a.v[j] = g.exp_avg_sq;
        a.numel[j] = g.numel;
        a.step_size[j] = state ? g.lr : (float)((double)g.lr / bc1);
        a.block_start[j] = blocks;
        blocks += (g.numel + (int64_t)ADAM_BLOCK * ADAM_VEC - 1) / ((int64_t)ADAM_BLOCK * ADAM_VEC);
    }
    a.block_start[a.n] = blocks;
    if (blocks == 0) return ISOGS_OK;
    adam_kernel<<<(unsigned)blocks, ADAM_BLOCK, 0, stream>>>(a);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_prune_mask(int64_t N, const float* logit_opac, const float* log_scales, int scale_dim,
                                float opac_thres, float max_scale, uint8_t* keep, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N < 0 || (scale_dim != 1 && scale_dim != 3)) { set_error("isogs_prune_mask: bad argument"); return ISOGS_E_ARG; }
    if (N == 0) return ISOGS_OK;
    prune_mask_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, logit_opac, log_scales, scale_dim, opac_thres,
                                                                       max_scale, keep);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_compact(int64_t N, const uint8_t* keep, const isogs_rows* rows, int n_rows, int32_t* scan,
                             void* scan_temp, size_t scan_temp_bytes, int32_t* new_count, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N < 0 || n_rows < 0 || n_rows > MAX_ROWS || N >= (1ll << 31)) { set_error("isogs_compact: bad argument"); return ISOGS_E_ARG; }
    if (N == 0) { ISOGS_CUDA(cudaMemsetAsync(new_count, 0, sizeof(int32_t), stream)); return ISOGS_OK; }
    u8_to_i32_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, keep, scan);
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, scan, scan, (int)N, stream));
    if (need > scan_temp_bytes) { set_error("scan temp too small"); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, scan, scan, (int)N, stream));
    RowsArgs ra;
    ra.n = n_rows;
    for (int r = 0; r < n_rows; ++r) { ra.in[r] = rows[r].in; ra.out[r] = rows[r].out; ra.width[r] = rows[r].width; }
    compact_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, keep, scan, ra, new_count);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_backproject_append(const isogs_camera* cam, const isogs_pose* pose, const float* gt_rgb,
                                        const float* gt_depth, const float* render_depth, const float* alpha,
                                        float sil_thres, float depth_err_thres, float* means, float* rgb, float* quats,
                                        float* logit_opac, float* log_scales, int scale_dim, float noise_std,
                                        uint64_t seed, int64_t capacity, int32_t* scan, void* scan_temp,
                                        size_t scan_temp_bytes, int32_t* count, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || (scale_dim != 1 && scale_dim != 3)) { set_error("isogs_backproject_append: bad argument"); return ISOGS_E_ARG; }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P == 0) return ISOGS_OK;
    DevCam dc = make_devcam(*cam);
    DevPose dp = make_devpose(pose, nullptr);
    // flags live in the upper half of the scan scratch: scan must hold 2*P int32
    int32_t* flag = scan + P;
    densify_mask_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(P, gt_depth, render_depth, alpha, sil_thres,
                                                                         depth_err_thres, flag);
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, flag, scan, (int)P, stream));
    if (need > scan_temp_bytes) { set_error("scan temp too small"); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, flag, scan, (int)P, stream));
    densify_append_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(dc, dp, gt_rgb, gt_depth, flag, scan, means, rgb,
                                                                           quats, logit_opac, log_scales, scale_dim,
                                                                           noise_std, seed, capacity, count);
    densify_count_kernel<<<1, 32, 0, stream>>>(P, scan, capacity, count);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
