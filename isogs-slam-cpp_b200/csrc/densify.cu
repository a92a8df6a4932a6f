// densify.cu -- SURVEY 8(a) rows a12, a13, a17: the gradient-driven clone / split densification of
// python_ordinary/utils/slam_external.py (accumulate_mean2d_gradient :100-104, densify :191-243) and of
// GaussianModel::densify + clone/split kernels (cpp/src/core/gaussian_model.cpp:157-289,
// cpp/cuda/kernels/densify.cu:10-92).  Off in every reference config (use_gaussian_splatting_densification=False)
// but part of the reference's operator surface.  All kernels are HBM-streaming, one thread per Gaussian.
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace isogs {

// a12 + a13: means2D_gradient_accum[seen] += |v_xy|, denom[seen] += 1, max_2D_radius[seen] = max(., radius)
// (slam_external.py:100-104, splatam.py:751-753); seen = radius > 0.
__global__ void __launch_bounds__(256)
densify_stats_kernel(int64_t N, const int2* __restrict__ radii, const float* __restrict__ v_xy, int stride, float sx,
                     float sy, float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radius) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int2 r = radii[i];
    const int rad = max(r.x, r.y);
    if (rad <= 0) return;
    const float gx = v_xy[i * stride] * sx, gy = v_xy[i * stride + 1] * sy;
    accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.0f;
    max_radius[i] = fmaxf(max_radius[i], (float)rad);
}

// densify :196-200,207-210: grads = accum / denom (nan -> 0); clone if grads >= thr and max scale <= 0.01 R;
// split if grads >= thr and max scale > 0.01 R.
__global__ void __launch_bounds__(256)
densify_select_kernel(int64_t N, const float* __restrict__ accum, const float* __restrict__ denom,
                      const float* __restrict__ log_scales, int sd, float grad_thresh, float small_scale,
                      int32_t* __restrict__ flag_clone, int32_t* __restrict__ flag_split) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float g = accum[i] / denom[i];
    if (isnan(g)) g = 0.0f;
    float smax = det_expf(log_scales[sd * i]);
    for (int k = 1; k < sd; ++k) smax = fmaxf(smax, det_expf(log_scales[sd * i + k]));
    const bool hot = g >= grad_thresh;
    flag_clone[i] = (hot && smax <= small_scale) ? 1 : 0;
    flag_split[i] = (hot && smax > small_scale) ? 1 : 0;
}

__global__ void densify_counts_kernel(int64_t N, const int32_t* __restrict__ scan_clone,
                                      const int32_t* __restrict__ scan_split, int n_split_into, int32_t* counts) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const int32_t nc = scan_clone[N - 1], ns = scan_split[N - 1];
        counts[0] = (int32_t)N + nc + n_split_into * ns;
        counts[1] = nc;
        counts[2] = ns;
    }
}

__device__ __forceinline__ uint32_t dz_hash(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return (uint32_t)x;
}
__device__ __forceinline__ float dz_normal(uint64_t seed, uint64_t idx, int axis) {
    const uint64_t c = seed * 0x9E3779B97F4A7C15ull + idx * 6ull + (uint64_t)axis * 2ull + 0x5bd1e995ull;
    const float u1 = ((float)(dz_hash(c) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float u2 = ((float)(dz_hash(c + 1) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

struct DensifyParams { float* means; float* rgb; float* quats; float* logit_opac; float* log_scales; int sd; };

__device__ __forceinline__ void copy_row(const DensifyParams& p, int64_t src, int64_t dst) {
    for (int k = 0; k < 3; ++k) { p.means[3 * dst + k] = p.means[3 * src + k]; p.rgb[3 * dst + k] = p.rgb[3 * src + k]; }
    if (p.quats)
        for (int k = 0; k < 4; ++k) p.quats[4 * dst + k] = p.quats[4 * src + k];
    p.logit_opac[dst] = p.logit_opac[src];
    for (int k = 0; k < p.sd; ++k) p.log_scales[p.sd * dst + k] = p.log_scales[p.sd * src + k];
}

// densify :201-223: clones appended first (in index order), then n blocks of the split rows (`.repeat(n, 1)`
// = block-wise repetition), each with mean + R(q) (std * z) and log_scale = log(exp(ls) / (0.8 n));
// parents of splits are marked for removal (:226-227).
__global__ void __launch_bounds__(256)
densify_emit_kernel(int64_t N, const int32_t* __restrict__ flag_clone, const int32_t* __restrict__ flag_split,
                    const int32_t* __restrict__ scan_clone, const int32_t* __restrict__ scan_split, DensifyParams p,
                    int n_into, const float* __restrict__ noise, uint64_t seed, int64_t capacity,
                    uint8_t* __restrict__ keep) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int32_t nc = scan_clone[N - 1], ns = scan_split[N - 1];
    const bool split = flag_split[i] != 0;
    keep[i] = split ? 0 : 1;
    if (flag_clone[i]) {
        const int64_t dst = N + scan_clone[i] - 1;
        if (dst < capacity) { copy_row(p, i, dst); keep[dst] = 1; }
    }
    if (!split) return;
    const int64_t j = scan_split[i] - 1;
    float R[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
    if (p.quats) {
        const float q[4] = {p.quats[4 * i], p.quats[4 * i + 1], p.quats[4 * i + 2], p.quats[4 * i + 3]};
        quat_to_R(q, R, nullptr, nullptr);                     // build_rotation normalises (slam_external.py:25-42)
    }
    float s[3];
    for (int k = 0; k < 3; ++k) s[k] = det_expf(p.log_scales[p.sd * i + (p.sd == 3 ? k : 0)]);
    const float shrink = 0.8f * (float)n_into;
    for (int r = 0; r < n_into; ++r) {
        const int64_t row = (int64_t)r * ns + j;               // row in the reference's [n * n_split] block order
        const int64_t dst = N + nc + row;
        if (dst >= capacity) continue;
        copy_row(p, i, dst);
        float z[3];
        for (int k = 0; k < 3; ++k)
            z[k] = s[k] * (noise ? noise[3 * row + k] : dz_normal(seed, (uint64_t)i * (uint64_t)n_into + r, k));
        for (int k = 0; k < 3; ++k)
            p.means[3 * dst + k] = p.means[3 * i + k] + (R[3 * k] * z[0] + R[3 * k + 1] * z[1] + R[3 * k + 2] * z[2]);
        for (int k = 0; k < p.sd; ++k) p.log_scales[p.sd * dst + k] = det_logf(s[p.sd == 3 ? k : 0] / shrink);
        keep[dst] = 1;
    }
}

__global__ void __launch_bounds__(256) mask_and_kernel(int64_t N, uint8_t* __restrict__ dst, const uint8_t* __restrict__ src) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) dst[i] = (dst[i] && src[i]) ? 1 : 0;
}

} // namespace isogs

using namespace isogs;

extern "C" int isogs_densify_stats(int64_t N, const int32_t* radii, const float* v_xy, int v_stride, float scale_x,
                                   float scale_y, float* grad_accum, float* denom, float* max_radius, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N < 0 || !radii || !v_xy || v_stride < 2 || !grad_accum || !denom || !max_radius) {
        set_error("isogs_densify_stats: bad argument");
        return ISOGS_E_ARG;
    }
    if (N == 0) return ISOGS_OK;
    densify_stats_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, (const int2*)radii, v_xy, v_stride, scale_x,
                                                                          scale_y, grad_accum, denom, max_radius);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_densify_select(int64_t N, const float* grad_accum, const float* denom, const float* log_scales,
                                    int scale_dim, float grad_thresh, float small_scale, int n_split_into,
                                    int32_t* scratch, void* scan_temp, size_t scan_temp_bytes, int32_t* counts,
                                    void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N <= 0 || N >= (1ll << 29) || (scale_dim != 1 && scale_dim != 3) || n_split_into < 1 || !scratch || !counts) {
        set_error("isogs_densify_select: bad argument");
        return ISOGS_E_ARG;
    }
    int32_t *fc = scratch, *fs = scratch + N, *sc = scratch + 2 * N, *ssp = scratch + 3 * N;
    const unsigned grid = (unsigned)((N + 255) / 256);
    densify_select_kernel<<<grid, 256, 0, stream>>>(N, grad_accum, denom, log_scales, scale_dim, grad_thresh, small_scale,
                                                    fc, fs);
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, fc, sc, (int)N, stream));
    if (need > scan_temp_bytes) { set_error("scan temp too small"); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, fc, sc, (int)N, stream));
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, fs, ssp, (int)N, stream));
    densify_counts_kernel<<<1, 32, 0, stream>>>(N, sc, ssp, n_split_into, counts);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_densify_emit(int64_t N, int64_t capacity, const int32_t* scratch, float* means, float* rgb,
                                  float* quats, float* logit_opac, float* log_scales, int scale_dim, int n_split_into,
                                  const float* noise, uint64_t seed, uint8_t* keep, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N <= 0 || (scale_dim != 1 && scale_dim != 3) || n_split_into < 1 || !scratch || !keep || capacity < N) {
        set_error("isogs_densify_emit: bad argument");
        return ISOGS_E_ARG;
    }
    DensifyParams p{means, rgb, quats, logit_opac, log_scales, scale_dim};
    densify_emit_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, scratch, scratch + N, scratch + 2 * N,
                                                                         scratch + 3 * N, p, n_split_into, noise, seed,
                                                                         capacity, keep);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_mask_and(int64_t N, uint8_t* dst, const uint8_t* src, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N < 0 || (N > 0 && (!dst || !src))) { set_error("isogs_mask_and: bad argument"); return ISOGS_E_ARG; }
    if (N == 0) return ISOGS_OK;
    mask_and_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, dst, src);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
