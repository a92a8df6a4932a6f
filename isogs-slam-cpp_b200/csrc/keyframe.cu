// keyframe.cu -- SURVEY 8(f)-4: keyframe selection by re-projection overlap
// (python_ordinary/utils/keyframe_selection.py:9-95; cpp/src/utils/keyframe_selection.cpp): sample `n` valid-depth
// pixels of the current frame, back-project them to the world, drop points that coincide (after rounding to 1e-4,
// in absolute value) with another sampled point or with the origin -- the reference's torch.unique / isin trick --
// project the rest into every keyframe and report the fraction that lands inside the image minus a 20-pixel edge.
// The reference does this with ~25 torch ops per keyframe in a Python loop; here: flag + scan of the valid pixels,
// one sampling / back-projection kernel, one O(n^2) duplicate kernel (n = 1600), one CTA per keyframe.
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace isogs {

struct Rigid { float R[9], t[3]; };

__global__ void __launch_bounds__(256)
kf_valid_kernel(int64_t P, const float* __restrict__ depth, int32_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) flag[i] = depth[i] > 0.f ? 1 : 0;
}

__device__ __forceinline__ uint32_t kf_hash(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return (uint32_t)x;
}

// one thread per sample: choose the pixel (given, or the r-th valid one by binary search in the inclusive scan),
// back-project with integer pixel coordinates (keyframe_selection.py:16-25), store the point and its rounded key
__global__ void __launch_bounds__(256)
kf_sample_kernel(DevCam cam, Rigid c2w, const float* __restrict__ depth, const int32_t* __restrict__ scan,
                 const int32_t* __restrict__ sample_idx, int n, uint64_t seed, float* __restrict__ pts,
                 float* __restrict__ key) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int64_t P = (int64_t)cam.width * cam.height;
    int64_t pix;
    if (sample_idx) {
        pix = sample_idx[s];
    } else {
        const int32_t n_valid = scan[P - 1];
        if (n_valid <= 0) { pts[3 * s] = pts[3 * s + 1] = pts[3 * s + 2] = 0.f; key[3 * s] = key[3 * s + 1] = key[3 * s + 2] = 0.f; return; }
        const uint64_t h = ((uint64_t)kf_hash(seed * 0x9E3779B97F4A7C15ull + 2ull * s) << 32) |
                           kf_hash(seed * 0x9E3779B97F4A7C15ull + 2ull * s + 1);
        const int32_t r = (int32_t)(h % (uint64_t)n_valid) + 1;        // want the first pixel with scan == r
        int64_t lo = 0, hi = P - 1;
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (scan[mid] >= r) hi = mid; else lo = mid + 1; }
        pix = lo;
    }
    const int px = (int)(pix % cam.width), py = (int)(pix / cam.width);
    const float z = depth[pix];
    const float pc[3] = {((float)px - cam.cx) / cam.fx * z, ((float)py - cam.cy) / cam.fy * z, z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float w = c2w.R[3 * k] * pc[0] + c2w.R[3 * k + 1] * pc[1] + c2w.R[3 * k + 2] * pc[2] + c2w.t[k];
        pts[3 * s + k] = w;
        key[3 * s + k] = fabsf(rintf(w * 1e4f) / 1e4f);                // torch.abs(torch.round(pts, decimals=4))
    }
}

// valid[i] = key[i] != 0 and key[i] != key[j] for every j != i  (unique(return_counts) + isin, :27-35)
__global__ void __launch_bounds__(256)
kf_unique_kernel(int n, const float* __restrict__ key, int32_t* __restrict__ valid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = key[3 * i], b = key[3 * i + 1], c = key[3 * i + 2];
    bool ok = !(a == 0.f && b == 0.f && c == 0.f);
    for (int j = 0; j < n && ok; ++j)
        if (j != i && key[3 * j] == a && key[3 * j + 1] == b && key[3 * j + 2] == c) ok = false;
    valid[i] = ok ? 1 : 0;
}

// one CTA per keyframe: fraction of the valid points that project inside [edge, W-edge] x [edge, H-edge], z > 0
__global__ void __launch_bounds__(256)
kf_overlap_kernel(DevCam cam, int n, const float* __restrict__ pts, const int32_t* __restrict__ valid,
                  const float* __restrict__ kf_w2c, int edge, float* __restrict__ percent) {
    const float* M = kf_w2c + 12 * blockIdx.x;
    float inside = 0.f, total = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (!valid[i]) continue;
        total += 1.f;
        const float x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
        const float cx_ = M[0] * x + M[1] * y + M[2] * z + M[3];
        const float cy_ = M[4] * x + M[5] * y + M[6] * z + M[7];
        const float cz_ = M[8] * x + M[9] * y + M[10] * z + M[11];
        const float pz = cz_ + 1e-5f;                                    // :73
        const float u = (cam.fx * cx_ + cam.cx * cz_) / pz, v = (cam.fy * cy_ + cam.cy * cz_) / pz;
        const bool in = (u < (float)(cam.width - edge)) && (u > (float)edge) && (v < (float)(cam.height - edge)) &&
                        (v > (float)edge) && (pz > 0.f);
        inside += in ? 1.f : 0.f;
    }
    __shared__ float s_in[8], s_tot[8];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        inside += __shfl_xor_sync(0xffffffffu, inside, off);
        total += __shfl_xor_sync(0xffffffffu, total, off);
    }
    if ((threadIdx.x & 31) == 0) { s_in[threadIdx.x >> 5] = inside; s_tot[threadIdx.x >> 5] = total; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int w = 0; w < 8; ++w) { a += s_in[w]; b += s_tot[w]; }
        percent[blockIdx.x] = a / b;                                    // mask.sum() / projected_pts.shape[0]
    }
}

} // namespace isogs

using namespace isogs;

extern "C" int isogs_keyframe_overlap(const isogs_camera* cam, const float* gt_depth, const float* cur_w2c,
                                      const int32_t* sample_idx, int n_samples, uint64_t seed, const float* kf_w2c,
                                      int n_keyframes, int edge, int32_t* scratch_i, float* scratch_f, void* scan_temp,
                                      size_t scan_temp_bytes, float* percent_inside, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || !gt_depth || !cur_w2c || n_samples <= 0 || n_keyframes < 0 || !scratch_i || !scratch_f ||
        (n_keyframes > 0 && (!kf_w2c || !percent_inside))) {
        set_error("isogs_keyframe_overlap: bad argument");
        return ISOGS_E_ARG;
    }
    const int64_t P = (int64_t)cam->width * cam->height;
    if (P <= 0 || P >= (1ll << 31)) { set_error("isogs_keyframe_overlap: bad image size"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    // rigid inverse of the current w2c (row-major 3x4 [R|t]): c2w = [R^T | -R^T t]
    Rigid c2w;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) c2w.R[3 * r + c] = cur_w2c[4 * c + r];
    for (int r = 0; r < 3; ++r)
        c2w.t[r] = -(c2w.R[3 * r] * cur_w2c[3] + c2w.R[3 * r + 1] * cur_w2c[7] + c2w.R[3 * r + 2] * cur_w2c[11]);
    int32_t* scan = scratch_i;              // [P]   (flags are scanned in place)
    int32_t* valid = scratch_i + P;         // [n]
    float* pts = scratch_f;                 // [3n]
    float* key = scratch_f + 3 * (int64_t)n_samples;   // [3n]
    if (!sample_idx) {
        kf_valid_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(P, gt_depth, scan);
        size_t need = 0;
        ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, scan, scan, (int)P, stream));
        if (need > scan_temp_bytes) { set_error("scan temp too small"); return ISOGS_E_ARG; }
        ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, scan, scan, (int)P, stream));
    }
    const unsigned gs = (unsigned)((n_samples + 255) / 256);
    kf_sample_kernel<<<gs, 256, 0, stream>>>(dc, c2w, gt_depth, scan, sample_idx, n_samples, seed, pts, key);
    kf_unique_kernel<<<gs, 256, 0, stream>>>(n_samples, key, valid);
    if (n_keyframes > 0)
        kf_overlap_kernel<<<(unsigned)n_keyframes, 256, 0, stream>>>(dc, n_samples, pts, valid, kf_w2c, edge,
                                                                     percent_inside);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
