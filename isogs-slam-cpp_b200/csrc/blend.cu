// blend.cu -- north_star (3) renderCUDA forward: per 16x16 tile alpha-blend of colour, depth,
// depth^2 and silhouette with shared-memory staging of Gaussian batches and block-vote early
// termination; north_star (4) backward: reverse per-pixel traversal with a transposed
// warp reduction and one no-return atomic (RED) per (warp, Gaussian, component).
// Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188, RasterizeToPixels3DGSBwd.cu:15-278.
//
// Compiled with -fmad=false: every fmaf below is deliberate, and the sigma expression is the
// one the oracle evaluates, so the contribute/skip decision is identical on CPU and GPU.
#include "common.cuh"

namespace isogs {

constexpr int BLOCK = TILE * TILE; // 256 threads, one per pixel

struct __align__(16) Batch {
    float4 g0[BLOCK]; // x, y, 0.5*conic_a, conic_b
    float4 g1[BLOCK]; // 0.5*conic_c, L (= sigma_max), opacity, c0
    float4 g2[BLOCK]; // c1, c2, c3 (depth), c4 (depth^2)
};

__device__ __forceinline__ float sigma_of(const float4& g0, const float4& g1, float px, float py, float& dx,
                                          float& dy) {
    dx = g0.x - px;
    dy = g0.y - py;
    float u = fmaf(g0.w, dy, g0.z * dx);
    return fmaf(u, dx, (g1.x * dy) * dy);
}


// Each warp owns an 8x4 pixel block of the tile (8 blocks).  While a batch of 256 Gaussians is
// staged, every thread tests ITS Gaussian's sigma<=L bounding box against the 8 blocks and the
// warp ballots are published as bit masks; a warp then visits only the Gaussians that can touch
// its block (about 1/3 of the tile list for ~9x9 px footprints), in sorted order.
constexpr int NWARP = BLOCK / 32;
constexpr int RW = 8, RH = 4; // pixel block per warp

struct PixelMap {
    int px_i, py_i;
};
__device__ __forceinline__ PixelMap pixel_of(int tid) {
    const int wid = tid >> 5, lane = tid & 31;
    PixelMap p;
    p.px_i = blockIdx.x * TILE + (wid & 1) * RW + (lane & (RW - 1));
    p.py_i = blockIdx.y * TILE + (wid >> 1) * RH + (lane >> 3);
    return p;
}

// conservative half extents of {sigma <= L}: sqrt(2 L cov_xx), sqrt(2 L cov_yy), padded for rounding
__device__ __forceinline__ void footprint(const float4& g0, const float4& g1, float& hx, float& hy) {
    const float a = 2.0f * g0.z, b = g0.w, c = 2.0f * g1.x;
    const float inv_det = __fdividef(1.0f, fmaf(a, c, -(b * b)));
    const float twoL = 2.0f * g1.y;
    hx = fmaf(sqrtf(twoL * c * inv_det), 1.001f, 0.01f);
    hy = fmaf(sqrtf(twoL * a * inv_det), 1.001f, 0.01f);
}

// publish, for the Gaussian this thread staged, which of the 8 warp blocks it can touch
__device__ __forceinline__ void publish_masks(bool have, const float4& g0, const float4& g1,
                                              uint32_t (*s_mask)[NWARP], int wid, int lane) {
    float hx = 0.f, hy = 0.f;
    if (have) footprint(g0, g1, hx, hy);
    const float x0 = (float)(blockIdx.x * TILE), y0 = (float)(blockIdx.y * TILE);
    const float lo_x = g0.x - hx - x0, hi_x = g0.x + hx - x0; // tile-local
    const float lo_y = g0.y - hy - y0, hi_y = g0.y + hy - y0;
#pragma unroll
    for (int r = 0; r < NWARP; ++r) {
        const float rx0 = (float)((r & 1) * RW) + 0.5f, ry0 = (float)((r >> 1) * RH) + 0.5f;
        const bool hit = have && hi_x >= rx0 && lo_x <= rx0 + (float)(RW - 1) && hi_y >= ry0 &&
                         lo_y <= ry0 + (float)(RH - 1);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (lane == r) s_mask[r][wid] = m;
    }
}

template <int CH>
__global__ void __launch_bounds__(BLOCK)
blend_fwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                 int32_t* __restrict__ last) {
    __shared__ Batch sb;
    __shared__ uint32_t s_mask[NWARP][NWARP]; // [pixel block][staging warp]
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(tid);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const int start = ranges[tile], end = ranges[tile + 1];

    float T = 1.0f;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) acc[k] = 0.f;
    int last_idx = start - 1;
    bool done = !inside;

    for (int b = start; b < end; b += BLOCK) {
        // barrier (protects the staging buffers) + early exit when the whole tile is saturated
        if (__syncthreads_count(done) == BLOCK) break;
        const int idx = b + tid;
        const bool have = idx < end;
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        if (have) {
            const float4* r = rec + 3 * (int64_t)vals[idx];
            g0 = __ldg(r);
            g1 = __ldg(r + 1);
            sb.g0[tid] = g0;
            sb.g1[tid] = g1;
            sb.g2[tid] = __ldg(r + 2);
        }
        publish_masks(have, g0, g1, s_mask, wid, lane);
        __syncthreads();
        if (__all_sync(0xffffffffu, done)) continue;
#pragma unroll 1
        for (int j = 0; j < NWARP; ++j) {
            uint32_t bits = s_mask[wid][j];
            while (bits) {
                const int t = j * 32 + __ffs(bits) - 1;
                bits &= bits - 1;
                const float4 g0s = sb.g0[t];
                const float4 g1s = sb.g1[t];
                float dx, dy;
                const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
                if (done || sigma < 0.f || sigma > g1s.y) continue;
                const float a = fminf(MAX_ALPHA, g1s.z * __expf(-sigma));
                const float nT = T * (1.0f - a);
                if (nT <= T_CUTOFF) { done = true; continue; }
                const float vis = a * T;
                const float4 g2 = sb.g2[t];
                acc[0] = fmaf(g1s.w, vis, acc[0]);
                acc[1] = fmaf(g2.x, vis, acc[1]);
                acc[2] = fmaf(g2.y, vis, acc[2]);
                acc[3] = fmaf(g2.z, vis, acc[3]);
                if (CH == 5) acc[CH - 1] = fmaf(g2.w, vis, acc[CH - 1]);
                last_idx = b + t;
                T = nT;
            }
            if (__all_sync(0xffffffffu, done)) break;
        }
    }
    if (inside) {
        const int64_t P = (int64_t)cam.width * cam.height;
        const int64_t pix = (int64_t)pm.py_i * cam.width + pm.px_i;
#pragma unroll
        for (int k = 0; k < CH; ++k) render[k * P + pix] = (k < 3) ? fmaf(T, cam.bg[k], acc[k]) : acc[k];
        alpha[pix] = 1.0f - T;
        last[pix] = last_idx;
    }
}

// Sum NV (<= 16) per-lane values across the warp with 16 shuffles instead of 5*NV: at each
// halving step a lane keeps one half of its values and trades the other half with its partner.
// On return lane L holds in v[0] the warp total of component
//   j = 8*bit4(L) + 4*bit3(L) + 2*bit2(L) + bit1(L)   (both lanes of an even/odd pair hold it).
__device__ __forceinline__ void warp_reduce16(float (&v)[16], int lane) {
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float send = up ? v[k] : v[k + 8];
            float keep = up ? v[k + 8] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float send = up ? v[k] : v[k + 4];
            float keep = up ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            float send = up ? v[k] : v[k + 2];
            float keep = up ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        float send = up ? v[0] : v[1];
        float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

template <int CH>
__global__ void __launch_bounds__(BLOCK)
blend_bwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                 const int32_t* __restrict__ last, const float* __restrict__ v_render,
                 const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    __shared__ Batch sb;
    __shared__ int s_ids[BLOCK];
    __shared__ uint32_t s_mask[NWARP][NWARP];
    __shared__ int s_warp_last[NWARP];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(tid);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const int start = ranges[tile];
    const int64_t P = (int64_t)cam.width * cam.height;
    const int64_t pix = inside ? (int64_t)pm.py_i * cam.width + pm.px_i : 0;

    const int my_last = inside ? last[pix] : start - 1;
    const float T_final = inside ? 1.0f - alpha[pix] : 1.0f;
    float T = T_final;
    float vr[CH], buffer[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) {
        vr[k] = inside ? v_render[k * P + pix] : 0.f;
        buffer[k] = 0.f;
    }
    const float va = (inside && v_alpha) ? v_alpha[pix] : 0.f;
    float bg_dot = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) bg_dot += cam.bg[k] * vr[k];
    const float va_eff = T_final * (va - bg_dot); // (T_final * ra) * va - (T_final * ra) * bg_dot

    const int warp_last = __reduce_max_sync(0xffffffffu, my_last);
    if (lane == 0) s_warp_last[wid] = warp_last;
    __syncthreads();
    int block_last = s_warp_last[0];
#pragma unroll
    for (int w = 1; w < NWARP; ++w) block_last = max(block_last, s_warp_last[w]);
    if (block_last < start) return;

    // which component this lane owns after warp_reduce16
    const int comp = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    const bool writer = ((lane & 1) == 0) && comp < 6 + CH;

    // batches run back to front; inside a batch slot t holds sorted index batch_end - t
    for (int batch_end = block_last; batch_end >= start; batch_end -= BLOCK) {
        __syncthreads();
        const int idx = batch_end - tid;
        const bool have = idx >= start;
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        if (have) {
            const int g = vals[idx];
            s_ids[tid] = g;
            const float4* r = rec + 3 * (int64_t)g;
            g0 = __ldg(r);
            g1 = __ldg(r + 1);
            sb.g0[tid] = g0;
            sb.g1[tid] = g1;
            sb.g2[tid] = __ldg(r + 2);
        }
        publish_masks(have, g0, g1, s_mask, wid, lane);
        __syncthreads();
        const int t_first = batch_end - warp_last; // slots below this are behind every pixel's last id
        if (t_first >= BLOCK) continue;
#pragma unroll 1
        for (int j = max(0, t_first) >> 5; j < NWARP; ++j) {
            uint32_t bits = s_mask[wid][j];
            if (j * 32 < t_first) bits &= ~((1u << (t_first - j * 32)) - 1u);
            while (bits) {
                const int t = j * 32 + __ffs(bits) - 1;
                bits &= bits - 1;
                const float4 g0s = sb.g0[t];
                const float4 g1s = sb.g1[t];
                float dx, dy;
                const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
                const bool valid = (batch_end - t <= my_last) && !(sigma < 0.f || sigma > g1s.y);
                if (!__any_sync(0xffffffffu, valid)) continue;
                float v[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) v[k] = 0.f;
                if (valid) {
                    const float4 g2 = sb.g2[t];
                    const float c[5] = {g1s.w, g2.x, g2.y, g2.z, g2.w};
                    const float opac = g1s.z;
                    const float vis = __expf(-sigma);
                    const float ov = opac * vis;
                    const float a = fminf(MAX_ALPHA, ov);
                    const float ra = __fdividef(1.0f, 1.0f - a);
                    T *= ra;
                    const float fac = a * T;
                    float v_a = 0.f;
#pragma unroll
                    for (int k = 0; k < CH; ++k) {
                        v[6 + k] = fac * vr[k];
                        v_a = fmaf(fmaf(c[k], T, -(buffer[k] * ra)), vr[k], v_a);
                        buffer[k] = fmaf(c[k], fac, buffer[k]);
                    }
                    v_a = fmaf(va_eff, ra, v_a);
                    if (ov <= MAX_ALPHA) {
                        const float v_sigma = -ov * v_a;
                        const float ca = 2.0f * g0s.z, cb = g0s.w, cc = 2.0f * g1s.x;
                        v[0] = v_sigma * fmaf(ca, dx, cb * dy);
                        v[1] = v_sigma * fmaf(cb, dx, cc * dy);
                        v[2] = 0.5f * v_sigma * dx * dx;
                        v[3] = v_sigma * dx * dy;
                        v[4] = 0.5f * v_sigma * dy * dy;
                        v[5] = vis * v_a;
                    }
                }
                warp_reduce16(v, lane);
                if (writer) atomicAdd(v_rec + (int64_t)s_ids[t] * REC + comp, v[0]);
            }
        }
    }
}

} // namespace isogs

using namespace isogs;

extern "C" int isogs_blend_fwd(const isogs_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                               const int32_t* tile_ranges, float* render, float* alpha, int32_t* last,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_fwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    dim3 grid(dc.tw, dc.th);
    if (ch == 4)
        blend_fwd_kernel<4><<<grid, BLOCK, 0, stream>>>(dc, (const float4*)rec, vals_sorted, tile_ranges, render, alpha, last);
    else
        blend_fwd_kernel<5><<<grid, BLOCK, 0, stream>>>(dc, (const float4*)rec, vals_sorted, tile_ranges, render, alpha, last);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_blend_bwd(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                               const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                               const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    (void)N;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_bwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    dim3 grid(dc.tw, dc.th);
    if (ch == 4)
        blend_bwd_kernel<4><<<grid, BLOCK, 0, stream>>>(dc, (const float4*)rec, vals_sorted, tile_ranges, alpha, last,
                                                         v_render, v_alpha, v_rec);
    else
        blend_bwd_kernel<5><<<grid, BLOCK, 0, stream>>>(dc, (const float4*)rec, vals_sorted, tile_ranges, alpha, last,
                                                         v_render, v_alpha, v_rec);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
