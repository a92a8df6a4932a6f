// ssim.cu -- SURVEY 8(f)-1: the SSIM term of the mapping colour loss 0.8*L1 + 0.2*(1 - SSIM)
// (python_ordinary/scripts/splatam.py:568; calc_ssim / _ssim, python_ordinary/utils/slam_external.py:54-97:
// 11x11 Gaussian window, sigma 1.5, zero padding, c1 = 0.01^2, c2 = 0.03^2, mean over [3,H,W]).
//
// The reference runs 5 depthwise 11x11 conv2d forward (+ their autograd backward) on [3,H,W].  Here:
//   ssim_fwd_kernel : one CTA per 32x32 tile of one colour plane; stages the (32+10)^2 halo tile of both images
//                     in shared memory, runs the separable window (horizontal then vertical) on the five
//                     moments {x, y, xx, yy, xy}, evaluates the SSIM map, block-reduces its sum, and writes the
//                     three partial-derivative maps dS/dmu_x, dS/dE[xx], dS/dE[xy] (the only state the backward
//                     needs).
//   ssim_bwd_kernel : the same tiling on the three derivative maps (the window is symmetric, so the adjoint of
//                     the convolution is the convolution), combined with x(p), y(p):
//                         dS/dx(p) = (w * dS/dmu_x)(p) + 2 x(p) (w * dS/dE[xx])(p) + y(p) (w * dS/dE[xy])(p)
//                     and ACCUMULATED into the colour planes of v_render that the L1 kernel wrote.
// Both are HBM-streaming: 2 planes read + 3 written, then 5 read + 1 read-modify-written per pixel-channel
// (algorithmic 20 B + 28 B per pixel-channel).  Compiled with FMA contraction on (this is not key arithmetic).
#include <cmath>

#include "common.cuh"

namespace isogs {

namespace {
constexpr int SR = 5;               // window radius (window_size 11)
constexpr int STX = 32, STY = 32;   // output tile
constexpr int SIX = STX + 2 * SR;   // 42
constexpr int SIY = STY + 2 * SR;   // 42
constexpr int SPAD = SIX + 1;       // row pitch of the staged tiles
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;

// the 11 window weights travel as a kernel argument (constant bank, statically indexed after unrolling)
struct Window {
    float g[2 * SR + 1];
    Window() {
        // gaussian(11, 1.5) of slam_external.py:54-56: float32 exponentials, normalised in float32
        float s = 0.f;
        for (int i = 0; i <= 2 * SR; ++i) {
            g[i] = (float)std::exp(-(double)((i - SR) * (i - SR)) / (2.0 * 1.5 * 1.5));
            s += g[i];
        }
        for (int i = 0; i <= 2 * SR; ++i) g[i] /= s;
    }
};

__device__ __forceinline__ float block_sum_256(float v, float* red) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x == 0)
        for (int w = 0; w < 8; ++w) t += red[w];
    return t;
}
} // namespace

// Pass structure shared by both kernels (shared-memory bandwidth, one warp-wide LDS per clock per SM, is what
// bounds a naive 11-tap loop): pass 1 gives each thread ONE staged row and a run of SEG consecutive outputs, so the
// SEG+10 inputs are read once into registers and reused across the run (lanes differ in the row: pitch 43 / 33
// keeps reads and writes conflict-free); pass 2 gives each thread one column and 4 consecutive rows (14 reads for
// 4 x 11 taps, lanes differ in the column).
constexpr int SEG = 8;                      // outputs per thread in pass 1
constexpr int HP = STX + 1;                 // pitch of the horizontally filtered rows
constexpr int P1_ITEMS = SIY * (STX / SEG); // 42 rows x 4 runs = 168 threads busy in pass 1

__global__ void __launch_bounds__(256)
ssim_fwd_kernel(int W, int H, const float* __restrict__ img_x, const float* __restrict__ img_y,
                float* __restrict__ dmaps, float add_const, float scale, float* __restrict__ losses,
                float* __restrict__ ssim_out, const Window win, const isogs_frame* __restrict__ fr) {
    if (fr != nullptr) img_y = fr->gt_rgb;   // current keyframe through the device-side frame slot (graph replay)
    __shared__ float sx[SIY][SPAD], sy[SIY][SPAD];
    __shared__ float sh[5][SIY][HP];
    __shared__ float red[8];
    const int64_t P = (int64_t)W * H;
    const int c = blockIdx.z;
    const float* __restrict__ X = img_x + c * P;
    const float* __restrict__ Y = img_y + c * P;
    const int x0 = blockIdx.x * STX - SR, y0 = blockIdx.y * STY - SR;
    for (int i = threadIdx.x; i < SIY * SIX; i += 256) {
        const int r = i / SIX, q = i - r * SIX;
        const int gx = x0 + q, gy = y0 + r;
        const bool in = (gx >= 0) && (gx < W) && (gy >= 0) && (gy < H);
        sx[r][q] = in ? X[(int64_t)gy * W + gx] : 0.f;     // zero padding (conv2d padding=5)
        sy[r][q] = in ? Y[(int64_t)gy * W + gx] : 0.f;
    }
    __syncthreads();
    if (threadIdx.x < P1_ITEMS) {
        const int r = threadIdx.x % SIY, c0 = (threadIdx.x / SIY) * SEG;
        float vx[SEG + 2 * SR], vy[SEG + 2 * SR];
#pragma unroll
        for (int k = 0; k < SEG + 2 * SR; ++k) { vx[k] = sx[r][c0 + k]; vy[k] = sy[r][c0 + k]; }
#pragma unroll
        for (int o = 0; o < SEG; ++o) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * SR; ++k) {
                const float w = win.g[k], px = vx[o + k], py = vy[o + k];
                const float wx = w * px, wy = w * py;
                a0 += wx; a1 += wy; a2 += wx * px; a3 += wy * py; a4 += wx * py;
            }
            sh[0][r][c0 + o] = a0; sh[1][r][c0 + o] = a1; sh[2][r][c0 + o] = a2; sh[3][r][c0 + o] = a3;
            sh[4][r][c0 + o] = a4;
        }
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * 4;
    float mom[5][4];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
        float v[4 + 2 * SR];
#pragma unroll
        for (int k = 0; k < 4 + 2 * SR; ++k) v[k] = sh[m][r0 + k][tx];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * SR; ++k) a += win.g[k] * v[j + k];
            mom[m][j] = a;
        }
    }
    float local = 0.f;
    const int gx = blockIdx.x * STX + tx;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int gy = blockIdx.y * STY + r0 + j;
        if (gx < W && gy < H) {
            const float mx = mom[0][j], my = mom[1][j], exx = mom[2][j], eyy = mom[3][j], exy = mom[4][j];
            const float mxx = mx * mx, myy = my * my, mxy = mx * my;
            const float A = 2.f * mxy + SSIM_C1, B = 2.f * (exy - mxy) + SSIM_C2;
            const float Cc = mxx + myy + SSIM_C1, D = (exx - mxx) + (eyy - myy) + SSIM_C2;
            const float inv_c = 1.f / Cc, inv_d = 1.f / D;
            const float inv_cd = inv_c * inv_d;
            const float S = A * B * inv_cd;
            local += S;
            const int64_t o = c * P + (int64_t)gy * W + gx;
            dmaps[o] = 2.f * my * (B - A) * inv_cd + 2.f * mx * S * (inv_d - inv_c);  // dS/dmu_x at fixed E[.]
            dmaps[3 * P + o] = -S * inv_d;                                              // dS/dE[xx]
            dmaps[6 * P + o] = 2.f * A * inv_cd;                                        // dS/dE[xy]
        }
    }
    const float tot = block_sum_256(local, red);
    if (threadIdx.x == 0) {
        const bool first = (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
        atomicAdd(ssim_out, tot * scale);                                   // mean SSIM
        if (losses) atomicAdd(losses, (first ? add_const : 0.f) - add_const * tot * scale);  // += w (1 - SSIM)
    }
}

__global__ void __launch_bounds__(256)
ssim_bwd_kernel(int W, int H, const float* __restrict__ img_x, const float* __restrict__ img_y,
                const float* __restrict__ dmaps, float coef, float* __restrict__ v_x, const Window win,
                const isogs_frame* __restrict__ fr) {
    if (fr != nullptr) img_y = fr->gt_rgb;
    __shared__ float sd[3][SIY][SPAD];
    __shared__ float sh[3][SIY][HP];
    const int64_t P = (int64_t)W * H;
    const int c = blockIdx.z;
    const int x0 = blockIdx.x * STX - SR, y0 = blockIdx.y * STY - SR;
    for (int i = threadIdx.x; i < SIY * SIX; i += 256) {
        const int r = i / SIX, q = i - r * SIX;
        const int gx = x0 + q, gy = y0 + r;
        const bool in = (gx >= 0) && (gx < W) && (gy >= 0) && (gy < H);
        const int64_t o = c * P + (int64_t)gy * W + gx;
        sd[0][r][q] = in ? dmaps[o] : 0.f;
        sd[1][r][q] = in ? dmaps[3 * P + o] : 0.f;
        sd[2][r][q] = in ? dmaps[6 * P + o] : 0.f;
    }
    __syncthreads();
    if (threadIdx.x < P1_ITEMS) {
        const int r = threadIdx.x % SIY, c0 = (threadIdx.x / SIY) * SEG;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            float v[SEG + 2 * SR];
#pragma unroll
            for (int k = 0; k < SEG + 2 * SR; ++k) v[k] = sd[m][r][c0 + k];
#pragma unroll
            for (int o = 0; o < SEG; ++o) {
                float a = 0.f;
#pragma unroll
                for (int k = 0; k <= 2 * SR; ++k) a += win.g[k] * v[o + k];
                sh[m][r][c0 + o] = a;
            }
        }
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * 4;
    float acc[3][4];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        float v[4 + 2 * SR];
#pragma unroll
        for (int k = 0; k < 4 + 2 * SR; ++k) v[k] = sh[m][r0 + k][tx];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * SR; ++k) a += win.g[k] * v[j + k];
            acc[m][j] = a;
        }
    }
    const int gx = blockIdx.x * STX + tx;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int gy = blockIdx.y * STY + r0 + j;
        if (gx < W && gy < H) {
            const int64_t o = c * P + (int64_t)gy * W + gx;
            const float x = img_x[o], y = img_y[o];
            v_x[o] += coef * (acc[0][j] + 2.f * x * acc[1][j] + y * acc[2][j]);
        }
    }
}

} // namespace isogs

using namespace isogs;

namespace isogs {
int launch_ssim_loss(const isogs_camera* cam, const float* render, const float* gt_rgb, const isogs_frame* frame_dev,
                     float weight, float* dmaps, float* v_render, float* losses, float* ssim_out, cudaStream_t stream);
}

extern "C" int isogs_ssim_loss_fused(const isogs_camera* cam, const float* render, const float* gt_rgb, float weight,
                                     float* dmaps, float* v_render, float* losses, float* ssim_out, void* stream_) {
    return launch_ssim_loss(cam, render, gt_rgb, nullptr, weight, dmaps, v_render, losses, ssim_out, (cudaStream_t)stream_);
}

int isogs::launch_ssim_loss(const isogs_camera* cam, const float* render, const float* gt_rgb,
                            const isogs_frame* frame_dev, float weight, float* dmaps, float* v_render, float* losses,
                            float* ssim_out, cudaStream_t stream) {
    if (!cam || !render || (!gt_rgb && !frame_dev) || !dmaps || !ssim_out || cam->width <= 0 || cam->height <= 0) {
        set_error("isogs_ssim_loss_fused: bad argument");
        return ISOGS_E_ARG;
    }
    static const Window win;
    const int W = cam->width, H = cam->height;
    const dim3 grid((W + STX - 1) / STX, (H + STY - 1) / STY, 3);
    const float inv_n = 1.0f / (3.0f * (float)W * (float)H);
    ssim_fwd_kernel<<<grid, 256, 0, stream>>>(W, H, render, gt_rgb, dmaps, weight, inv_n, losses, ssim_out, win, frame_dev);
    ISOGS_CUDA(cudaGetLastError());
    if (v_render) {
        ssim_bwd_kernel<<<grid, 256, 0, stream>>>(W, H, render, gt_rgb, dmaps, -weight * inv_n, v_render, win, frame_dev);
        ISOGS_CUDA(cudaGetLastError());
    }
    return ISOGS_OK;
}
