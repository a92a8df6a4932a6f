// common.cuh -- shared device helpers for the sm_100a kernels.
//
// Arithmetic contract: this translation unit family is compiled with -fmad=false, so
// a*b+c is never contracted; a fused multiply-add happens only where fmaf/__fmaf_rn is
// written.  Division, sqrt and reciprocal are the IEEE-rounded CUDA defaults (no
// --use_fast_math).  det_logf / det_expf are the deterministic elementary functions of
// DESIGN.md ("bit-exact keys"): the same operation sequence is specified for the CPU
// oracle, which makes tile keys and ranges bit-exact across CPU and GPU.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/isogs_rast.h"

namespace isogs {

constexpr float ALPHA_THRESHOLD = 1.0f / 255.0f; // gsplat include/Common.h:54
constexpr float MAX_ALPHA = 0.999f;              // RasterizeToPixels3DGSFwd.cu:148
constexpr float T_CUTOFF = 1e-4f;                // RasterizeToPixels3DGSFwd.cu:154
constexpr float MAX_EXTENT = 3.33f;              // ProjectionEWA3DGSFused.cu:164
constexpr int TILE = ISOGS_TILE;
constexpr int REC = ISOGS_REC;

void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);

#define ISOGS_CUDA(call)                                          \
    do {                                                          \
        int _rc = ::isogs::check_cuda((call), #call);             \
        if (_rc != 0) return _rc;                                 \
    } while (0)

// camera as passed to kernels (by value): the C struct plus the host-precomputed fov limits
struct DevCam {
    int width, height, tw, th;
    float fx, fy, cx, cy;
    float R[9], t[3];
    float near_plane, far_plane, eps2d;
    float bg[3];
    float lim_x_pos, lim_x_neg, lim_y_pos, lim_y_neg;
};

inline DevCam make_devcam(const isogs_camera& c) {
    DevCam d;
    d.width = c.width; d.height = c.height;
    d.tw = (c.width + TILE - 1) / TILE; d.th = (c.height + TILE - 1) / TILE;
    d.fx = c.fx; d.fy = c.fy; d.cx = c.cx; d.cy = c.cy;
    for (int i = 0; i < 9; ++i) d.R[i] = c.R[i];
    for (int i = 0; i < 3; ++i) { d.t[i] = c.t[i]; d.bg[i] = c.bg[i]; }
    d.near_plane = c.near_plane; d.far_plane = c.far_plane; d.eps2d = c.eps2d;
    // gsplat include/Utils.cuh:514-519 (host code, built with -Xcompiler -ffp-contract=off)
    float W = (float)c.width, H = (float)c.height;
    float tan_fovx = (0.5f * W) / c.fx;
    float tan_fovy = (0.5f * H) / c.fy;
    d.lim_x_pos = (W - c.cx) / c.fx + 0.3f * tan_fovx;
    d.lim_x_neg = c.cx / c.fx + 0.3f * tan_fovx;
    d.lim_y_pos = (H - c.cy) / c.fy + 0.3f * tan_fovy;
    d.lim_y_neg = c.cy / c.fy + 0.3f * tan_fovy;
    return d;
}

// pose by value (host-supplied) or by device pointer
struct DevPose {
    int mode; // 0 none, 1 by value, 2 device pointer
    float q[4], t[3];
    const float* dev;
};

inline DevPose make_devpose(const isogs_pose* p, const float* pose_dev) {
    DevPose d;
    d.mode = 0; d.dev = nullptr;
    for (int i = 0; i < 4; ++i) d.q[i] = 0.f;
    for (int i = 0; i < 3; ++i) d.t[i] = 0.f;
    if (pose_dev) { d.mode = 2; d.dev = pose_dev; }
    else if (p) {
        d.mode = 1;
        for (int i = 0; i < 4; ++i) d.q[i] = p->q[i];
        for (int i = 0; i < 3; ++i) d.t[i] = p->t[i];
    }
    return d;
}

__device__ __forceinline__ float det_logf(float x) {
    uint32_t bits = __float_as_uint(x);
    int e = (int)(bits >> 23) - 127;
    float m = __uint_as_float((bits & 0x007fffffu) | 0x3f800000u);
    if (m > 1.41421356f) { m = m * 0.5f; e += 1; }
    float f = m - 1.0f;
    float z = f * f;
    float p = 7.0376836292E-2f;
    p = fmaf(p, f, -1.1514610310E-1f);
    p = fmaf(p, f, 1.1676998740E-1f);
    p = fmaf(p, f, -1.2420140846E-1f);
    p = fmaf(p, f, 1.4249322787E-1f);
    p = fmaf(p, f, -1.6668057665E-1f);
    p = fmaf(p, f, 2.0000714765E-1f);
    p = fmaf(p, f, -2.4999993993E-1f);
    p = fmaf(p, f, 3.3333331174E-1f);
    float y = (f * z) * p;
    float fe = (float)e;
    y = fmaf(fe, -2.12194440e-4f, y);
    y = fmaf(-0.5f, z, y);
    float r = f + y;
    r = fmaf(fe, 0.693359375f, r);
    return r;
}

__device__ __forceinline__ float det_expf(float x) {
    x = fminf(88.0f, fmaxf(-87.0f, x));
    float n = floorf(fmaf(x, 1.44269504088896341f, 0.5f));
    float r = fmaf(n, -0.693359375f, x);
    r = fmaf(n, 2.12194440e-4f, r);
    float z = r * r;
    float p = 1.9875691500E-4f;
    p = fmaf(p, r, 1.3981999507E-3f);
    p = fmaf(p, r, 8.3334519073E-3f);
    p = fmaf(p, r, 4.1665795894E-2f);
    p = fmaf(p, r, 1.6666665459E-1f);
    p = fmaf(p, r, 5.0000001201E-1f);
    float y = fmaf(p, z, r) + 1.0f;
    int ni = (int)n;
    return y * __uint_as_float((uint32_t)(ni + 127) << 23);
}

__device__ __forceinline__ float det_sigmoidf(float x) { return 1.0f / (1.0f + det_expf(-x)); }

// rotation (row-major) of an un-normalised quaternion; same operation order as the oracle
__device__ __forceinline__ void quat_to_R(const float* q, float* R, float* qn, float* inv_out) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float n = sqrtf(((w * w + x * x) + y * y) + z * z);
    float inv = 1.0f / n;
    w = w * inv; x = x * inv; y = y * inv; z = z * inv;
    if (qn) { qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z; }
    if (inv_out) *inv_out = inv;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.0f - 2.0f * (y2 + z2); R[1] = 2.0f * (xy - wz);        R[2] = 2.0f * (xz + wy);
    R[3] = 2.0f * (xy + wz);        R[4] = 1.0f - 2.0f * (x2 + z2); R[5] = 2.0f * (yz - wx);
    R[6] = 2.0f * (xz - wy);        R[7] = 2.0f * (yz + wx);        R[8] = 1.0f - 2.0f * (x2 + y2);
}

// VJP of quat_to_R w.r.t. the UN-normalised quaternion (accumulates into v_q)
__device__ __forceinline__ void quat_to_R_vjp(const float* q, const float* vR, float* v_q) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float n = sqrtf(((w * w + x * x) + y * y) + z * z);
    float inv = 1.0f / n;
    w *= inv; x *= inv; y *= inv; z *= inv;
#define G(r, c) vR[(r) * 3 + (c)]
    float vw = 2.f * (x * (G(2, 1) - G(1, 2)) + y * (G(0, 2) - G(2, 0)) + z * (G(1, 0) - G(0, 1)));
    float vx = 2.f * (-2.f * x * (G(1, 1) + G(2, 2)) + y * (G(1, 0) + G(0, 1)) + z * (G(2, 0) + G(0, 2)) +
                      w * (G(2, 1) - G(1, 2)));
    float vy = 2.f * (x * (G(1, 0) + G(0, 1)) - 2.f * y * (G(0, 0) + G(2, 2)) + z * (G(2, 1) + G(1, 2)) +
                      w * (G(0, 2) - G(2, 0)));
    float vz = 2.f * (x * (G(2, 0) + G(0, 2)) + y * (G(2, 1) + G(1, 2)) - 2.f * z * (G(0, 0) + G(1, 1)) +
                      w * (G(1, 0) - G(0, 1)));
#undef G
    float dot = vw * w + vx * x + vy * y + vz * z;
    v_q[0] += (vw - dot * w) * inv;
    v_q[1] += (vx - dot * x) * inv;
    v_q[2] += (vy - dot * y) * inv;
    v_q[3] += (vz - dot * z) * inv;
}

__device__ __forceinline__ void xform(const float* R, const float* t, const float* p, float* o) {
    o[0] = ((R[0] * p[0] + R[1] * p[1]) + R[2] * p[2]) + t[0];
    o[1] = ((R[3] * p[0] + R[4] * p[1]) + R[5] * p[2]) + t[1];
    o[2] = ((R[6] * p[0] + R[7] * p[1]) + R[8] * p[2]) + t[2];
}

__device__ __forceinline__ void quat_mul(const float* a, const float* b, float* o) {
    o[0] = ((a[0] * b[0] - a[1] * b[1]) - a[2] * b[2]) - a[3] * b[3];
    o[1] = ((a[0] * b[1] + a[1] * b[0]) + a[2] * b[3]) - a[3] * b[2];
    o[2] = ((a[0] * b[2] - a[1] * b[3]) + a[2] * b[0]) + a[3] * b[1];
    o[3] = ((a[0] * b[3] + a[1] * b[2]) - a[2] * b[1]) + a[3] * b[0];
}

// One Adam element update with torch.optim.Adam semantics (bias corrections folded into step_size / inv_bc2_sqrt by the
// caller); the ONE definition used by adam_kernel (step.cu), the projection backward's inline optimizer step
// (preprocess.cu) and the fused exchange (peer.cu), so all three produce the same bits.
__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float one_minus_b1, float b2,
                                            float one_minus_b2, float inv_bc2_sqrt, float eps, float step_size) {
    const float mi = m + (g - m) * one_minus_b1;
    const float vi = v * b2 + one_minus_b2 * g * g;
    m = mi; v = vi;
    const float denom = sqrtf(vi) * inv_bc2_sqrt + eps;
    p = p - step_size * (mi / denom);
}

struct Rect { int x0, y0, x1, y1; };
// IntersectTile.cu:66-77; float->int conversion saturates on the GPU, written as the signed clamp
__device__ __forceinline__ Rect tile_rect(float mx, float my, int rx, int ry, int tw, int th) {
    float trx = (float)rx / (float)TILE, try_ = (float)ry / (float)TILE;
    float tx = mx / (float)TILE, ty = my / (float)TILE;
    Rect r;
    r.x0 = min(max(0, __float2int_rd(tx - trx)), tw);
    r.y0 = min(max(0, __float2int_rd(ty - try_)), th);
    r.x1 = min(max(0, __float2int_ru(tx + trx)), tw);
    r.y1 = min(max(0, __float2int_ru(ty + try_)), th);
    return r;
}

} // namespace isogs
