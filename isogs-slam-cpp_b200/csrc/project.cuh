// project.cuh -- per-Gaussian projection shared by preprocess forward and backward.
// Mirrors (operation for operation) oracle/isogs_oracle.cpp::project_one, which follows
// gsplat ProjectionEWA3DGSFused.cu:56-208 and include/Utils.cuh:142-205,380-388,498-537.
#pragma once
#include "common.cuh"

namespace isogs {

struct FrameXf {
    bool has_pose;
    float Rp[9], tp[3], qn[4], q_raw[4];
};

__device__ __forceinline__ void load_frame(const DevPose& pose, FrameXf& xf) {
    xf.has_pose = pose.mode != 0;
    if (!xf.has_pose) return;
    float q[4], t[3];
    if (pose.mode == 2) {
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = __ldg(pose.dev + i);
#pragma unroll
        for (int i = 0; i < 3; ++i) t[i] = __ldg(pose.dev + 4 + i);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = pose.q[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) t[i] = pose.t[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) xf.q_raw[i] = q[i];
    quat_to_R(q, xf.Rp, xf.qn, nullptr);
#pragma unroll
    for (int i = 0; i < 3; ++i) xf.tp[i] = t[i];
}

template <bool ISO>
struct Proj {
    bool valid;
    float pw[3], mc[3];
    float s[ISO ? 1 : 3];
    float opac;
    float Rq[ISO ? 1 : 9];
    float qg[ISO ? 1 : 4];
    float cov3[ISO ? 1 : 9];
    float rz, rz2, tx, ty, J00, J11, J02, J12;
    bool clip_x, clip_y;
    float c00, c01, c11;
    float a, b, c;
    float mx, my, L;
    float rx, ry;
};

template <bool ISO>
__device__ __forceinline__ void project_one(const DevCam& cam, const FrameXf& xf, const float* mean,
                                            const float* scale, const float* quat, const float* opacity,
                                            int flags, Proj<ISO>& P) {
    P.valid = false;
    P.pw[0] = mean[0]; P.pw[1] = mean[1]; P.pw[2] = mean[2];
    float pf[3];
    if (xf.has_pose) xform(xf.Rp, xf.tp, P.pw, pf);
    else { pf[0] = P.pw[0]; pf[1] = P.pw[1]; pf[2] = P.pw[2]; }
    xform(cam.R, cam.t, pf, P.mc);
    if (P.mc[2] < cam.near_plane || P.mc[2] > cam.far_plane) return;

    constexpr int SD = ISO ? 1 : 3;
#pragma unroll
    for (int k = 0; k < SD; ++k) P.s[k] = (flags & ISOGS_F_LOG_SCALES) ? det_expf(scale[k]) : scale[k];
    const bool have_opac = opacity != nullptr;
    P.opac = have_opac ? ((flags & ISOGS_F_LOGIT_OPAC) ? det_sigmoidf(opacity[0]) : opacity[0]) : 1.0f;

    float x = P.mc[0], y = P.mc[1], z = P.mc[2];
    P.rz = 1.0f / z;
    P.rz2 = P.rz * P.rz;
    float xr = x * P.rz, yr = y * P.rz;
    P.clip_x = (xr <= cam.lim_x_pos) && (xr >= -cam.lim_x_neg);
    P.clip_y = (yr <= cam.lim_y_pos) && (yr >= -cam.lim_y_neg);
    P.tx = z * fminf(cam.lim_x_pos, fmaxf(-cam.lim_x_neg, xr));
    P.ty = z * fminf(cam.lim_y_pos, fmaxf(-cam.lim_y_neg, yr));
    P.J00 = cam.fx * P.rz;
    P.J11 = cam.fy * P.rz;
    P.J02 = -((cam.fx * P.tx) * P.rz2);
    P.J12 = -((cam.fy * P.ty) * P.rz2);

    if constexpr (ISO) {
        float s2 = P.s[0] * P.s[0];
        P.c00 = s2 * (P.J00 * P.J00 + P.J02 * P.J02);
        P.c01 = s2 * (P.J02 * P.J12);
        P.c11 = s2 * (P.J11 * P.J11 + P.J12 * P.J12);
    } else {
        float qg[4];
        if (xf.has_pose) {
            float qn[4], Rtmp[9];
            float qin[4] = {quat[0], quat[1], quat[2], quat[3]};
            quat_to_R(qin, Rtmp, qn, nullptr);
            quat_mul(xf.qn, qn, qg);
        } else {
            qg[0] = quat[0]; qg[1] = quat[1]; qg[2] = quat[2]; qg[3] = quat[3];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) P.qg[i] = qg[i];
        quat_to_R(qg, P.Rq, nullptr, nullptr);
        float M[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int k = 0; k < 3; ++k) M[i * 3 + k] = P.Rq[i * 3 + k] * P.s[k];
        float covw[9], tmp[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                covw[i * 3 + j] = (M[i * 3 + 0] * M[j * 3 + 0] + M[i * 3 + 1] * M[j * 3 + 1]) + M[i * 3 + 2] * M[j * 3 + 2];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                tmp[i * 3 + j] = (cam.R[i * 3 + 0] * covw[0 * 3 + j] + cam.R[i * 3 + 1] * covw[1 * 3 + j]) +
                                 cam.R[i * 3 + 2] * covw[2 * 3 + j];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                P.cov3[i * 3 + j] = (tmp[i * 3 + 0] * cam.R[j * 3 + 0] + tmp[i * 3 + 1] * cam.R[j * 3 + 1]) +
                                    tmp[i * 3 + 2] * cam.R[j * 3 + 2];
        const float* S = P.cov3;
        float a0 = P.J00 * S[0] + P.J02 * S[6];
        float a1 = P.J00 * S[1] + P.J02 * S[7];
        float a2 = P.J00 * S[2] + P.J02 * S[8];
        float b1 = P.J11 * S[4] + P.J12 * S[7];
        float b2 = P.J11 * S[5] + P.J12 * S[8];
        P.c00 = a0 * P.J00 + a2 * P.J02;
        P.c01 = a1 * P.J11 + a2 * P.J12;
        P.c11 = b1 * P.J11 + b2 * P.J12;
    }
    P.mx = (cam.fx * x) * P.rz + cam.cx;
    P.my = (cam.fy * y) * P.rz + cam.cy;
    P.c00 = P.c00 + cam.eps2d;
    P.c11 = P.c11 + cam.eps2d;
    float det = P.c00 * P.c11 - P.c01 * P.c01;
    if (det <= 0.0f) return;
    float inv_det = 1.0f / det;
    P.a = P.c11 * inv_det;
    P.b = -(P.c01 * inv_det);
    P.c = P.c00 * inv_det;

    float extent = MAX_EXTENT;
    P.L = 5.5412635f;
    if (have_opac) {
        if (P.opac < ALPHA_THRESHOLD) return;
        P.L = det_logf(P.opac / ALPHA_THRESHOLD);
        if (!(flags & ISOGS_F_NO_OPAC_EXTENT)) extent = fminf(extent, sqrtf(2.0f * P.L));
    }
    P.rx = ceilf(extent * sqrtf(P.c00));
    P.ry = ceilf(extent * sqrtf(P.c11));
    if (P.rx <= 0.0f && P.ry <= 0.0f) return;
    float W = (float)cam.width, H = (float)cam.height;
    if (P.mx + P.rx <= 0.0f || P.mx - P.rx >= W || P.my + P.ry <= 0.0f || P.my - P.ry >= H) return;
    P.valid = true;
}

} // namespace isogs
