// preprocess.cu -- north_star (1): project the 3D means, isotropic / anisotropic 2D footprint,
// radius, tile rect and tile count; plus its backward (north_star (4), the chain into the 3D
// means and the camera pose).  Reference: gsplat ProjectionEWA3DGSFused.cu:16-212,294-531.
#include <cub/device/device_scan.cuh>

#include "project.cuh"

namespace isogs {

template <bool ISO>
__global__ void __launch_bounds__(256)
preprocess_fwd_kernel(DevCam cam, DevPose pose, int64_t N, const float* __restrict__ means,
                      const float* __restrict__ scales, const float* __restrict__ quats,
                      const float* __restrict__ opacities, const float* __restrict__ colors, int ch, int flags,
                      float4* __restrict__ rec, int2* __restrict__ radii, int32_t* __restrict__ tiles,
                      float4* __restrict__ v_rec_zero, uint4* __restrict__ tile_rects, int32_t* __restrict__ tile_counts,
                      int n_bins) {
    // fused tile count (intersect_tile pass 1 + the histogram of the counting sort): CTA-private bins in shared
    // memory, ONE global atomic per (CTA, touched tile); n_bins = 0: no shared bins (more tiles than fit)
    extern __shared__ __align__(16) int32_t s_bins[];   // n_bins rounded up to a multiple of 4 (16-byte clears / reads)
    const int n_bins4 = (n_bins + 3) >> 2;
    if (tile_counts != nullptr && n_bins > 0) {
        for (int b = threadIdx.x; b < n_bins4; b += blockDim.x) reinterpret_cast<int4*>(s_bins)[b] = make_int4(0, 0, 0, 0);
        __syncthreads();
    }
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    Rect tr = {0, 0, 0, 0};
    uint32_t depth_bits = 0u;
    if (i < N) {
    if (v_rec_zero != nullptr) { // the gradient record the blend backward accumulates into (saves a memset launch)
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        v_rec_zero[3 * i] = z; v_rec_zero[3 * i + 1] = z; v_rec_zero[3 * i + 2] = z;
    }
    FrameXf xf;
    load_frame(pose, xf);
    constexpr int SD = ISO ? 1 : 3;
    float m[3] = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    float s[SD];
#pragma unroll
    for (int k = 0; k < SD; ++k) s[k] = scales[SD * i + k];
    float q[4] = {1.f, 0.f, 0.f, 0.f};
    if (!ISO) {
        float4 qq = reinterpret_cast<const float4*>(quats)[i];
        q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
    }
    float o = opacities ? opacities[i] : 0.f;
    Proj<ISO> P;
    project_one<ISO>(cam, xf, m, s, q, opacities ? &o : nullptr, flags, P);
    float4* r = rec + 3 * i;
    if (!P.valid) {
        r[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        r[1] = make_float4(0.f, 0.f, 0.f, 0.f);
        r[2] = make_float4(0.f, 0.f, 0.f, 0.f);
        radii[i] = make_int2(0, 0);
        if (tiles) tiles[i] = 0;
    } else {
        float z = P.mc[2];
        r[0] = make_float4(P.mx, P.my, 0.5f * P.a, P.b);
        r[1] = make_float4(0.5f * P.c, P.L, P.opac, colors[3 * i]);
        r[2] = make_float4(colors[3 * i + 1], colors[3 * i + 2], z, (ch == 5) ? z * z : 0.0f);
        int rx = (int)P.rx, ry = (int)P.ry;
        radii[i] = make_int2(rx, ry);
        if (tiles || tile_rects || tile_counts) {
            const Rect t = tile_rect(P.mx, P.my, rx, ry, cam.tw, cam.th);
            if (tiles) tiles[i] = (t.y1 - t.y0) * (t.x1 - t.x0);
            if (rx > 0 && ry > 0) tr = t;          // the binning kernels skip a splat unless both radii are positive
            depth_bits = __float_as_uint(z);
        }
    }
    // the tile rect + depth bits, 16 bytes per Gaussian: what the binning kernels need, so they stream this array
    // instead of gathering 28 bytes out of the 56-byte (rec, radii) rows
    if (tile_rects) tile_rects[i] = make_uint4((uint32_t)tr.x0 | ((uint32_t)tr.x1 << 16), (uint32_t)tr.y0 | ((uint32_t)tr.y1 << 16),
                                               depth_bits, 0u);
    } // i < N
    if (tile_counts != nullptr) {
        if (n_bins > 0) {
            for (int y = tr.y0; y < tr.y1; ++y)
                for (int x = tr.x0; x < tr.x1; ++x) atomicAdd(s_bins + (y * cam.tw + x), 1);
            __syncthreads();
            // a CTA's 256 Gaussians touch a few dozen of the bins: read four at a time, most groups are all zero
            for (int b = threadIdx.x; b < n_bins4; b += blockDim.x) {
                const int4 c = reinterpret_cast<const int4*>(s_bins)[b];
                if ((c.x | c.y | c.z | c.w) != 0) {
                    if (c.x) atomicAdd(tile_counts + 4 * b, c.x);
                    if (c.y) atomicAdd(tile_counts + 4 * b + 1, c.y);
                    if (c.z) atomicAdd(tile_counts + 4 * b + 2, c.z);
                    if (c.w) atomicAdd(tile_counts + 4 * b + 3, c.w);
                }
            }
        } else {
            for (int y = tr.y0; y < tr.y1; ++y)
                for (int x = tr.x0; x < tr.x1; ++x) atomicAdd(tile_counts + (y * cam.tw + x), 1);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------

// pose_acc layout: [0..8] dL/dRp, [9..11] dL/dtp, [12..15] dL/dqn (through quat_mul, aniso only)
template <bool ISO>
__global__ void __launch_bounds__(256)
preprocess_bwd_kernel(DevCam cam, DevPose pose, int64_t N, const float* __restrict__ means,
                      const float* __restrict__ scales, const float* __restrict__ quats,
                      const float* __restrict__ opacities, int ch, int flags, const int2* __restrict__ radii,
                      const float4* __restrict__ v_rec, float* __restrict__ v_means, float* __restrict__ v_scales,
                      float* __restrict__ v_quats, float* __restrict__ v_opac, float* __restrict__ v_colors,
                      float* __restrict__ v_means2d, float* __restrict__ pose_acc) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    constexpr int SD = ISO ? 1 : 3;
    constexpr int NACC = ISO ? 12 : 16;
    float acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) acc[k] = 0.f;
    FrameXf xf;
    load_frame(pose, xf);
    bool active = false;
    if (i < N) {
        int2 rad = radii[i];
        active = rad.x > 0 && rad.y > 0;
    }
    float vm[3] = {0.f, 0.f, 0.f}, vs[SD], vq[4] = {0.f, 0.f, 0.f, 0.f}, vo = 0.f, vc[3] = {0.f, 0.f, 0.f};
    float vxy[2] = {0.f, 0.f};
#pragma unroll
    for (int k = 0; k < SD; ++k) vs[k] = 0.f;
    if (active) {
        float m[3] = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
        float s[SD];
#pragma unroll
        for (int k = 0; k < SD; ++k) s[k] = scales[SD * i + k];
        float q[4] = {1.f, 0.f, 0.f, 0.f};
        if (!ISO) {
            float4 qq = reinterpret_cast<const float4*>(quats)[i];
            q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
        }
        float o = opacities ? opacities[i] : 0.f;
        Proj<ISO> P;
        project_one<ISO>(cam, xf, m, s, q, opacities ? &o : nullptr, flags, P);
        float4 g0 = v_rec[3 * i], g1 = v_rec[3 * i + 1], g2 = v_rec[3 * i + 2];
        // g0 = v_x v_y v_a v_b ; g1 = v_c v_opac v_c0 v_c1 ; g2 = v_c2 v_c3 v_c4 -
        vc[0] = g1.z; vc[1] = g1.w; vc[2] = g2.x;
        vo = g1.y;
        if (flags & ISOGS_F_LOGIT_OPAC) vo = vo * (P.opac * (1.0f - P.opac));
        float v_depth = g2.y;
        if (ch == 5) v_depth += 2.0f * P.mc[2] * g2.z;
        vxy[0] = g0.x; vxy[1] = g0.y;
        float a = P.a, b = P.b, c = P.c;
        float G00 = g0.z, G01 = 0.5f * g0.w, G11 = g1.x;
        float AG00 = a * G00 + b * G01, AG01 = a * G01 + b * G11;
        float AG10 = b * G00 + c * G01, AG11 = b * G01 + c * G11;
        float V00 = -(AG00 * a + AG01 * b);
        float V01 = -(AG00 * b + AG01 * c);
        float V10 = -(AG10 * a + AG11 * b);
        float V11 = -(AG10 * b + AG11 * c);
        float fx = cam.fx, fy = cam.fy;
        float x = P.mc[0], y = P.mc[1];
        float rz = P.rz, rz2 = P.rz2, rz3 = rz2 * rz;
        float vmc[3];
        vmc[0] = fx * rz * g0.x;
        vmc[1] = fy * rz * g0.y;
        vmc[2] = -(fx * x * g0.x + fy * y * g0.y) * rz2;
        float vJ00, vJ02, vJ11, vJ12;
        float v_cov3[ISO ? 1 : 9];
        if constexpr (ISO) {
            float sv = P.s[0], s2 = sv * sv;
            float jj00 = P.J00 * P.J00 + P.J02 * P.J02;
            float jj01 = P.J02 * P.J12;
            float jj11 = P.J11 * P.J11 + P.J12 * P.J12;
            float v_s2 = V00 * jj00 + (V01 + V10) * jj01 + V11 * jj11;
            float v_s = 2.0f * sv * v_s2;
            if (flags & ISOGS_F_LOG_SCALES) v_s = v_s * sv;
            vs[0] = v_s;
            float S00 = 2.0f * V00, S01 = V01 + V10, S11 = 2.0f * V11;
            vJ00 = s2 * (S00 * P.J00);
            vJ02 = s2 * (S00 * P.J02 + S01 * P.J12);
            vJ11 = s2 * (S11 * P.J11);
            vJ12 = s2 * (S01 * P.J02 + S11 * P.J12);
        } else {
            float J[6] = {P.J00, 0.0f, P.J02, 0.0f, P.J11, P.J12};
            float VJ[6], VtJ[6];
#pragma unroll
            for (int cc = 0; cc < 3; ++cc) {
                VJ[cc] = V00 * J[cc] + V01 * J[3 + cc];
                VJ[3 + cc] = V10 * J[cc] + V11 * J[3 + cc];
                VtJ[cc] = V00 * J[cc] + V10 * J[3 + cc];
                VtJ[3 + cc] = V01 * J[cc] + V11 * J[3 + cc];
            }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) v_cov3[r * 3 + cc] = J[r] * VJ[cc] + J[3 + r] * VJ[3 + cc];
            const float* S = P.cov3;
            float vJ[6];
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) {
                    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        s1 += VJ[r * 3 + k] * S[cc * 3 + k];
                        s2 += VtJ[r * 3 + k] * S[k * 3 + cc];
                    }
                    vJ[r * 3 + cc] = s1 + s2;
                }
            vJ00 = vJ[0]; vJ02 = vJ[2]; vJ11 = vJ[4]; vJ12 = vJ[5];
        }
        if (P.clip_x) vmc[0] += -fx * rz2 * vJ02;
        else vmc[2] += -fx * rz3 * vJ02 * P.tx;
        if (P.clip_y) vmc[1] += -fy * rz2 * vJ12;
        else vmc[2] += -fy * rz3 * vJ12 * P.ty;
        vmc[2] += -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.0f * fx * P.tx * rz3 * vJ02 + 2.0f * fy * P.ty * rz3 * vJ12;
        vmc[2] += v_depth;
        float vpf[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) vpf[k] = cam.R[k] * vmc[0] + cam.R[3 + k] * vmc[1] + cam.R[6 + k] * vmc[2];
        float vqg[4] = {0.f, 0.f, 0.f, 0.f};
        if constexpr (!ISO) {
            float t1[9], v_covw[9];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    t1[r * 3 + cc] = cam.R[r] * v_cov3[cc] + cam.R[3 + r] * v_cov3[3 + cc] + cam.R[6 + r] * v_cov3[6 + cc];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    v_covw[r * 3 + cc] = t1[r * 3] * cam.R[cc] + t1[r * 3 + 1] * cam.R[3 + cc] + t1[r * 3 + 2] * cam.R[6 + cc];
            float M[9], vM[9], vRq[9];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int k = 0; k < 3; ++k) M[r * 3 + k] = P.Rq[r * 3 + k] * P.s[k];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) {
                    float sacc = 0.f;
#pragma unroll
                    for (int k = 0; k < 3; ++k) sacc += (v_covw[r * 3 + k] + v_covw[k * 3 + r]) * M[k * 3 + cc];
                    vM[r * 3 + cc] = sacc;
                }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int k = 0; k < 3; ++k) vRq[r * 3 + k] = vM[r * 3 + k] * P.s[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float v_sk = P.Rq[k] * vM[k] + P.Rq[3 + k] * vM[3 + k] + P.Rq[6 + k] * vM[6 + k];
                if (flags & ISOGS_F_LOG_SCALES) v_sk = v_sk * P.s[k];
                vs[k] = v_sk;
            }
            quat_to_R_vjp(P.qg, vRq, vqg);
        }
        if (xf.has_pose) {
#pragma unroll
            for (int k = 0; k < 3; ++k) vm[k] = xf.Rp[k] * vpf[0] + xf.Rp[3 + k] * vpf[1] + xf.Rp[6 + k] * vpf[2];
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                acc[9 + r] = vpf[r];
#pragma unroll
                for (int k = 0; k < 3; ++k) acc[r * 3 + k] = vpf[r] * P.pw[k];
            }
            if constexpr (!ISO) {
                float qn[4], inv, Rtmp[9];
                quat_to_R(q, Rtmp, qn, &inv);
                const float* A = xf.qn;
                const float* B = qn;
                const float* ov = vqg;
                float va_[4], vb_[4];
                va_[0] = ov[0] * B[0] + ov[1] * B[1] + ov[2] * B[2] + ov[3] * B[3];
                va_[1] = -ov[0] * B[1] + ov[1] * B[0] - ov[2] * B[3] + ov[3] * B[2];
                va_[2] = -ov[0] * B[2] + ov[1] * B[3] + ov[2] * B[0] - ov[3] * B[1];
                va_[3] = -ov[0] * B[3] - ov[1] * B[2] + ov[2] * B[1] + ov[3] * B[0];
                vb_[0] = ov[0] * A[0] + ov[1] * A[1] + ov[2] * A[2] + ov[3] * A[3];
                vb_[1] = -ov[0] * A[1] + ov[1] * A[0] + ov[2] * A[3] - ov[3] * A[2];
                vb_[2] = -ov[0] * A[2] - ov[1] * A[3] + ov[2] * A[0] + ov[3] * A[1];
                vb_[3] = -ov[0] * A[3] + ov[1] * A[2] - ov[2] * A[1] + ov[3] * A[0];
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[12 + k] = va_[k];
                float dot = vb_[0] * qn[0] + vb_[1] * qn[1] + vb_[2] * qn[2] + vb_[3] * qn[3];
#pragma unroll
                for (int k = 0; k < 4; ++k) vq[k] = (vb_[k] - dot * qn[k]) * inv;
            }
        } else {
            vm[0] = vpf[0]; vm[1] = vpf[1]; vm[2] = vpf[2];
            if constexpr (!ISO) { vq[0] = vqg[0]; vq[1] = vqg[1]; vq[2] = vqg[2]; vq[3] = vqg[3]; }
        }
    }
    if (i < N) {
        if (v_means) { v_means[3 * i] = vm[0]; v_means[3 * i + 1] = vm[1]; v_means[3 * i + 2] = vm[2]; }
        if (v_scales) {
#pragma unroll
            for (int k = 0; k < SD; ++k) v_scales[SD * i + k] = vs[k];
        }
        if (!ISO && v_quats) reinterpret_cast<float4*>(v_quats)[i] = make_float4(vq[0], vq[1], vq[2], vq[3]);
        if (v_opac) v_opac[i] = vo;
        if (v_colors) { v_colors[3 * i] = vc[0]; v_colors[3 * i + 1] = vc[1]; v_colors[3 * i + 2] = vc[2]; }
        if (v_means2d) {
            // Inria-style NDC-scaled screen gradient side channel (parity unpinned)
            v_means2d[3 * i] = vxy[0] * (0.5f * (float)cam.width);
            v_means2d[3 * i + 1] = vxy[1] * (0.5f * (float)cam.height);
            v_means2d[3 * i + 2] = 0.f;
        }
    }
    // N -> 7 pose reduction: warp shuffle, then shared, then one atomic per block per component
    if (pose_acc != nullptr && xf.has_pose) {
        __shared__ float sm[8][NACC];
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < NACC; ++k) {
            float v = acc[k];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
            if (lane == 0) sm[wid][k] = v;
        }
        __syncthreads();
        if (threadIdx.x < NACC) {
            float v = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) v += sm[w][threadIdx.x];
            if (v != 0.f) atomicAdd(pose_acc + threadIdx.x, v);
        }
    }
}

// finishing kernel: chain dL/dRp (+ dL/dqn) into the un-normalised pose quaternion and translation
__global__ void pose_finish_kernel(DevPose pose, const float* __restrict__ pose_acc, float* __restrict__ v_pose) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    FrameXf xf;
    load_frame(pose, xf);
    float vR[9];
    for (int k = 0; k < 9; ++k) vR[k] = pose_acc[k];
    float vq[4] = {0.f, 0.f, 0.f, 0.f};
    quat_to_R_vjp(xf.q_raw, vR, vq);
    const float* q = xf.q_raw;
    float inv = 1.0f / sqrtf(((q[0] * q[0] + q[1] * q[1]) + q[2] * q[2]) + q[3] * q[3]);
    float aq[4] = {pose_acc[12], pose_acc[13], pose_acc[14], pose_acc[15]};
    float dot = aq[0] * xf.qn[0] + aq[1] * xf.qn[1] + aq[2] * xf.qn[2] + aq[3] * xf.qn[3];
    for (int k = 0; k < 4; ++k) v_pose[k] += vq[k] + (aq[k] - dot * xf.qn[k]) * inv;
    for (int k = 0; k < 3; ++k) v_pose[4 + k] += pose_acc[9 + k];
}

} // namespace isogs

using namespace isogs;

extern "C" int isogs_temp_bytes(int64_t N, int64_t cap, size_t* scan_bytes, size_t* sort_bytes);

extern "C" int isogs_preprocess_fwd(const isogs_camera* cam, const isogs_pose* pose, const float* pose_dev,
                                    int64_t N, const float* means, const float* scales, int scale_dim,
                                    const float* quats, const float* opacities, const float* colors, int ch,
                                    int flags, float* rec, int32_t* radii, int32_t* tiles, int32_t* cum,
                                    void* scan_temp, size_t scan_temp_bytes, int64_t* n_isects_host,
                                    float* v_rec_zero, uint32_t* tile_rects, int32_t* tile_counts, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || N < 0 || (scale_dim != 1 && scale_dim != 3) || (ch != 4 && ch != 5) || (scale_dim == 3 && !quats)) {
        set_error("isogs_preprocess_fwd: bad argument (scale_dim=%d ch=%d)", scale_dim, ch);
        return ISOGS_E_ARG;
    }
    if (N == 0) {
        if (n_isects_host) *n_isects_host = 0;
        return ISOGS_OK;
    }
    if (N >= (1ll << 31)) { set_error("N too large"); return ISOGS_E_ARG; }
    if ((((uintptr_t)rec | (uintptr_t)quats | (uintptr_t)v_rec_zero | (uintptr_t)tile_rects) & 15) || ((uintptr_t)radii & 7)) {
        set_error("isogs_preprocess_fwd: rec / quats / v_rec_zero / tile_rects must be 16-byte aligned, radii 8-byte aligned");
        return ISOGS_E_ARG;
    }
    DevCam dc = make_devcam(*cam);
    DevPose dp = make_devpose(pose, pose_dev);
    const int block = 256;
    const unsigned grid = (unsigned)((N + block - 1) / block);
    const int n_tiles = dc.tw * dc.th;
    const int n_bins = (tile_counts != nullptr && n_tiles <= 11264) ? n_tiles : 0; // 44 KB of bins: default 48 KB limit
    const size_t smem = sizeof(int32_t) * (size_t)((n_bins + 3) & ~3);
    if (scale_dim == 1)
        preprocess_fwd_kernel<true><<<grid, block, smem, stream>>>(dc, dp, N, means, scales, quats, opacities, colors,
                                                                    ch, flags, (float4*)rec, (int2*)radii, tiles,
                                                                    (float4*)v_rec_zero, (uint4*)tile_rects, tile_counts, n_bins);
    else
        preprocess_fwd_kernel<false><<<grid, block, smem, stream>>>(dc, dp, N, means, scales, quats, opacities, colors,
                                                                     ch, flags, (float4*)rec, (int2*)radii, tiles,
                                                                     (float4*)v_rec_zero, (uint4*)tile_rects, tile_counts,
                                                                     n_bins);
    ISOGS_CUDA(cudaGetLastError());
    if (!tiles || !cum) { // sync-free pipeline (isogs_tile_bin_async counts per tile itself)
        if (n_isects_host) { set_error("n_isects_host needs tiles and cum"); return ISOGS_E_ARG; }
        return ISOGS_OK;
    }
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, tiles, cum, (int)N, stream));
    if (need > scan_temp_bytes) { set_error("scan temp too small: need %zu have %zu", need, scan_temp_bytes); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceScan::InclusiveSum(scan_temp, scan_temp_bytes, tiles, cum, (int)N, stream));
    if (n_isects_host) {
        int32_t m = 0;
        ISOGS_CUDA(cudaMemcpyAsync(&m, cum + (N - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
        ISOGS_CUDA(cudaStreamSynchronize(stream));
        *n_isects_host = (int64_t)m;
    }
    return ISOGS_OK;
}

extern "C" int isogs_preprocess_bwd(const isogs_camera* cam, const isogs_pose* pose, const float* pose_dev,
                                    int64_t N, const float* means, const float* scales, int scale_dim,
                                    const float* quats, const float* opacities, int ch, int flags,
                                    const int32_t* radii, const float* v_rec, float* v_means, float* v_scales,
                                    float* v_quats, float* v_opac, float* v_colors, float* v_means2d,
                                    float* v_pose, float* pose_acc, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || N < 0 || (scale_dim != 1 && scale_dim != 3) || (ch != 4 && ch != 5) || (scale_dim == 3 && !quats)) {
        set_error("isogs_preprocess_bwd: bad argument");
        return ISOGS_E_ARG;
    }
    if (N == 0) return ISOGS_OK;
    if ((((uintptr_t)v_rec | (uintptr_t)quats | (uintptr_t)v_quats) & 15) || ((uintptr_t)radii & 7)) {
        set_error("isogs_preprocess_bwd: v_rec / quats / v_quats must be 16-byte aligned, radii 8-byte aligned");
        return ISOGS_E_ARG;
    }
    DevCam dc = make_devcam(*cam);
    DevPose dp = make_devpose(pose, pose_dev);
    const bool want_pose = (v_pose != nullptr) && dp.mode != 0;
    if (want_pose && !pose_acc) { set_error("pose_acc scratch required with v_pose"); return ISOGS_E_ARG; }
    const int block = 256;
    const unsigned grid = (unsigned)((N + block - 1) / block);
    if (scale_dim == 1)
        preprocess_bwd_kernel<true><<<grid, block, 0, stream>>>(dc, dp, N, means, scales, quats, opacities, ch, flags,
                                                                 (const int2*)radii, (const float4*)v_rec, v_means,
                                                                 v_scales, v_quats, v_opac, v_colors, v_means2d,
                                                                 want_pose ? pose_acc : nullptr);
    else
        preprocess_bwd_kernel<false><<<grid, block, 0, stream>>>(dc, dp, N, means, scales, quats, opacities, ch, flags,
                                                                  (const int2*)radii, (const float4*)v_rec, v_means,
                                                                  v_scales, v_quats, v_opac, v_colors, v_means2d,
                                                                  want_pose ? pose_acc : nullptr);
    ISOGS_CUDA(cudaGetLastError());
    if (want_pose) {
        pose_finish_kernel<<<1, 32, 0, stream>>>(dp, pose_acc, v_pose);
        ISOGS_CUDA(cudaGetLastError());
    }
    return ISOGS_OK;
}
