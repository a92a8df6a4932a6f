// isogs_oracle.cpp -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
// See isogs_oracle.h for the reference file:line map and the parity status.
// Build: g++ -O2 -std=c++17 -fopenmp -ffp-contract=off -fno-fast-math -shared -fPIC
//
// Arithmetic contract (shared by specification with the CUDA kernels, which are
// compiled with -fmad=false): every expression below is evaluated in IEEE fp32,
// left to right as parenthesised, with a fused multiply-add ONLY where fmaf is
// written.  That is what makes tile keys / ranges bit-exact between CPU and GPU.
#include "isogs_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr float ALPHA_THRESHOLD = 1.0f / 255.0f; // gsplat include/Common.h:54
constexpr float MAX_ALPHA = 0.999f;              // RasterizeToPixels3DGSFwd.cu:148
constexpr float T_CUTOFF = 1e-4f;                // RasterizeToPixels3DGSFwd.cu:154
constexpr float MAX_EXTENT = 3.33f;              // ProjectionEWA3DGSFused.cu:164
constexpr int TILE = 16;

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

} // namespace

extern "C" float orc_logf(float x) {
    // Cephes-style logf; x must be a positive normal float.
    uint32_t bits = f2u(x);
    int e = (int)(bits >> 23) - 127;
    float m = u2f((bits & 0x007fffffu) | 0x3f800000u); // [1,2)
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

extern "C" float orc_expf(float x) {
    x = std::min(88.0f, std::max(-87.0f, x));
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
    return y * u2f((uint32_t)(ni + 127) << 23);
}

namespace {

inline float sigmoidf_det(float x) { return 1.0f / (1.0f + orc_expf(-x)); }

struct CamLims {
    float lim_x_pos, lim_x_neg, lim_y_pos, lim_y_neg;
};
// gsplat include/Utils.cuh:514-519
inline CamLims cam_lims(const orc_camera& c) {
    float W = (float)c.width, H = (float)c.height;
    float tan_fovx = (0.5f * W) / c.fx;
    float tan_fovy = (0.5f * H) / c.fy;
    CamLims l;
    l.lim_x_pos = (W - c.cx) / c.fx + 0.3f * tan_fovx;
    l.lim_x_neg = c.cx / c.fx + 0.3f * tan_fovx;
    l.lim_y_pos = (H - c.cy) / c.fy + 0.3f * tan_fovy;
    l.lim_y_neg = c.cy / c.fy + 0.3f * tan_fovy;
    return l;
}

// rotation matrix (row-major) of an un-normalised quaternion w,x,y,z.
// Same element formulas as gsplat Utils.cuh:142-165 and build_rotation
// (python_ordinary/utils/slam_external.py:25-42).
inline void quat_to_R(const float* q, float* R, float* qn /*normalised, nullable*/, float* inv_norm_out) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float n = sqrtf(((w * w + x * x) + y * y) + z * z);
    float inv = 1.0f / n;
    w = w * inv; x = x * inv; y = y * inv; z = z * inv;
    if (qn) { qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z; }
    if (inv_norm_out) *inv_norm_out = inv;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.0f - 2.0f * (y2 + z2); R[1] = 2.0f * (xy - wz);        R[2] = 2.0f * (xz + wy);
    R[3] = 2.0f * (xy + wz);        R[4] = 1.0f - 2.0f * (x2 + z2); R[5] = 2.0f * (yz - wx);
    R[6] = 2.0f * (xz - wy);        R[7] = 2.0f * (yz + wx);        R[8] = 1.0f - 2.0f * (x2 + y2);
}

// gradient of L w.r.t. un-normalised quaternion given dL/dR (row-major) -- Utils.cuh:167-192
inline void quat_to_R_vjp(const float* q, const float* vR, float* v_q) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float n = sqrtf(((w * w + x * x) + y * y) + z * z);
    float inv = 1.0f / n;
    w *= inv; x *= inv; y *= inv; z *= inv;
    // vR[r*3+c] = dL/dR(r,c)
    auto G = [&](int r, int c) { return vR[r * 3 + c]; };
    float vw = 2.f * (x * (G(2, 1) - G(1, 2)) + y * (G(0, 2) - G(2, 0)) + z * (G(1, 0) - G(0, 1)));
    float vx = 2.f * (-2.f * x * (G(1, 1) + G(2, 2)) + y * (G(1, 0) + G(0, 1)) + z * (G(2, 0) + G(0, 2)) +
                      w * (G(2, 1) - G(1, 2)));
    float vy = 2.f * (x * (G(1, 0) + G(0, 1)) - 2.f * y * (G(0, 0) + G(2, 2)) + z * (G(2, 1) + G(1, 2)) +
                      w * (G(0, 2) - G(2, 0)));
    float vz = 2.f * (x * (G(2, 0) + G(0, 2)) + y * (G(2, 1) + G(1, 2)) - 2.f * z * (G(0, 0) + G(1, 1)) +
                      w * (G(1, 0) - G(0, 1)));
    float dot = vw * w + vx * x + vy * y + vz * z;
    v_q[0] += (vw - dot * w) * inv;
    v_q[1] += (vx - dot * x) * inv;
    v_q[2] += (vy - dot * y) * inv;
    v_q[3] += (vz - dot * z) * inv;
}

inline void mat3_mul(const float* A, const float* B, float* C) { // C = A*B row-major
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            C[i * 3 + j] = (A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j]) + A[i * 3 + 2] * B[2 * 3 + j];
}
inline void mat3_mul_bt(const float* A, const float* B, float* C) { // C = A*B^T
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            C[i * 3 + j] = (A[i * 3 + 0] * B[j * 3 + 0] + A[i * 3 + 1] * B[j * 3 + 1]) + A[i * 3 + 2] * B[j * 3 + 2];
}
inline void mat3_mul_at(const float* A, const float* B, float* C) { // C = A^T*B
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            C[i * 3 + j] = (A[0 * 3 + i] * B[0 * 3 + j] + A[1 * 3 + i] * B[1 * 3 + j]) + A[2 * 3 + i] * B[2 * 3 + j];
}

// Everything the forward computes for one Gaussian; reused by the backward.
struct Proj {
    bool valid;
    float pw[3];   // input mean
    float pf[3];   // after frame pose
    float mc[3];   // camera space
    float s[3];    // activated scales (s[0] only when iso)
    float opac;    // activated opacity
    float Rq[9];   // gaussian rotation in camera frame (aniso)
    float qg[4];   // the quaternion handed to the rasterizer (pose-composed when pose given)
    float cov3[9]; // camera-space covariance (aniso)
    float rz, rz2, tx, ty, J00, J11, J02, J12;
    bool clip_x, clip_y; // true when x/z inside the fov limits
    float c00, c01, c11; // blurred 2d covariance
    float det, a, b, c;  // conic
    float mx, my, L;
    float rx, ry;
};

struct FrameXf {
    bool has_pose;
    float Rp[9], tp[3], qn[4]; // frame pose
    float Rv[9], tv[3];        // view
    CamLims lims;
};

inline FrameXf make_xf(const orc_camera& cam, const orc_pose* pose) {
    FrameXf x;
    x.has_pose = pose != nullptr;
    if (pose) {
        quat_to_R(pose->q, x.Rp, x.qn, nullptr);
        for (int i = 0; i < 3; ++i) x.tp[i] = pose->t[i];
    }
    for (int i = 0; i < 9; ++i) x.Rv[i] = cam.R[i];
    for (int i = 0; i < 3; ++i) x.tv[i] = cam.t[i];
    x.lims = cam_lims(cam);
    return x;
}

inline void xform(const float* R, const float* t, const float* p, float* o) {
    o[0] = ((R[0] * p[0] + R[1] * p[1]) + R[2] * p[2]) + t[0];
    o[1] = ((R[3] * p[0] + R[4] * p[1]) + R[5] * p[2]) + t[1];
    o[2] = ((R[6] * p[0] + R[7] * p[1]) + R[8] * p[2]) + t[2];
}

// quaternion product a (x) b, python_ordinary/utils/slam_helpers.py:21-28
inline void quat_mul(const float* a, const float* b, float* o) {
    o[0] = ((a[0] * b[0] - a[1] * b[1]) - a[2] * b[2]) - a[3] * b[3];
    o[1] = ((a[0] * b[1] + a[1] * b[0]) + a[2] * b[3]) - a[3] * b[2];
    o[2] = ((a[0] * b[2] - a[1] * b[3]) + a[2] * b[0]) + a[3] * b[1];
    o[3] = ((a[0] * b[3] + a[1] * b[2]) - a[2] * b[1]) + a[3] * b[0];
}

inline Proj project_one(const orc_camera& cam, const FrameXf& xf, const float* mean, const float* scale,
                        int scale_dim, const float* quat, const float* opacity, int flags) {
    Proj P;
    P.valid = false;
    for (int i = 0; i < 3; ++i) P.pw[i] = mean[i];
    if (xf.has_pose) xform(xf.Rp, xf.tp, P.pw, P.pf);
    else for (int i = 0; i < 3; ++i) P.pf[i] = P.pw[i];
    xform(xf.Rv, xf.tv, P.pf, P.mc);
    // ProjectionEWA3DGSFused.cu:72
    if (P.mc[2] < cam.near_plane || P.mc[2] > cam.far_plane) return P;

    for (int k = 0; k < scale_dim; ++k) P.s[k] = (flags & ORC_F_LOG_SCALES) ? orc_expf(scale[k]) : scale[k];
    bool have_opac = opacity != nullptr;
    P.opac = have_opac ? ((flags & ORC_F_LOGIT_OPAC) ? sigmoidf_det(opacity[0]) : opacity[0]) : 1.0f;

    const CamLims& l = xf.lims;
    float x = P.mc[0], y = P.mc[1], z = P.mc[2];
    // Utils.cuh:521-534
    P.rz = 1.0f / z;
    P.rz2 = P.rz * P.rz;
    float xr = x * P.rz, yr = y * P.rz;
    P.clip_x = (xr <= l.lim_x_pos) && (xr >= -l.lim_x_neg);
    P.clip_y = (yr <= l.lim_y_pos) && (yr >= -l.lim_y_neg);
    P.tx = z * std::min(l.lim_x_pos, std::max(-l.lim_x_neg, xr));
    P.ty = z * std::min(l.lim_y_pos, std::max(-l.lim_y_neg, yr));
    P.J00 = cam.fx * P.rz;
    P.J11 = cam.fy * P.rz;
    P.J02 = -((cam.fx * P.tx) * P.rz2);
    P.J12 = -((cam.fy * P.ty) * P.rz2);

    if (scale_dim == 1) {
        // isotropic closed form: Sigma_c = s^2 I  =>  cov2d = s^2 J J^T
        float s2 = P.s[0] * P.s[0];
        P.c00 = s2 * (P.J00 * P.J00 + P.J02 * P.J02);
        P.c01 = s2 * (P.J02 * P.J12);
        P.c11 = s2 * (P.J11 * P.J11 + P.J12 * P.J12);
    } else {
        // Utils.cuh:194-205 (Sigma = M M^T, M = R_q S), :51-57 (R Sigma R^T)
        float qg[4];
        if (xf.has_pose) {
            float qn[4], inv;
            float Rtmp[9];
            quat_to_R(quat, Rtmp, qn, &inv);
            quat_mul(xf.qn, qn, qg); // slam_helpers.py:297-300
        } else {
            for (int i = 0; i < 4; ++i) qg[i] = quat[i];
        }
        for (int i = 0; i < 4; ++i) P.qg[i] = qg[i];
        quat_to_R(qg, P.Rq, nullptr, nullptr);
        float M[9];
        for (int i = 0; i < 3; ++i)
            for (int k = 0; k < 3; ++k) M[i * 3 + k] = P.Rq[i * 3 + k] * P.s[k];
        float covw[9];
        mat3_mul_bt(M, M, covw);
        float tmp[9];
        mat3_mul(xf.Rv, covw, tmp);
        mat3_mul_bt(tmp, xf.Rv, P.cov3);
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
    // add_blur, Utils.cuh:380-388
    P.c00 = P.c00 + cam.eps2d;
    P.c11 = P.c11 + cam.eps2d;
    P.det = P.c00 * P.c11 - P.c01 * P.c01;
    if (P.det <= 0.0f) return P;
    float inv_det = 1.0f / P.det;
    P.a = P.c11 * inv_det;
    P.b = -(P.c01 * inv_det);
    P.c = P.c00 * inv_det;

    float extent = MAX_EXTENT;
    P.L = 5.5412635f; // log(255): only used when opacities are absent
    if (have_opac) {
        if (P.opac < ALPHA_THRESHOLD) return P; // ProjectionEWA3DGSFused.cu:171
        P.L = orc_logf(P.opac / ALPHA_THRESHOLD);
        if (!(flags & ORC_F_NO_OPAC_EXTENT)) extent = std::min(extent, sqrtf(2.0f * P.L)); // :178
    }
    P.rx = ceilf(extent * sqrtf(P.c00));
    P.ry = ceilf(extent * sqrtf(P.c11));
    if (P.rx <= 0.0f && P.ry <= 0.0f) return P; // radius_clip = 0
    float W = (float)cam.width, H = (float)cam.height;
    if (P.mx + P.rx <= 0.0f || P.mx - P.rx >= W || P.my + P.ry <= 0.0f || P.my - P.ry >= H) return P; // :193-198
    P.valid = true;
    return P;
}

struct Rect { int x0, y0, x1, y1; };
// IntersectTile.cu:66-77 (signed-clamp form)
inline Rect tile_rect(float mx, float my, int rx, int ry, int tw, int th) {
    float trx = (float)rx / (float)TILE, try_ = (float)ry / (float)TILE;
    float tx = mx / (float)TILE, ty = my / (float)TILE;
    auto cl = [](float v, int hi) {
        if (!(v > 0.0f)) return 0;
        if (v >= (float)hi) return hi;
        return (int)v;
    };
    Rect r;
    r.x0 = cl(floorf(tx - trx), tw);
    r.y0 = cl(floorf(ty - try_), th);
    r.x1 = cl(ceilf(tx + trx), tw);
    r.y1 = cl(ceilf(ty + try_), th);
    return r;
}

} // namespace

extern "C" void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

extern "C" int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

extern "C" void orc_preprocess_fwd(const orc_camera* cam, const orc_pose* pose, int64_t N, const float* means,
                                   const float* scales, int scale_dim, const float* quats, const float* opacities,
                                   const float* colors, int ch, int flags, float* rec, int32_t* radii,
                                   int32_t* tiles) {
    FrameXf xf = make_xf(*cam, pose);
    const int tw = (cam->width + TILE - 1) / TILE, th = (cam->height + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        Proj P = project_one(*cam, xf, means + 3 * i, scales + (int64_t)scale_dim * i,
                             scale_dim, quats ? quats + 4 * i : nullptr,
                             opacities ? opacities + i : nullptr, flags);
        float* r = rec + ORC_REC * i;
        if (!P.valid) {
            for (int k = 0; k < ORC_REC; ++k) r[k] = 0.0f;
            radii[2 * i] = 0; radii[2 * i + 1] = 0; tiles[i] = 0;
            continue;
        }
        r[0] = P.mx; r[1] = P.my; r[2] = 0.5f * P.a; r[3] = P.b; r[4] = 0.5f * P.c; r[5] = P.L; r[6] = P.opac;
        r[7] = colors[3 * i]; r[8] = colors[3 * i + 1]; r[9] = colors[3 * i + 2];
        r[10] = P.mc[2];
        r[11] = (ch == 5) ? P.mc[2] * P.mc[2] : 0.0f;
        radii[2 * i] = (int32_t)P.rx; radii[2 * i + 1] = (int32_t)P.ry;
        Rect q = tile_rect(P.mx, P.my, (int)P.rx, (int)P.ry, tw, th);
        tiles[i] = (q.y1 - q.y0) * (q.x1 - q.x0);
    }
}

extern "C" int64_t orc_tile_bin(const orc_camera* cam, int64_t N, const float* rec, const int32_t* radii,
                                const int32_t* tiles, int64_t cap, uint64_t* keys_unsorted, int32_t* vals_unsorted,
                                uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges) {
    const int tw = (cam->width + TILE - 1) / TILE, th = (cam->height + TILE - 1) / TILE;
    const int n_tiles = tw * th;
    int64_t M = 0;
    for (int64_t i = 0; i < N; ++i) M += tiles[i];
    if (keys_sorted == nullptr || M > cap) return M;
    std::vector<uint64_t> ku((size_t)M);
    std::vector<int32_t> vu((size_t)M);
    // duplicateWithKeys: IntersectTile.cu:87-113 (image id bits are zero for one image)
    int64_t cur = 0;
    for (int64_t i = 0; i < N; ++i) {
        if (radii[2 * i] <= 0 || radii[2 * i + 1] <= 0) continue;
        const float* r = rec + ORC_REC * i;
        Rect q = tile_rect(r[0], r[1], radii[2 * i], radii[2 * i + 1], tw, th);
        uint64_t depth_bits = (uint64_t)f2u(r[10]);
        for (int y = q.y0; y < q.y1; ++y)
            for (int x = q.x0; x < q.x1; ++x) {
                uint64_t tile_id = (uint64_t)(y * tw + x);
                ku[cur] = (tile_id << 32) | depth_bits;
                vu[cur] = (int32_t)i;
                ++cur;
            }
    }
    if (keys_unsorted) std::memcpy(keys_unsorted, ku.data(), sizeof(uint64_t) * M);
    if (vals_unsorted) std::memcpy(vals_unsorted, vu.data(), sizeof(int32_t) * M);
    // stable LSD radix sort == std::stable_sort on the key (IntersectTile.cu:296-339)
    std::vector<int64_t> perm((size_t)M);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int64_t a, int64_t b) { return ku[a] < ku[b]; });
    for (int64_t j = 0; j < M; ++j) { keys_sorted[j] = ku[perm[j]]; vals_sorted[j] = vu[perm[j]]; }
    // identifyTileRanges: IntersectTile.cu:227-256; ranges[t] = first index with tile >= t, ranges[n_tiles] = M
    int64_t j = 0;
    for (int t = 0; t <= n_tiles; ++t) {
        while (j < M && (int64_t)(keys_sorted[j] >> 32) < t) ++j;
        tile_ranges[t] = (int32_t)j;
    }
    return M;
}

namespace {
inline float sigma_of(const float* r, float px, float py, float& dx, float& dy) {
    dx = r[0] - px;
    dy = r[1] - py;
    float u = fmaf(r[3], dy, r[2] * dx);
    return fmaf(u, dx, (r[4] * dy) * dy);
}
} // namespace

extern "C" void orc_blend_fwd(const orc_camera* cam, int ch, const float* rec, const int32_t* vals,
                              const int32_t* ranges, float* render, float* alpha, int32_t* last) {
    const int W = cam->width, H = cam->height;
    const int tw = (W + TILE - 1) / TILE;
    const int64_t P = (int64_t)W * H;
#pragma omp parallel for schedule(dynamic, 4)
    for (int py_i = 0; py_i < H; ++py_i) {
        for (int px_i = 0; px_i < W; ++px_i) {
            int tile = (py_i / TILE) * tw + (px_i / TILE);
            int start = ranges[tile], end = ranges[tile + 1];
            float px = (float)px_i + 0.5f, py = (float)py_i + 0.5f; // RasterizeToPixels3DGSFwd.cu:62-63
            float T = 1.0f;
            float acc[5] = {0, 0, 0, 0, 0};
            int last_idx = start - 1;
            for (int idx = start; idx < end; ++idx) {
                const float* r = rec + ORC_REC * (int64_t)vals[idx];
                float dx, dy;
                float sigma = sigma_of(r, px, py, dx, dy);
                if (sigma < 0.0f || sigma > r[5]) continue; // == alpha < 1/255, :149
                float a = std::min(MAX_ALPHA, r[6] * expf(-sigma));
                float nT = T * (1.0f - a);
                if (nT <= T_CUTOFF) break; // :153-157
                float vis = a * T;
                for (int k = 0; k < ch; ++k) acc[k] = fmaf(r[7 + k], vis, acc[k]);
                last_idx = idx;
                T = nT;
            }
            int64_t pix = (int64_t)py_i * W + px_i;
            for (int k = 0; k < ch; ++k) render[k * P + pix] = (k < 3) ? fmaf(T, cam->bg[k], acc[k]) : acc[k];
            alpha[pix] = 1.0f - T; // :178
            last[pix] = last_idx;
        }
    }
}

extern "C" void orc_blend_bwd(const orc_camera* cam, int ch, int64_t N, const float* rec, const int32_t* vals,
                              const int32_t* ranges, const float* alpha, const int32_t* last,
                              const float* v_render, const float* v_alpha_img, float* v_rec) {
    const int W = cam->width, H = cam->height;
    const int tw = (W + TILE - 1) / TILE;
    const int64_t P = (int64_t)W * H;
    std::vector<double> acc((size_t)N * ORC_REC, 0.0);
#pragma omp parallel for schedule(dynamic, 4)
    for (int py_i = 0; py_i < H; ++py_i) {
        for (int px_i = 0; px_i < W; ++px_i) {
            int tile = (py_i / TILE) * tw + (px_i / TILE);
            int start = ranges[tile];
            int64_t pix = (int64_t)py_i * W + px_i;
            float px = (float)px_i + 0.5f, py = (float)py_i + 0.5f;
            float T_final = 1.0f - alpha[pix];
            float T = T_final;
            float buffer[5] = {0, 0, 0, 0, 0};
            float vr[5];
            for (int k = 0; k < ch; ++k) vr[k] = v_render[k * P + pix];
            float va = v_alpha_img ? v_alpha_img[pix] : 0.0f;
            float bg_dot = 0.0f;
            for (int k = 0; k < 3 && k < ch; ++k) bg_dot += cam->bg[k] * vr[k];
            for (int idx = last[pix]; idx >= start; --idx) {
                int64_t g = vals[idx];
                const float* r = rec + ORC_REC * g;
                float dx, dy;
                float sigma = sigma_of(r, px, py, dx, dy);
                if (sigma < 0.0f || sigma > r[5]) continue;
                float vis = expf(-sigma);
                float opac = r[6];
                float a = std::min(MAX_ALPHA, opac * vis);
                // RasterizeToPixels3DGSBwd.cu:195-237
                float ra = 1.0f / (1.0f - a);
                T *= ra;
                float fac = a * T;
                double* o = &acc[(size_t)g * ORC_REC];
                float v_a = 0.0f;
                for (int k = 0; k < ch; ++k) {
                    float vc = fac * vr[k];
#pragma omp atomic
                    o[6 + k] += (double)vc;
                    v_a += (r[7 + k] * T - buffer[k] * ra) * vr[k];
                }
                v_a += T_final * ra * va;
                v_a += -T_final * ra * bg_dot;
                if (opac * vis <= MAX_ALPHA) {
                    float v_sigma = -opac * vis * v_a;
                    float ca = 2.0f * r[2], cb = r[3], cc = 2.0f * r[4];
                    double g0 = v_sigma * (ca * dx + cb * dy);
                    double g1 = v_sigma * (cb * dx + cc * dy);
                    double g2 = 0.5f * v_sigma * dx * dx;
                    double g3 = v_sigma * dx * dy;
                    double g4 = 0.5f * v_sigma * dy * dy;
                    double g5 = vis * v_a;
#pragma omp atomic
                    o[0] += g0;
#pragma omp atomic
                    o[1] += g1;
#pragma omp atomic
                    o[2] += g2;
#pragma omp atomic
                    o[3] += g3;
#pragma omp atomic
                    o[4] += g4;
#pragma omp atomic
                    o[5] += g5;
                }
                for (int k = 0; k < ch; ++k) buffer[k] += r[7 + k] * fac;
            }
        }
    }
    for (size_t i = 0; i < acc.size(); ++i) v_rec[i] = (float)acc[i];
}

extern "C" void orc_preprocess_bwd(const orc_camera* cam, const orc_pose* pose, int64_t N, const float* means,
                                   const float* scales, int scale_dim, const float* quats, const float* opacities,
                                   const float* colors, int ch, int flags, const int32_t* radii, const float* rec,
                                   const float* v_rec, float* v_means, float* v_scales, float* v_quats,
                                   float* v_opac, float* v_colors, double* v_pose) {
    (void)colors; (void)rec;
    FrameXf xf = make_xf(*cam, pose);
    // pose accumulators: dL/dRp (9), dL/dtp (3), dL/dqn (4, through quat_mul for aniso)
    double aR[9] = {0}, at[3] = {0}, aq[4] = {0};
#pragma omp parallel
    {
        double lR[9] = {0}, lt[3] = {0}, lq[4] = {0};
#pragma omp for schedule(static)
        for (int64_t i = 0; i < N; ++i) {
            float* vm = v_means + 3 * i;
            float* vs = v_scales + (int64_t)scale_dim * i;
            float* vq = v_quats ? v_quats + 4 * i : nullptr;
            vm[0] = vm[1] = vm[2] = 0.0f;
            for (int k = 0; k < scale_dim; ++k) vs[k] = 0.0f;
            if (vq) vq[0] = vq[1] = vq[2] = vq[3] = 0.0f;
            v_opac[i] = 0.0f;
            v_colors[3 * i] = v_colors[3 * i + 1] = v_colors[3 * i + 2] = 0.0f;
            if (radii[2 * i] <= 0 || radii[2 * i + 1] <= 0) continue; // ProjectionEWA3DGSFused.cu:327
            Proj P = project_one(*cam, xf, means + 3 * i, scales + (int64_t)scale_dim * i, scale_dim,
                                 quats ? quats + 4 * i : nullptr, opacities ? opacities + i : nullptr, flags);
            const float* g = v_rec + ORC_REC * i;
            v_colors[3 * i] = g[6]; v_colors[3 * i + 1] = g[7]; v_colors[3 * i + 2] = g[8];
            // opacity (activation chain: sigmoid' = o(1-o))
            float vo = g[5];
            if (flags & ORC_F_LOGIT_OPAC) vo = vo * (P.opac * (1.0f - P.opac));
            v_opac[i] = vo;
            // depth channels: c3 = z, c4 = z^2
            float v_depth = g[9];
            if (ch == 5) v_depth += 2.0f * P.mc[2] * g[10];
            // conic -> cov2d : inverse_vjp, Utils.cuh:373-378 with the symmetric packing of :346-348
            float a = P.a, b = P.b, c = P.c;
            float G00 = g[2], G01 = 0.5f * g[3], G11 = g[4];
            float AG00 = a * G00 + b * G01, AG01 = a * G01 + b * G11;
            float AG10 = b * G00 + c * G01, AG11 = b * G01 + c * G11;
            float V00 = -(AG00 * a + AG01 * b);
            float V01 = -(AG00 * b + AG01 * c);
            float V10 = -(AG10 * a + AG11 * b);
            float V11 = -(AG10 * b + AG11 * c);
            // persp_proj_vjp, Utils.cuh:539-616
            float fx = cam->fx, fy = cam->fy;
            float x = P.mc[0], y = P.mc[1];
            float rz = P.rz, rz2 = P.rz2, rz3 = rz2 * rz;
            float vmc[3];
            vmc[0] = fx * rz * g[0];
            vmc[1] = fy * rz * g[1];
            vmc[2] = -(fx * x * g[0] + fy * y * g[1]) * rz2;
            float vJ00, vJ02, vJ11, vJ12;
            float v_cov3[9];
            if (scale_dim == 1) {
                float s = P.s[0], s2 = s * s;
                float jj00 = P.J00 * P.J00 + P.J02 * P.J02;
                float jj01 = P.J02 * P.J12;
                float jj11 = P.J11 * P.J11 + P.J12 * P.J12;
                float v_s2 = V00 * jj00 + (V01 + V10) * jj01 + V11 * jj11;
                float v_s = 2.0f * s * v_s2;
                if (flags & ORC_F_LOG_SCALES) v_s = v_s * s;
                vs[0] = v_s;
                // v_J = s2 (V + V^T) J
                float S00 = 2.0f * V00, S01 = V01 + V10, S11 = 2.0f * V11;
                vJ00 = s2 * (S00 * P.J00);
                vJ02 = s2 * (S00 * P.J02 + S01 * P.J12);
                vJ11 = s2 * (S11 * P.J11);
                vJ12 = s2 * (S01 * P.J02 + S11 * P.J12);
            } else {
                // v_cov3 = J^T V J  (3x3), J = [[J00,0,J02],[0,J11,J12]]
                float J[6] = {P.J00, 0.0f, P.J02, 0.0f, P.J11, P.J12}; // 2x3 row-major
                float VJ[6];
                for (int cidx = 0; cidx < 3; ++cidx) {
                    VJ[0 * 3 + cidx] = V00 * J[cidx] + V01 * J[3 + cidx];
                    VJ[1 * 3 + cidx] = V10 * J[cidx] + V11 * J[3 + cidx];
                }
                for (int r = 0; r < 3; ++r)
                    for (int cidx = 0; cidx < 3; ++cidx)
                        v_cov3[r * 3 + cidx] = J[r] * VJ[cidx] + J[3 + r] * VJ[3 + cidx];
                // v_J = V J Sigma^T + V^T J Sigma  (2x3)
                const float* S = P.cov3;
                float VtJ[6];
                for (int cidx = 0; cidx < 3; ++cidx) {
                    VtJ[0 * 3 + cidx] = V00 * J[cidx] + V10 * J[3 + cidx];
                    VtJ[1 * 3 + cidx] = V01 * J[cidx] + V11 * J[3 + cidx];
                }
                float vJ[6];
                for (int r = 0; r < 2; ++r)
                    for (int cidx = 0; cidx < 3; ++cidx) {
                        float s1 = 0.0f, s2 = 0.0f;
                        for (int k = 0; k < 3; ++k) {
                            s1 += VJ[r * 3 + k] * S[cidx * 3 + k];  // (VJ) Sigma^T
                            s2 += VtJ[r * 3 + k] * S[k * 3 + cidx]; // (V^T J) Sigma
                        }
                        vJ[r * 3 + cidx] = s1 + s2;
                    }
                vJ00 = vJ[0]; vJ02 = vJ[2]; vJ11 = vJ[4]; vJ12 = vJ[5];
            }
            if (P.clip_x) vmc[0] += -fx * rz2 * vJ02;
            else vmc[2] += -fx * rz3 * vJ02 * P.tx;
            if (P.clip_y) vmc[1] += -fy * rz2 * vJ12;
            else vmc[2] += -fy * rz3 * vJ12 * P.ty;
            vmc[2] += -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.0f * fx * P.tx * rz3 * vJ02 +
                      2.0f * fy * P.ty * rz3 * vJ12;
            vmc[2] += v_depth; // ProjectionEWA3DGSFused.cu:458-459 (v_depths)
            // view transform: v_pf = Rv^T vmc
            float vpf[3];
            for (int k = 0; k < 3; ++k) vpf[k] = xf.Rv[0 * 3 + k] * vmc[0] + xf.Rv[3 + k] * vmc[1] + xf.Rv[6 + k] * vmc[2];
            float vqg[4] = {0, 0, 0, 0};
            if (scale_dim == 3) {
                // covarW2C_VJP: v_covw = Rv^T v_cov3 Rv, Utils.cuh:63-80
                float t1[9], v_covw[9];
                mat3_mul_at(xf.Rv, v_cov3, t1);
                mat3_mul(t1, xf.Rv, v_covw);
                // quat_scale_to_covar_vjp, Utils.cuh:224-261
                float M[9];
                for (int r = 0; r < 3; ++r)
                    for (int k = 0; k < 3; ++k) M[r * 3 + k] = P.Rq[r * 3 + k] * P.s[k];
                float Gs[9];
                for (int r = 0; r < 3; ++r)
                    for (int cidx = 0; cidx < 3; ++cidx) Gs[r * 3 + cidx] = v_covw[r * 3 + cidx] + v_covw[cidx * 3 + r];
                float vM[9];
                mat3_mul(Gs, M, vM);
                float vRq[9];
                for (int r = 0; r < 3; ++r)
                    for (int k = 0; k < 3; ++k) vRq[r * 3 + k] = vM[r * 3 + k] * P.s[k];
                for (int k = 0; k < 3; ++k) {
                    float v_sk = P.Rq[0 * 3 + k] * vM[0 * 3 + k] + P.Rq[3 + k] * vM[3 + k] + P.Rq[6 + k] * vM[6 + k];
                    if (flags & ORC_F_LOG_SCALES) v_sk = v_sk * P.s[k];
                    vs[k] = v_sk;
                }
                quat_to_R_vjp(P.qg, vRq, vqg);
            }
            if (xf.has_pose) {
                // p_f = Rp p_w + tp
                for (int k = 0; k < 3; ++k) vm[k] = xf.Rp[0 * 3 + k] * vpf[0] + xf.Rp[3 + k] * vpf[1] + xf.Rp[6 + k] * vpf[2];
                for (int r = 0; r < 3; ++r) {
                    lt[r] += vpf[r];
                    for (int k = 0; k < 3; ++k) lR[r * 3 + k] += (double)vpf[r] * P.pw[k];
                }
                if (scale_dim == 3 && vq) {
                    // qg = qcam (x) normalize(q): gradient to both factors
                    float qn[4], inv, Rtmp[9];
                    quat_to_R(quats + 4 * i, Rtmp, qn, &inv);
                    const float* A = xf.qn; // a
                    const float* B = qn;    // b
                    const float* o = vqg;
                    float va_[4], vb_[4];
                    va_[0] = o[0] * B[0] + o[1] * B[1] + o[2] * B[2] + o[3] * B[3];
                    va_[1] = -o[0] * B[1] + o[1] * B[0] - o[2] * B[3] + o[3] * B[2];
                    va_[2] = -o[0] * B[2] + o[1] * B[3] + o[2] * B[0] - o[3] * B[1];
                    va_[3] = -o[0] * B[3] - o[1] * B[2] + o[2] * B[1] + o[3] * B[0];
                    vb_[0] = o[0] * A[0] + o[1] * A[1] + o[2] * A[2] + o[3] * A[3];
                    vb_[1] = -o[0] * A[1] + o[1] * A[0] + o[2] * A[3] - o[3] * A[2];
                    vb_[2] = -o[0] * A[2] - o[1] * A[3] + o[2] * A[0] + o[3] * A[1];
                    vb_[3] = -o[0] * A[3] + o[1] * A[2] - o[2] * A[1] + o[3] * A[0];
                    for (int k = 0; k < 4; ++k) lq[k] += va_[k];
                    // through normalize(q): v_q = (vb - (vb.qn) qn) * inv
                    float dot = vb_[0] * qn[0] + vb_[1] * qn[1] + vb_[2] * qn[2] + vb_[3] * qn[3];
                    for (int k = 0; k < 4; ++k) vq[k] = (vb_[k] - dot * qn[k]) * inv;
                }
            } else {
                for (int k = 0; k < 3; ++k) vm[k] = vpf[k];
                if (scale_dim == 3 && vq) for (int k = 0; k < 4; ++k) vq[k] = vqg[k];
            }
        }
#pragma omp critical
        {
            for (int k = 0; k < 9; ++k) aR[k] += lR[k];
            for (int k = 0; k < 3; ++k) at[k] += lt[k];
            for (int k = 0; k < 4; ++k) aq[k] += lq[k];
        }
    }
    if (v_pose && pose) {
        // chain dL/dRp (+ dL/dqn) into the un-normalised pose quaternion
        // (F.normalize + build_rotation backward, slam_helpers.py:266-271)
        float vR[9];
        for (int k = 0; k < 9; ++k) vR[k] = (float)aR[k];
        float vq[4] = {0, 0, 0, 0};
        quat_to_R_vjp(pose->q, vR, vq);
        float inv = 1.0f / sqrtf(((pose->q[0] * pose->q[0] + pose->q[1] * pose->q[1]) + pose->q[2] * pose->q[2]) +
                                 pose->q[3] * pose->q[3]);
        double dot = aq[0] * xf.qn[0] + aq[1] * xf.qn[1] + aq[2] * xf.qn[2] + aq[3] * xf.qn[3];
        for (int k = 0; k < 4; ++k) v_pose[k] = (double)vq[k] + (aq[k] - dot * xf.qn[k]) * inv;
        for (int k = 0; k < 3; ++k) v_pose[4 + k] = at[k];
    }
}

// 10 * torch.median(|gt - d| * (gt > 0)) (splatam.py:552-553; torch.median = lower median)
extern "C" float orc_depth_outlier_threshold(int64_t P, const float* render_depth, const float* gt_depth, float factor) {
    std::vector<float> e(P);
    for (int64_t i = 0; i < P; ++i) e[i] = fabsf(gt_depth[i] - render_depth[i]) * (gt_depth[i] > 0.0f ? 1.0f : 0.0f);
    std::nth_element(e.begin(), e.begin() + (P - 1) / 2, e.end());
    return factor * e[(P - 1) / 2];
}

extern "C" void orc_loss_l1(const orc_camera* cam, int mode, const float* render, const float* alpha,
                            const float* gt_rgb, const float* gt_depth, float w_im, float w_depth, float sil_thres,
                            int use_sil, float outlier_thresh, float* v_render, double* losses) {
    const int64_t P = (int64_t)cam->width * cam->height;
    double sum_d = 0.0, sum_c = 0.0;
    int64_t n_d = 0;
    // pass 1: counts for the mean normalisers (mapping).  With ignore_outlier_depth_loss the mean runs over the
    // masked pixels (splatam.py:545-549, 556-559: |err|[mask].mean()).
    if (mode == 1)
        for (int64_t i = 0; i < P; ++i) {
            bool m = gt_depth[i] > 0.0f;
            if (outlier_thresh >= 0.0f) {
                float d = render[3 * P + i], unc = render[4 * P + i] - d * d;
                m = m && (fabsf(d - gt_depth[i]) < outlier_thresh) && !(std::isnan(d) || std::isnan(unc));
            }
            n_d += m;
        }
    const float inv_nd = (mode == 1) ? (n_d > 0 ? 1.0f / (float)n_d : 0.0f) : 1.0f;
    const float inv_nc = (mode == 1) ? 1.0f / (float)(3 * P) : 1.0f;
#pragma omp parallel for reduction(+ : sum_d, sum_c) schedule(static)
    for (int64_t i = 0; i < P; ++i) {
        float d = render[3 * P + i], d2 = render[4 * P + i];
        float unc = d2 - d * d;
        bool nan_ok = !(std::isnan(d) || std::isnan(unc)); // splatam.py:541
        bool m = (gt_depth[i] > 0.0f) && nan_ok;
        bool mc = true;
        float dd = d - gt_depth[i];
        const bool outl = outlier_thresh >= 0.0f;          // ignore_outlier_depth_loss, :551-554 (tracking)
        if (outl) m = m && (fabsf(dd) < outlier_thresh);
        if (mode == 0) {
            if (use_sil) m = m && (alpha[i] > sil_thres); // :551-552
            mc = (use_sil || outl) ? m : true;            // :563-568
        }
        float sgn_d = (dd > 0.0f) ? 1.0f : ((dd < 0.0f) ? -1.0f : 0.0f);
        if (m) sum_d += (double)fabsf(dd);
        v_render[3 * P + i] = m ? (w_depth * inv_nd) * sgn_d : 0.0f;
        v_render[4 * P + i] = 0.0f;
        for (int k = 0; k < 3; ++k) {
            float dc = render[k * P + i] - gt_rgb[k * P + i];
            float sg = (dc > 0.0f) ? 1.0f : ((dc < 0.0f) ? -1.0f : 0.0f);
            if (mc) sum_c += (double)fabsf(dc);
            v_render[k * P + i] = mc ? (w_im * inv_nc) * sg : 0.0f;
        }
    }
    double ld = sum_d * inv_nd, lc = sum_c * inv_nc;
    losses[0] = w_im * lc + w_depth * ld;
    losses[1] = lc;
    losses[2] = ld;
}

// calc_ssim / _ssim (python_ordinary/utils/slam_external.py:54-97): separable 11-tap Gaussian (sigma 1.5, float32
// weights normalised in float32), zero padding, on x, y, x*x, y*y, x*y; SSIM map; mean.  The backward is the
// analytic adjoint of the same expression (the reference gets it from autograd).
extern "C" double orc_ssim(int W, int H, const float* x, const float* y, float coef, float* v_x) {
    const int R = 5;
    float g[11];
    {
        float s = 0.f;
        for (int i = 0; i < 11; ++i) { g[i] = (float)std::exp(-(double)((i - R) * (i - R)) / (2.0 * 1.5 * 1.5)); s += g[i]; }
        for (int i = 0; i < 11; ++i) g[i] /= s;
    }
    const int64_t P = (int64_t)W * H;
    const double c1 = 0.01 * 0.01, c2 = 0.03 * 0.03;
    std::vector<double> tmp(5 * P), mom(5 * P), dm(3 * P);
    double total = 0.0;
    for (int c = 0; c < 3; ++c) {
        const float* X = x + c * P;
        const float* Y = y + c * P;
        // horizontal then vertical pass over the five moments
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < P; ++i) {
            const int yy = (int)(i / W), xx = (int)(i % W);
            double a[5] = {0, 0, 0, 0, 0};
            for (int k = -R; k <= R; ++k) {
                const int q = xx + k;
                if (q < 0 || q >= W) continue;
                const double w = g[k + R], vx = X[(int64_t)yy * W + q], vy = Y[(int64_t)yy * W + q];
                a[0] += w * vx; a[1] += w * vy; a[2] += w * vx * vx; a[3] += w * vy * vy; a[4] += w * vx * vy;
            }
            for (int m = 0; m < 5; ++m) tmp[m * P + i] = a[m];
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < P; ++i) {
            const int yy = (int)(i / W), xx = (int)(i % W);
            double a[5] = {0, 0, 0, 0, 0};
            for (int k = -R; k <= R; ++k) {
                const int q = yy + k;
                if (q < 0 || q >= H) continue;
                for (int m = 0; m < 5; ++m) a[m] += (double)g[k + R] * tmp[m * P + (int64_t)q * W + xx];
            }
            for (int m = 0; m < 5; ++m) mom[m * P + i] = a[m];
        }
        double sum = 0.0;
#pragma omp parallel for reduction(+ : sum) schedule(static)
        for (int64_t i = 0; i < P; ++i) {
            const double mx = mom[i], my = mom[P + i], exx = mom[2 * P + i], eyy = mom[3 * P + i], exy = mom[4 * P + i];
            const double A = 2 * mx * my + c1, B = 2 * (exy - mx * my) + c2;
            const double Cc = mx * mx + my * my + c1, D = (exx - mx * mx) + (eyy - my * my) + c2;
            const double S = A * B / (Cc * D);
            sum += S;
            dm[i] = 2 * my * (B - A) / (Cc * D) + 2 * mx * S * (1 / D - 1 / Cc);
            dm[P + i] = -S / D;
            dm[2 * P + i] = 2 * A / (Cc * D);
        }
        total += sum;
        if (!v_x) continue;
        // adjoint: the window is symmetric, so convolve the three derivative maps with it
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < P; ++i) {
            const int yy = (int)(i / W), xx = (int)(i % W);
            double a[3] = {0, 0, 0};
            for (int k = -R; k <= R; ++k) {
                const int q = xx + k;
                if (q < 0 || q >= W) continue;
                for (int m = 0; m < 3; ++m) a[m] += (double)g[k + R] * dm[m * P + (int64_t)yy * W + q];
            }
            for (int m = 0; m < 3; ++m) tmp[m * P + i] = a[m];
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < P; ++i) {
            const int yy = (int)(i / W), xx = (int)(i % W);
            double a[3] = {0, 0, 0};
            for (int k = -R; k <= R; ++k) {
                const int q = yy + k;
                if (q < 0 || q >= H) continue;
                for (int m = 0; m < 3; ++m) a[m] += (double)g[k + R] * tmp[m * P + (int64_t)q * W + xx];
            }
            const double d = a[0] + 2.0 * X[i] * a[1] + (double)Y[i] * a[2];
            v_x[c * P + i] += (float)((double)coef * d / (3.0 * (double)P));
        }
    }
    return total / (3.0 * (double)P);
}

// IsoGS geometric terms (python_ordinary/scripts/splatam.py: compute_flat_loss :42-54, get_loss :572-675,
// compute_iso_surface_loss_sampled :162-262): brute-force exact 16-NN (the reference: cdist + topk), density of the
// neighbours at the query centre, (D - target)^2, flat = mean min clamp(exp(ls), 1e-5).  Double arithmetic; the
// gradient is the analytic adjoint (the reference gets it from autograd).  out = {flat, iso, mean density};
// returns w_flat * flat + w_iso * iso.  g_* (all or none) are ACCUMULATED.
extern "C" double orc_iso_flat_loss(int64_t N, const float* means, const float* log_scales, const float* quats,
                                    const float* logit_opac, int Q, const int32_t* query_idx, float w_iso, float w_flat,
                                    float target, float* g_means, float* g_ls, float* g_quats, float* g_logit,
                                    double* out, int32_t* nbr_out) {
    const int K = 16;
    std::vector<double> gm(g_means ? 3 * N : 0, 0.0), gl(g_means ? 3 * N : 0, 0.0), gq(g_means ? 4 * N : 0, 0.0),
        go(g_means ? N : 0, 0.0);
    double iso = 0.0, dens = 0.0;
    for (int q = 0; q < Q; ++q) {
        const int64_t qi = query_idx ? query_idx[q] : q;
        std::vector<std::pair<double, int>> dist(N);
        for (int64_t j = 0; j < N; ++j) {
            double d2 = 0;
            for (int k = 0; k < 3; ++k) { double d = (double)means[3 * qi + k] - (double)means[3 * j + k]; d2 += d * d; }
            dist[j] = {d2, (int)j};
        }
        const int kk = (int)std::min<int64_t>(K, N);
        std::partial_sort(dist.begin(), dist.begin() + kk, dist.end());
        struct Nb { int j; double d[3], R[9], y[3], s[3], w[3], alpha, e, qn[4], inv; };
        std::vector<Nb> nb(kk);
        double D = 0;
        for (int t = 0; t < kk; ++t) {
            Nb& b = nb[t];
            b.j = dist[t].second;
            if (nbr_out) nbr_out[(int64_t)q * K + t] = b.j;
            const float* qq = quats + 4 * (int64_t)b.j;
            double n = std::sqrt((double)qq[0] * qq[0] + (double)qq[1] * qq[1] + (double)qq[2] * qq[2] + (double)qq[3] * qq[3]);
            b.inv = 1.0 / n;
            double w = qq[0] * b.inv, x = qq[1] * b.inv, y = qq[2] * b.inv, z = qq[3] * b.inv;
            b.qn[0] = w; b.qn[1] = x; b.qn[2] = y; b.qn[3] = z;
            double Rm[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
            for (int k = 0; k < 9; ++k) b.R[k] = Rm[k];
            for (int k = 0; k < 3; ++k) b.d[k] = (double)means[3 * qi + k] - (double)means[3 * (int64_t)b.j + k];
            double quad = 0;
            for (int k = 0; k < 3; ++k) {
                b.y[k] = b.R[k] * b.d[0] + b.R[3 + k] * b.d[1] + b.R[6 + k] * b.d[2];
                b.s[k] = std::max(std::exp((double)log_scales[3 * (int64_t)b.j + k]), 1e-5);
                b.w[k] = 1.0 / (b.s[k] * b.s[k] + 1e-8);
                quad += b.w[k] * b.y[k] * b.y[k];
            }
            b.alpha = 1.0 / (1.0 + std::exp(-(double)logit_opac[b.j]));
            b.e = std::exp(-0.5 * quad);
            D += b.alpha * b.e;
        }
        for (int t = kk; t < K && nbr_out; ++t) nbr_out[(int64_t)q * K + t] = -1;
        const double diff = D - (double)target;
        iso += diff * diff;
        dens += D;
        if (!g_means) continue;
        const double c = (double)w_iso * 2.0 * diff / (double)Q;
        for (const Nb& b : nb) {
            const double g = c * b.alpha * b.e * -0.5;
            const double u[3] = {b.w[0] * b.y[0], b.w[1] * b.y[1], b.w[2] * b.y[2]};
            double vR[9];
            for (int a = 0; a < 3; ++a) {
                const double dd = 2.0 * g * (b.R[3 * a] * u[0] + b.R[3 * a + 1] * u[1] + b.R[3 * a + 2] * u[2]);
                gm[3 * qi + a] += dd;
                gm[3 * (int64_t)b.j + a] -= dd;
                for (int k = 0; k < 3; ++k) vR[3 * a + k] = 2.0 * g * b.d[a] * u[k];
            }
            // adjoint of quaternion -> rotation through the normalisation
            const double w = b.qn[0], x = b.qn[1], y = b.qn[2], z = b.qn[3];
#define G(r, c) vR[(r) * 3 + (c)]
            double vq[4];
            vq[0] = 2 * (x * (G(2, 1) - G(1, 2)) + y * (G(0, 2) - G(2, 0)) + z * (G(1, 0) - G(0, 1)));
            vq[1] = 2 * (-2 * x * (G(1, 1) + G(2, 2)) + y * (G(1, 0) + G(0, 1)) + z * (G(2, 0) + G(0, 2)) + w * (G(2, 1) - G(1, 2)));
            vq[2] = 2 * (x * (G(1, 0) + G(0, 1)) - 2 * y * (G(0, 0) + G(2, 2)) + z * (G(2, 1) + G(1, 2)) + w * (G(0, 2) - G(2, 0)));
            vq[3] = 2 * (x * (G(2, 0) + G(0, 2)) + y * (G(2, 1) + G(1, 2)) - 2 * z * (G(0, 0) + G(1, 1)) + w * (G(1, 0) - G(0, 1)));
#undef G
            const double dot = vq[0] * w + vq[1] * x + vq[2] * y + vq[3] * z;
            for (int k = 0; k < 4; ++k) gq[4 * (int64_t)b.j + k] += (vq[k] - dot * b.qn[k]) * b.inv;
            for (int k = 0; k < 3; ++k)
                if (std::exp((double)log_scales[3 * (int64_t)b.j + k]) >= 1e-5)
                    gl[3 * (int64_t)b.j + k] += g * b.y[k] * b.y[k] * (-2.0 * b.s[k] * b.w[k] * b.w[k]) * b.s[k];
            go[b.j] += c * b.e * b.alpha * (1.0 - b.alpha);
        }
    }
    double flat = 0;
    for (int64_t i = 0; i < N; ++i) {
        double best = 1e300, raw = 0;
        int arg = 0;
        for (int k = 0; k < 3; ++k) {
            const double ex = std::exp((double)log_scales[3 * i + k]), sc = std::max(ex, 1e-5);
            if (sc < best) { best = sc; arg = k; raw = ex; }
        }
        flat += best;
        if (g_means && raw >= 1e-5) gl[3 * i + arg] += (double)w_flat * raw / (double)N;
    }
    flat /= (double)N;
    iso = Q > 0 ? iso / Q : 0.0;
    if (g_means) {
        for (int64_t i = 0; i < 3 * N; ++i) { g_means[i] += (float)gm[i]; g_ls[i] += (float)gl[i]; }
        for (int64_t i = 0; i < 4 * N; ++i) g_quats[i] += (float)gq[i];
        for (int64_t i = 0; i < N; ++i) g_logit[i] += (float)go[i];
    }
    out[0] = flat; out[1] = iso; out[2] = Q > 0 ? dens / Q : 0.0;
    return (double)w_flat * flat + (double)w_iso * iso;
}

extern "C" void orc_adam(int64_t n, float* p, const float* g, float* m, float* v, float lr, float b1, float b2,
                         float eps, int step) {
    // torch.optim.Adam (single tensor, no amsgrad / weight decay): step counted from 1
    const double bc1 = 1.0 - std::pow((double)b1, step);
    const double bc2 = 1.0 - std::pow((double)b2, step);
    const float step_size = (float)(lr / bc1);
    const float bc2_sqrt = (float)std::sqrt(bc2);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        float gi = g[i];
        float mi = m[i] + (gi - m[i]) * (1.0f - b1); // lerp form used by torch
        float vi = v[i] * b2 + (1.0f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = p[i] - step_size * (mi / denom);
    }
}

extern "C" int64_t orc_compact_rows(int64_t N, int width, const uint8_t* keep, const float* in, float* out) {
    int64_t j = 0;
    for (int64_t i = 0; i < N; ++i) {
        if (!keep[i]) continue;
        std::memcpy(out + j * width, in + i * width, sizeof(float) * width);
        ++j;
    }
    return j;
}
