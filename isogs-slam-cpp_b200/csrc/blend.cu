// blend.cu -- north_star (3) renderCUDA forward: per 16x16 tile alpha-blend of colour, depth,
// depth^2 and silhouette; north_star (4) backward: reverse per-pixel traversal with quad-reduced
// no-return atomics (RED).  Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188,
// RasterizeToPixels3DGSBwd.cu:15-278.
//
// Design: warp-autonomous, no CTA barrier.  A CTA is one tile, a warp owns an 8x4
// pixel block of it and walks the tile's depth-sorted list on its own, 32*NB Gaussians per batch:
// every lane stages NB Gaussians of the batch into the warp's private shared-memory slab and tests
// their {sigma <= L} footprints against the warp's EIGHT 2x2 pixel quads; eight ballots per 32
// Gaussians give every quad (4 lanes) its own hit mask (NB words).  The quads then consume their masks INDEPENDENTLY: in one warp
// iteration each quad blends a different Gaussian (its next hit), so SplaTAM-sized splats (about
// 7 px wide) keep ~70 % of the lanes on pixels they really touch instead of ~23 % with one
// Gaussian per warp iteration.  Backward: the 11 partial gradients are summed over the 4 lanes of
// a quad with a transposed 9-shuffle reduction and every lane issues 3 RED.ADD.F32.
//
// This file is compiled with FMA contraction ON; the sigma expression uses explicit
// round-to-nearest intrinsics and is the one the oracle evaluates, so the contribute/skip
// decision (and therefore `last`) is identical on CPU and GPU.
#include <cstdlib>

#include "common.cuh"

namespace isogs {

constexpr int BLOCK = TILE * TILE; // 256 threads, one per pixel
constexpr int NWARP = BLOCK / 32;
constexpr int RW = 8, RH = 4;      // pixel block per warp
constexpr unsigned FULL = 0xffffffffu;

#ifdef ISOGS_STATS
// debug build only (make STATS=1): occupancy-of-work counters read by scripts/blend_stats.py
__device__ unsigned long long g_stats[16];
#define STAT(i, n) do { const unsigned long long n_ = (unsigned long long)(n); if (lane == 0) atomicAdd(&g_stats[i], n_); } while (0)
#else
#define STAT(i, n) do { } while (0)
#endif

// Warp-private staging slab for one batch of 32*NB Gaussians (dynamic shared memory).
template <int NB, bool WITH_ID>
struct __align__(16) WarpStage {
    float4 g0[32 * NB]; // x, y, 0.5*conic_a, conic_b
    float4 g1[32 * NB]; // 0.5*conic_c, L (= sigma_max), opacity, c0
    float4 g2[32 * NB]; // c1, c2, c3 (depth), c4 (depth^2)
    int id[WITH_ID ? 32 * NB : 4];
};

__device__ __forceinline__ float sigma_of(const float4& g0, const float4& g1, float px, float py, float& dx,
                                          float& dy) {
    dx = __fsub_rn(g0.x, px);
    dy = __fsub_rn(g0.y, py);
    const float u = __fmaf_rn(g0.w, dy, __fmul_rn(g0.z, dx));
    return __fmaf_rn(u, dx, __fmul_rn(__fmul_rn(g1.x, dy), dy));
}

struct PixelMap {
    int px_i, py_i;
};
// lane -> pixel: quad q = lane>>2 (4 across x 2 down inside the 8x4 block), i = lane&3 inside the quad
__device__ __forceinline__ PixelMap pixel_of(int wid, int lane) {
    const int q = lane >> 2, i = lane & 3;
    PixelMap p;
    p.px_i = blockIdx.x * TILE + (wid & 1) * RW + (q & 3) * 2 + (i & 1);
    p.py_i = blockIdx.y * TILE + (wid >> 1) * RH + (q >> 2) * 2 + (i >> 1);
    return p;
}

// Hit mask of 32 staged Gaussians (one per lane) against THIS lane's quad.  The footprint test is
// conservative: axis-aligned half extents sqrt(2 L cov_xx), sqrt(2 L cov_yy) AND the bounding
// circle sqrt(2 L lambda_max), each padded for rounding, against the quad's pixel-centre square.
__device__ __forceinline__ uint32_t quad_hits(bool have, const float4& g0, const float4& g1, float bx0, float by0,
                                              int lane) {
    float hx = -1.f, hy = -1.f, R2 = -1.f;
    if (have) {
        const float a = 2.0f * g0.z, b = g0.w, c = 2.0f * g1.x;
        const float inv_det = __fdividef(1.0f, fmaf(a, c, -(b * b)));
        const float twoL = 2.0f * g1.y;
        const float cxx = c * inv_det, cyy = a * inv_det, cxy = b * inv_det;
        hx = fmaf(sqrtf(twoL * cxx), 1.001f, 0.01f);
        hy = fmaf(sqrtf(twoL * cyy), 1.001f, 0.01f);
        const float mid = 0.5f * (cxx + cyy), dif = 0.5f * (cxx - cyy);
        const float lam = mid + sqrtf(fmaf(dif, dif, cxy * cxy));
        const float R = fmaf(sqrtf(twoL * lam), 1.001f, 0.01f);
        R2 = R * R;
    }
    const float X = g0.x - bx0, Y = g0.y - by0; // relative to the block's first pixel centre
    float ddx[4], ddy[2];
#pragma unroll
    for (int c = 0; c < 4; ++c) ddx[c] = fmaxf(fmaxf((float)(2 * c) - X, X - (float)(2 * c + 1)), 0.f);
#pragma unroll
    for (int r = 0; r < 2; ++r) ddy[r] = fmaxf(fmaxf((float)(2 * r) - Y, Y - (float)(2 * r + 1)), 0.f);
    uint32_t bits = 0;
    const int myq = lane >> 2;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float ex = ddx[q & 3], ey = ddy[q >> 2];
        const bool hit = ex <= hx && ey <= hy && fmaf(ex, ex, ey * ey) <= R2;
        const uint32_t m = __ballot_sync(FULL, hit);
        if (myq == q) bits = m;
    }
    return bits;
}

// Next hit of this lane's quad: advances (j, bits) over the NB mask words; returns the slot or -1.
template <int NB>
__device__ __forceinline__ int next_hit(const uint32_t (&w)[NB], int& j, uint32_t& bits) {
    if (bits == 0u) {
#pragma unroll
        for (int k = 1; k < NB; ++k)
            if (bits == 0u && j < k) { j = k; bits = w[k]; }
    }
    if (bits == 0u) return -1;
    const int t = j * 32 + __ffs(bits) - 1;
    bits &= bits - 1u;
    return t;
}

template <int CH, int NB>
__global__ void __launch_bounds__(BLOCK)
blend_fwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                 int32_t* __restrict__ last) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using Stage = WarpStage<NB, false>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    Stage& st = reinterpret_cast<Stage*>(s_raw)[wid];
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(wid, lane);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const float bx0 = (float)(blockIdx.x * TILE + (wid & 1) * RW) + 0.5f;
    const float by0 = (float)(blockIdx.y * TILE + (wid >> 1) * RH) + 0.5f;
    const int start = ranges[tile], end = ranges[tile + 1];
    const int qshift = lane & 28;
    constexpr int B = 32 * NB;

    float T = 1.0f;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) acc[k] = 0.f;
    int last_idx = start - 1;
    bool done = !inside;

    for (int b = start; b < end; b += B) {
        if (__all_sync(FULL, done)) break;
        __syncwarp(); // the previous batch is fully consumed
        uint32_t w[NB];
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            const int slot = k * 32 + lane;
            const bool have = b + slot < end;
            float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
            if (have) {
                const float4* r = rec + 3 * (int64_t)__ldg(vals + b + slot);
                g0 = __ldg(r);
                g1 = __ldg(r + 1);
                st.g0[slot] = g0;
                st.g1[slot] = g1;
                st.g2[slot] = __ldg(r + 2);
            }
            w[k] = quad_hits(have, g0, g1, bx0, by0, lane);
        }
        __syncwarp();
        STAT(0, 1);
        int j = 0;
        uint32_t bits = w[0];
        bool qdone = ((__ballot_sync(FULL, !done) >> qshift) & 15u) == 0u;
        while (true) {
            const int t = qdone ? -1 : next_hit<NB>(w, j, bits);
            if (!__any_sync(FULL, t >= 0)) break;
            const int ts = max(t, 0);
            const float4 g0s = st.g0[ts];
            const float4 g1s = st.g1[ts];
            float dx, dy;
            const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
            STAT(1, 1);
            STAT(2, __popc(__ballot_sync(FULL, t >= 0)));
            STAT(3, __popc(__ballot_sync(FULL, t >= 0 && !done && sigma >= 0.f && sigma <= g1s.y)));
            bool newly = false;
            if (t >= 0 && !done && sigma >= 0.f && sigma <= g1s.y) {
                const float a = fminf(MAX_ALPHA, g1s.z * __expf(-sigma));
                const float nT = T * (1.0f - a);
                if (nT <= T_CUTOFF) {
                    done = true;
                    newly = true;
                } else {
                    const float vis = a * T;
                    const float4 g2 = st.g2[ts];
                    acc[0] = fmaf(g1s.w, vis, acc[0]);
                    acc[1] = fmaf(g2.x, vis, acc[1]);
                    acc[2] = fmaf(g2.y, vis, acc[2]);
                    acc[3] = fmaf(g2.z, vis, acc[3]);
                    if (CH == 5) acc[CH - 1] = fmaf(g2.w, vis, acc[CH - 1]);
                    last_idx = b + ts;
                    T = nT;
                }
            }
            if (__any_sync(FULL, newly)) qdone = ((__ballot_sync(FULL, !done) >> qshift) & 15u) == 0u;
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

// Sum 12 per-lane values over the 4 lanes of a quad with 9 shuffles: at each halving step a lane
// keeps one half of its values and trades the other half with its partner.  On return lane l
// holds in v[0..2] the quad totals of components 6*(l&1) + 3*((l>>1)&1) + {0,1,2}.
__device__ __forceinline__ void quad_reduce12(float (&v)[12], int lane) {
    {
        const bool up = lane & 1;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const float send = up ? v[k] : v[k + 6];
            const float keep = up ? v[k + 6] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 1);
        }
    }
    {
        const bool up = lane & 2;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float send = up ? v[k] : v[k + 3];
            const float keep = up ? v[k + 3] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 2);
        }
    }
}

__device__ __forceinline__ void red_v4(float* a, float x, float y, float z, float w) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(x), "f"(y), "f"(z), "f"(w)
                 : "memory");
}
__device__ __forceinline__ void red_v2(float* a, float x, float y) {
    asm volatile("red.relaxed.gpu.global.add.v2.f32 [%0], {%1, %2};" ::"l"(a), "f"(x), "f"(y) : "memory");
}

// RMODE: how the 11 partial gradients of a (quad, Gaussian) pair reach v_rec.
//   0  quad reduce (9 shuffles) + 3 scalar RED per lane
//   1  pair reduce (6 shuffles) + one v4 and one v2 RED per lane
//   2  no reduce: three v4 RED per contributing lane
template <int CH, int NB, int RMODE>
__global__ void __launch_bounds__(BLOCK)
blend_bwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                 const int32_t* __restrict__ last, const float* __restrict__ v_render,
                 const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using Stage = WarpStage<NB, true>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    Stage& st = reinterpret_cast<Stage*>(s_raw)[wid];
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(wid, lane);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const float bx0 = (float)(blockIdx.x * TILE + (wid & 1) * RW) + 0.5f;
    const float by0 = (float)(blockIdx.y * TILE + (wid >> 1) * RH) + 0.5f;
    const int start = ranges[tile];
    const int64_t P = (int64_t)cam.width * cam.height;
    const int64_t pix = inside ? (int64_t)pm.py_i * cam.width + pm.px_i : 0;
    const int qshift = lane & 28;
    constexpr int B = 32 * NB;

    const int my_last = inside ? last[pix] : start - 1;
    const int warp_last = __reduce_max_sync(FULL, my_last);
    if (warp_last < start) return; // nothing was blended into this warp's block
    int quad_last = max(my_last, __shfl_xor_sync(FULL, my_last, 1));
    quad_last = max(quad_last, __shfl_xor_sync(FULL, quad_last, 2));

    const float T_final = inside ? 1.0f - alpha[pix] : 1.0f;
    float T = T_final;
    float vr[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) vr[k] = inside ? v_render[k * P + pix] : 0.f;
    const float va = (inside && v_alpha) ? v_alpha[pix] : 0.f;
    float bg_dot = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) bg_dot += cam.bg[k] * vr[k];
    const float va_eff = T_final * (va - bg_dot); // (T_final * ra) * va - (T_final * ra) * bg_dot
    // sum_k buffer_k * vr_k: the colour accumulated BEHIND the current Gaussian enters the
    // gradient only through its dot product with the pixel's v_render, so one scalar carries it
    float bdot = 0.f;

    // component triple this lane owns after quad_reduce12
    const int comp0 = 6 * (lane & 1) + 3 * ((lane >> 1) & 1);
    const bool third = comp0 + 2 < 6 + CH;

    // batches run back to front; inside a batch slot t holds sorted index batch_end - t
    for (int batch_end = warp_last; batch_end >= start; batch_end -= B) {
        __syncwarp(); // the previous batch is fully consumed
        uint32_t w[NB];
        // slots in front of t_first lie behind this quad's last contributor
        const int t_first = batch_end - quad_last;
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            const int slot = k * 32 + lane;
            const bool have = batch_end - slot >= start;
            float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
            if (have) {
                const int g = __ldg(vals + batch_end - slot);
                const float4* r = rec + 3 * (int64_t)g;
                g0 = __ldg(r);
                g1 = __ldg(r + 1);
                st.g0[slot] = g0;
                st.g1[slot] = g1;
                st.g2[slot] = __ldg(r + 2);
                st.id[slot] = g;
            }
            uint32_t m = quad_hits(have, g0, g1, bx0, by0, lane);
            const int cut = t_first - 32 * k;
            if (cut > 0) m = (cut >= 32) ? 0u : (m & ~((1u << cut) - 1u));
            w[k] = m;
        }
        __syncwarp();
        STAT(8, 1);
        int j = 0;
        uint32_t bits = w[0];
        while (true) {
            const int t = next_hit<NB>(w, j, bits);
            if (!__any_sync(FULL, t >= 0)) break;
            const int ts = max(t, 0);
            const float4 g0s = st.g0[ts];
            const float4 g1s = st.g1[ts];
            float dx, dy;
            const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
            const bool valid = t >= 0 && (batch_end - ts <= my_last) && sigma >= 0.f && sigma <= g1s.y;
            const uint32_t vb = __ballot_sync(FULL, valid);
            STAT(9, 1);
            STAT(10, __popc(__ballot_sync(FULL, t >= 0)));
            STAT(11, __popc(vb));
            if (vb == 0u) continue;
            STAT(12, 1);
            // branch-free body: an invalid lane runs with sigma = 0 and contributes exact zeros
            const float4 g2 = st.g2[ts];
            const float opac = g1s.z;
            const float vis = __expf(valid ? -sigma : 0.f);
            const float ov = opac * vis;
            const float a = fminf(MAX_ALPHA, ov);
            const float ra = __fdividef(1.0f, 1.0f - a);
            const float Tn = T * ra; // transmittance in front of this Gaussian
            T = valid ? Tn : T;
            const float fac = valid ? a * Tn : 0.f;
            float cdot = g1s.w * vr[0];
            cdot = fmaf(g2.x, vr[1], cdot);
            cdot = fmaf(g2.y, vr[2], cdot);
            cdot = fmaf(g2.z, vr[3], cdot);
            if (CH == 5) cdot = fmaf(g2.w, vr[CH - 1], cdot);
            const float v_a = fmaf(Tn, cdot, ra * (va_eff - bdot));
            bdot = fmaf(fac, cdot, bdot);
            const float u = (valid && ov <= MAX_ALPHA) ? vis * v_a : 0.f;
            const float v_sigma = -opac * u;
            const float ca = 2.0f * g0s.z, cb = g0s.w, cc = 2.0f * g1s.x;
            const float wx = v_sigma * dx, wy = v_sigma * dy;
            float v[12];
            v[0] = fmaf(ca, wx, cb * wy);
            v[1] = fmaf(cb, wx, cc * wy);
            v[2] = 0.5f * wx * dx;
            v[3] = wx * dy;
            v[4] = 0.5f * wy * dy;
            v[5] = u;
#pragma unroll
            for (int k = 0; k < CH; ++k) v[6 + k] = fac * vr[k];
#pragma unroll
            for (int k = 6 + CH; k < 12; ++k) v[k] = 0.f;
            if (RMODE == 0) {
                quad_reduce12(v, lane);
                if ((vb >> qshift) & 15u) {
                    float* dst = v_rec + (int64_t)st.id[ts] * REC + comp0;
                    atomicAdd(dst, v[0]);
                    atomicAdd(dst + 1, v[1]);
                    if (third) atomicAdd(dst + 2, v[2]);
                }
            } else if (RMODE == 1) {
                const bool up = lane & 1;
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const float send = up ? v[k] : v[k + 6];
                    const float keep = up ? v[k + 6] : v[k];
                    v[k] = keep + __shfl_xor_sync(FULL, send, 1);
                }
                if ((vb >> (lane & 30)) & 3u) {
                    float* dst = v_rec + (int64_t)st.id[ts] * REC;
                    if (!up) {
                        red_v4(dst, v[0], v[1], v[2], v[3]);
                        red_v2(dst + 4, v[4], v[5]);
                    } else {
                        red_v2(dst + 6, v[0], v[1]);
                        red_v4(dst + 8, v[2], v[3], v[4], v[5]);
                    }
                }
            } else if (RMODE == 3) {
                // quad reduce that leaves lane 0 with comps 0-3, lane 1 with 4-7, lane 2 with 8-9 and
                // lane 3 with 10-11, so a quad issues two v4 and one or two v2 REDs (REDG throughput
                // is per lane-op, and a v4 lane-op costs ~1.7x a scalar one)
                const bool up = lane & 1, hi = lane & 2;
                constexpr int A[6] = {0, 1, 2, 3, 8, 9}, Bc[6] = {4, 5, 6, 7, 10, 11};
                float w6[6];
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const float send = up ? v[A[k]] : v[Bc[k]];
                    const float keep = up ? v[Bc[k]] : v[A[k]];
                    w6[k] = keep + __shfl_xor_sync(FULL, send, 1);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float send = hi ? w6[k] : (k < 2 ? w6[4 + k] : 0.f);
                    const float r = __shfl_xor_sync(FULL, send, 2);
                    if (k < 2) w6[4 + k] += hi ? r : 0.f;
                    w6[k] += hi ? 0.f : r;
                }
                if ((vb >> qshift) & 15u) {
                    float* dst = v_rec + (int64_t)st.id[ts] * REC;
                    if (!hi) red_v4(dst + 4 * (lane & 1), w6[0], w6[1], w6[2], w6[3]);
                    else if (CH == 5 || !up) red_v2(dst + 8 + 2 * (lane & 1), w6[4], w6[5]);
                }
            } else {
                if (valid) {
                    float* dst = v_rec + (int64_t)st.id[ts] * REC;
                    red_v4(dst, v[0], v[1], v[2], v[3]);
                    red_v4(dst + 4, v[4], v[5], v[6], v[7]);
                    red_v4(dst + 8, v[8], v[9], v[10], 0.f);
                }
            }
        }
    }
}

} // namespace isogs

using namespace isogs;

namespace {
constexpr int NB_BWD = 4;

template <typename K>
int set_smem(K kernel, size_t bytes) {
    return check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes),
                      "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
}
} // namespace

#ifdef ISOGS_STATS
extern "C" int isogs_debug_stats(unsigned long long* out16, int reset) {
    if (out16) cudaMemcpyFromSymbol(out16, g_stats, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_stats, z, sizeof(z)); }
    return 0;
}
#endif

extern "C" int isogs_blend_fwd(const isogs_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                               const int32_t* tile_ranges, float* render, float* alpha, int32_t* last,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_fwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    dim3 grid(dc.tw, dc.th);
    static const int nb = [] { const char* e = getenv("ISOGS_NBF"); return e ? atoi(e) : 1; }();
#define ISOGS_LAUNCH_FWD(CHN, NBF)                                                                          \
    do {                                                                                                   \
        constexpr size_t smem = sizeof(WarpStage<NBF, false>) * NWARP;                                     \
        static thread_local int attr_dev = -1;                                                             \
        int dev = 0;                                                                                       \
        ISOGS_CUDA(cudaGetDevice(&dev));                                                                   \
        if (attr_dev != dev) {                                                                             \
            int rc = set_smem(blend_fwd_kernel<CHN, NBF>, smem);                                           \
            if (rc != 0) return rc;                                                                        \
            attr_dev = dev;                                                                                \
        }                                                                                                  \
        blend_fwd_kernel<CHN, NBF><<<grid, BLOCK, smem, stream>>>(dc, (const float4*)rec, vals_sorted,     \
                                                                   tile_ranges, render, alpha, last);      \
    } while (0)
    if (ch == 4) {
        if (nb == 4) ISOGS_LAUNCH_FWD(4, 4); else if (nb == 2) ISOGS_LAUNCH_FWD(4, 2); else ISOGS_LAUNCH_FWD(4, 1);
    } else {
        if (nb == 4) ISOGS_LAUNCH_FWD(5, 4); else if (nb == 2) ISOGS_LAUNCH_FWD(5, 2); else ISOGS_LAUNCH_FWD(5, 1);
    }
#undef ISOGS_LAUNCH_FWD
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
    constexpr size_t smem = sizeof(WarpStage<NB_BWD, true>) * NWARP;
    static const int rmode = [] { const char* e = getenv("ISOGS_RED"); return e ? atoi(e) : 0; }();
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        int rc = set_smem(blend_bwd_kernel<4, NB_BWD, 0>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<5, NB_BWD, 0>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<4, NB_BWD, 1>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<5, NB_BWD, 1>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<4, NB_BWD, 2>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<5, NB_BWD, 2>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<4, NB_BWD, 3>, smem);
        if (rc == 0) rc = set_smem(blend_bwd_kernel<5, NB_BWD, 3>, smem);
        if (rc != 0) return rc;
        attr_dev = dev;
    }
#define ISOGS_LAUNCH_BWD(CHN, RM)                                                                              \
    blend_bwd_kernel<CHN, NB_BWD, RM><<<grid, BLOCK, smem, stream>>>(dc, (const float4*)rec, vals_sorted, tile_ranges, \
                                                                      alpha, last, v_render, v_alpha, v_rec)
    if (ch == 4) {
        if (rmode == 1) ISOGS_LAUNCH_BWD(4, 1); else if (rmode == 2) ISOGS_LAUNCH_BWD(4, 2); else if (rmode == 3) ISOGS_LAUNCH_BWD(4, 3); else ISOGS_LAUNCH_BWD(4, 0);
    } else {
        if (rmode == 1) ISOGS_LAUNCH_BWD(5, 1); else if (rmode == 2) ISOGS_LAUNCH_BWD(5, 2); else if (rmode == 3) ISOGS_LAUNCH_BWD(5, 3); else ISOGS_LAUNCH_BWD(5, 0);
    }
#undef ISOGS_LAUNCH_BWD
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
