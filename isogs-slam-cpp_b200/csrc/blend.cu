// blend.cu -- north_star (3) renderCUDA forward: per 16x16 tile alpha-blend of colour, depth,
// depth^2 and silhouette; north_star (4) backward: reverse per-pixel traversal with quad-reduced
// no-return atomics (RED).  Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188,
// RasterizeToPixels3DGSBwd.cu:15-278.
//
// Design: warp-autonomous, no CTA barrier.  A CTA is one tile, a warp owns an 8x4
// pixel block of it and walks the tile's depth-sorted list on its own, 32 Gaussians per batch:
// lane L stages Gaussian L of the batch into the warp's private shared-memory slab and tests its
// {sigma <= L} footprint against the warp's EIGHT 2x2 pixel quads; eight ballots give every quad
// (4 lanes) its own hit mask.  The quads then consume their masks INDEPENDENTLY: in one warp
// iteration each quad blends a different Gaussian (its next hit), so SplaTAM-sized splats (about
// 7 px wide) keep ~70 % of the lanes on pixels they really touch instead of ~23 % with one
// Gaussian per warp iteration.  Backward: the 11 partial gradients are summed over the 4 lanes of
// a quad with a transposed 9-shuffle reduction and every lane issues 3 RED.ADD.F32.
//
// This file is compiled with FMA contraction ON; the sigma expression uses explicit
// round-to-nearest intrinsics and is the one the oracle evaluates, so the contribute/skip
// decision (and therefore `last`) is identical on CPU and GPU.
#include "common.cuh"

namespace isogs {

constexpr int BLOCK = TILE * TILE; // 256 threads, one per pixel
constexpr int NWARP = BLOCK / 32;
constexpr int RW = 8, RH = 4;      // pixel block per warp
constexpr unsigned FULL = 0xffffffffu;

struct __align__(16) WarpStage {
    float4 g0[32]; // x, y, 0.5*conic_a, conic_b
    float4 g1[32]; // 0.5*conic_c, L (= sigma_max), opacity, c0
    float4 g2[32]; // c1, c2, c3 (depth), c4 (depth^2)
    int id[32];
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

// Hit mask of the warp's 32 staged Gaussians against THIS lane's quad.  The footprint test is
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

template <int CH>
__global__ void __launch_bounds__(BLOCK)
blend_fwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                 int32_t* __restrict__ last) {
    __shared__ WarpStage s_stage[NWARP];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    WarpStage& st = s_stage[wid];
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(wid, lane);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const float bx0 = (float)(blockIdx.x * TILE + (wid & 1) * RW) + 0.5f;
    const float by0 = (float)(blockIdx.y * TILE + (wid >> 1) * RH) + 0.5f;
    const int start = ranges[tile], end = ranges[tile + 1];
    const int qshift = lane & 28;

    float T = 1.0f;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) acc[k] = 0.f;
    int last_idx = start - 1;
    bool done = !inside;

    int g_next = (start + lane < end) ? __ldg(vals + start + lane) : 0;
    for (int b = start; b < end; b += 32) {
        if (__all_sync(FULL, done)) break;
        const bool have = b + lane < end;
        const int g = g_next;
        if (b + 32 + lane < end) g_next = __ldg(vals + b + 32 + lane);
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        __syncwarp(); // the previous batch is fully consumed
        if (have) {
            const float4* r = rec + 3 * (int64_t)g;
            g0 = __ldg(r);
            g1 = __ldg(r + 1);
            st.g0[lane] = g0;
            st.g1[lane] = g1;
            st.g2[lane] = __ldg(r + 2);
        }
        uint32_t bits = quad_hits(have, g0, g1, bx0, by0, lane);
        __syncwarp();
        while (true) {
            const uint32_t live = __ballot_sync(FULL, !done);
            if (((live >> qshift) & 15u) == 0u) bits = 0; // the whole quad is saturated
            if (!__any_sync(FULL, bits != 0)) break;
            const bool act = bits != 0;
            const int t = act ? (__ffs(bits) - 1) : 0;
            bits &= bits - 1;
            const float4 g0s = st.g0[t];
            const float4 g1s = st.g1[t];
            float dx, dy;
            const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
            if (act && !done && sigma >= 0.f && sigma <= g1s.y) {
                const float a = fminf(MAX_ALPHA, g1s.z * __expf(-sigma));
                const float nT = T * (1.0f - a);
                if (nT <= T_CUTOFF) {
                    done = true;
                } else {
                    const float vis = a * T;
                    const float4 g2 = st.g2[t];
                    acc[0] = fmaf(g1s.w, vis, acc[0]);
                    acc[1] = fmaf(g2.x, vis, acc[1]);
                    acc[2] = fmaf(g2.y, vis, acc[2]);
                    acc[3] = fmaf(g2.z, vis, acc[3]);
                    if (CH == 5) acc[CH - 1] = fmaf(g2.w, vis, acc[CH - 1]);
                    last_idx = b + t;
                    T = nT;
                }
            }
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

template <int CH>
__global__ void __launch_bounds__(BLOCK)
blend_bwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                 const int32_t* __restrict__ last, const float* __restrict__ v_render,
                 const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    __shared__ WarpStage s_stage[NWARP];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    WarpStage& st = s_stage[wid];
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

    const int my_last = inside ? last[pix] : start - 1;
    const int warp_last = __reduce_max_sync(FULL, my_last);
    if (warp_last < start) return; // nothing was blended into this warp's block
    int quad_last = max(my_last, __shfl_xor_sync(FULL, my_last, 1));
    quad_last = max(quad_last, __shfl_xor_sync(FULL, quad_last, 2));

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

    // component triple this lane owns after quad_reduce12
    const int comp0 = 6 * (lane & 1) + 3 * ((lane >> 1) & 1);

    // batches run back to front; inside a batch slot t holds sorted index batch_end - t
    int g_next = (warp_last - lane >= start) ? __ldg(vals + warp_last - lane) : 0;
    for (int batch_end = warp_last; batch_end >= start; batch_end -= 32) {
        const bool have = batch_end - lane >= start;
        const int g = g_next;
        if (batch_end - 32 - lane >= start) g_next = __ldg(vals + batch_end - 32 - lane);
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        __syncwarp(); // the previous batch is fully consumed
        if (have) {
            const float4* r = rec + 3 * (int64_t)g;
            g0 = __ldg(r);
            g1 = __ldg(r + 1);
            st.g0[lane] = g0;
            st.g1[lane] = g1;
            st.g2[lane] = __ldg(r + 2);
            st.id[lane] = g;
        }
        uint32_t bits = quad_hits(have, g0, g1, bx0, by0, lane);
        // slots in front of t_first lie behind every pixel of this quad's last contributor
        const int t_first = batch_end - quad_last;
        if (t_first > 0) bits = (t_first >= 32) ? 0u : (bits & ~((1u << t_first) - 1u));
        __syncwarp();
        while (__any_sync(FULL, bits != 0)) {
            const bool act = bits != 0;
            const int t = act ? (__ffs(bits) - 1) : 0;
            bits &= bits - 1;
            const float4 g0s = st.g0[t];
            const float4 g1s = st.g1[t];
            float dx, dy;
            const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
            const bool valid = act && (batch_end - t <= my_last) && sigma >= 0.f && sigma <= g1s.y;
            const uint32_t vb = __ballot_sync(FULL, valid);
            if (vb == 0u) continue;
            float v[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) v[k] = 0.f;
            if (valid) {
                const float4 g2 = st.g2[t];
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
                    const float wx = v_sigma * dx, wy = v_sigma * dy;
                    v[0] = fmaf(ca, wx, cb * wy);
                    v[1] = fmaf(cb, wx, cc * wy);
                    v[2] = 0.5f * wx * dx;
                    v[3] = wx * dy;
                    v[4] = 0.5f * wy * dy;
                    v[5] = vis * v_a;
                }
            }
            quad_reduce12(v, lane);
            if ((vb >> qshift) & 15u) {
                float* dst = v_rec + (int64_t)st.id[t] * REC + comp0;
                atomicAdd(dst, v[0]);
                atomicAdd(dst + 1, v[1]);
                if (comp0 + 2 < 6 + CH) atomicAdd(dst + 2, v[2]);
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
