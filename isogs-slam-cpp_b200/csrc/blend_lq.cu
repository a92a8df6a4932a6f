// blend_lq.cu -- north_star (3) renderCUDA forward (per 16x16 tile alpha-blend of colour, depth, depth^2 and silhouette)
// and (4) backward (reverse per-pixel traversal, vector no-return atomics): the "lane-per-quad" kernels of
// DESIGN.md 3.1 / 3.1b, and the C entry points isogs_blend_fwd / isogs_blend_bwd / isogs_blend_bwd_geom.
// Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188, RasterizeToPixels3DGSBwd.cu:15-278.
//
// Warp-autonomous ring walk (no CTA barrier): a LANE owns a whole 2x2 pixel quad (its four pixels
// live in registers and are processed one after the other) and a warp owns a 16x8 half tile = 32 quads; the warp
// walks the tile's depth-sorted list through a private shared-memory ring of NW words x 32 staged Gaussians with
// per-quad hit masks and independent per-quad cursors.
// In one warp iteration every lane blends a different Gaussian (its quad's next hit) into its four
// pixels.  Backward: the per-quad partial gradients are summed INSIDE the lane -- no shuffle reduction, no
// select network -- and one quad's 11 gradient components leave as three 16-byte vector REDs
// (red.global.add.v4.f32) into the 48-byte gradient record instead of 12 scalar REDs from four lanes.  The
// cursor / ring bookkeeping (about 45 instructions per warp iteration) is shared by up to 32 (quad, Gaussian)
// pairs instead of 8.
//
// The file is compiled with FMA contraction ON; the sigma expression uses explicit round-to-nearest intrinsics and is the
// one the oracle evaluates, so the contribute / skip decision of every pixel (and therefore `last`) is identical on CPU
// and GPU.  Measured alternatives that did NOT pay off are recorded in DESIGN.md 3.1c (four lanes per quad, rings of
// 4 words, scalar REDs, compacted rings, one 16-word ring per tile filled by cp.async.bulk / cp.async on mbarriers).

#include "common.cuh"

namespace isogs {
namespace lq {

constexpr unsigned FULL = 0xffffffffu;
constexpr int WARPS = 2;   // half tiles (16x8 pixels, one warp each) per tile = per CTA
constexpr int THREADS = 32 * WARPS;
constexpr int NW_DEFAULT = 8; // ring words per warp (256 staged Gaussians, 14 KB): 7-8 CTAs per SM
// (quad, Gaussian) pairs a lane evaluates per warp iteration (both from its current ring word): two in the backward
// amortise the cursor / retire bookkeeping and interleave two bodies (0.388 -> 0.379 ms, 124 registers); in the
// lighter forward the second slot idles too often and costs more than it saves (0.267 -> 0.305 ms)
constexpr int HITS_FWD = 1, HITS_BWD = 2;
constexpr int FAR = 0x01000000;

template <int NW, bool WITH_ID = true>
struct __align__(16) Ring {
    float4 g0[32 * NW]; // x, y, 0.5*conic_a, conic_b
    float4 g1[32 * NW]; // 0.5*conic_c, L (= sigma_max), opacity, c0
    float4 g2[32 * NW]; // c1, c2, c3 (depth), c4 (depth^2)
    uint32_t hits[32 * NW]; // [ring word][quad]: which of the word's 32 Gaussians can touch the quad
    int id[WITH_ID ? 32 * NW : 4];
};

struct Pf {
    float4 g0, g1, g2;
    int id;
    bool have;
};

// list entry e is sorted index first + DIR * e (backward walk: DIR = -1)
template <int DIR = -1>
__device__ __forceinline__ void prefetch(Pf& pf, int word, int first, int n_entries, const float4* __restrict__ rec,
                                         const int32_t* __restrict__ vals, int lane) {
    const int e = 32 * word + lane;
    const int id = (e < n_entries) ? __ldg(vals + first + DIR * e) : -1;
    pf.have = id >= 0;
    pf.id = pf.have ? id : 0;
    pf.g0 = make_float4(0.f, 0.f, 0.f, 0.f);
    pf.g1 = pf.g0;
    pf.g2 = pf.g0;
    if (pf.have) {
        const float4* r = rec + 3 * (int64_t)id;
        pf.g0 = __ldg(r);
        pf.g1 = __ldg(r + 1);
        pf.g2 = __ldg(r + 2);
    }
}

// Hit masks of the 32 Gaussians of one ring word (lane L holds Gaussian L's g0 / g1) against the warp's 32 quads,
// written to st.hits[wsl + q].  Footprint test as in blend.cu::quad_hits (squared half extents + Gershgorin circle,
// padded, conservative).
template <int NW, bool WITH_ID>
__device__ __forceinline__ void publish_hits(Ring<NW, WITH_ID>& st, bool have, const float4& g0, const float4& g1,
                                             unsigned wsl, float bx0, float by0, int lane) {
    float hx2 = -1.f, hy2 = -1.f, R2 = -1.f;
    if (have) {
        const float a = 2.0f * g0.z, b = g0.w, c = 2.0f * g1.x;
        const float inv_det = __fdividef(1.0f, fmaf(a, c, -(b * b)));
        const float twoL = 2.0f * g1.y * inv_det;
        const float cxx = c * twoL, cyy = a * twoL, cxy = fabsf(b * twoL);
        hx2 = fmaf(cxx, 1.01f, 0.05f);
        hy2 = fmaf(cyy, 1.01f, 0.05f);
        R2 = fmaf(fmaxf(cxx, cyy) + cxy, 1.01f, 0.05f);
    }
    const float X = g0.x - bx0, Y = g0.y - by0; // relative to the region's first pixel centre
    // per quad row r the three tests (x extent, y extent, circle) collapse into one threshold on ex2:
    // ex2 <= min(hx2, R2 - ey2[r]) if ey2[r] <= hy2, never otherwise
    float ex2[8], thr[4];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float e = fmaxf(fmaxf((float)(2 * c) - X, X - (float)(2 * c + 1)), 0.f);
        ex2[c] = e * e;
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const float e = fmaxf(fmaxf((float)(2 * r) - Y, Y - (float)(2 * r + 1)), 0.f);
        const float e2 = e * e;
        thr[r] = (e2 <= hy2) ? fminf(hx2, R2 - e2) : -1.f;
    }
#pragma unroll
    for (int q = 0; q < 32; ++q) {
        const uint32_t m = __ballot_sync(FULL, ex2[q & 7] <= thr[q >> 3]);
        if (lane == 0) st.hits[wsl + q] = m;
    }
}

// Commit the prefetched word into ring slot word % NW and publish its hit masks.
template <int NW, bool WITH_ID>
__device__ __forceinline__ void commit(Ring<NW, WITH_ID>& st, const Pf& pf, int word, float bx0, float by0, int lane) {
    const unsigned wsl = ((unsigned)word & (NW - 1)) * 32u;
    const unsigned slot = wsl + lane;
    __syncwarp();
    if (pf.have) {
        st.g0[slot] = pf.g0;
        st.g1[slot] = pf.g1;
        st.g2[slot] = pf.g2;
        if (WITH_ID) st.id[slot] = pf.id;
    }
    publish_hits<NW, WITH_ID>(st, pf.have, pf.g0, pf.g1, wsl, bx0, by0, lane);
    __syncwarp();
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}

// Initial fill of the ring (words 0 .. min(NW, n_words) - 1) with every load in flight at once: the NW ids of a lane
// first, then 3 x 16 B asynchronous copies per entry straight into the ring -- two global round trips instead of the
// 2 * NW a word-by-word fill serialises at the start of every warp.  Returns the number of words staged.
template <int NW, bool WITH_ID, int DIR>
__device__ __forceinline__ int initial_fill(Ring<NW, WITH_ID>& st, int first, int n_entries, int n_words,
                                            const float4* __restrict__ rec, const int32_t* __restrict__ vals, float bx0,
                                            float by0, int lane) {
    const int nfill = min(NW, n_words);
    int ids[NW];
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const int e = 32 * w + lane;
        ids[w] = (e < n_entries) ? __ldg(vals + first + DIR * e) : -1;
    }
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        if (ids[w] >= 0) {
            const float4* r = rec + 3 * (int64_t)ids[w];
            const unsigned slot = 32u * w + lane;
            cp_async16(&st.g0[slot], r);
            cp_async16(&st.g1[slot], r + 1);
            cp_async16(&st.g2[slot], r + 2);
            if (WITH_ID) st.id[slot] = ids[w];
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
#pragma unroll 1
    for (int w = 0; w < nfill; ++w) {
        const unsigned slot = 32u * w + lane;
        const bool have = 32 * w + lane < n_entries;
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        if (have) {
            g0 = st.g0[slot];
            g1 = st.g1[slot];
        }
        publish_hits<NW, WITH_ID>(st, have, g0, g1, 32u * w, bx0, by0, lane);
    }
    __syncwarp();
    return nfill;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// sigma in [0, L]  <=>  bits(sigma) < bits(L) + 1 as UNSIGNED integers (negative and NaN sigmas have the sign bit set
// or exceed every finite L; a negative L admits nothing): ONE integer compare per pixel instead of two float compares
// and a predicate AND.  on = false (the lane has no hit this iteration) admits nothing either.  (-0.0 cannot come out
// of the sigma expression: its last addend is a square, +0 at least, and x + (+0) rounds to +0.)
__device__ __forceinline__ uint32_t sigma_limit(float L, bool on) {
    return (on && L >= 0.f) ? __float_as_uint(L) + 1u : 0u;
}

// one (quad, Gaussian) pair being evaluated by a lane
struct Hit {
    float4 g0, g1;
    float dx[2], dy[2], sigma[4];
    unsigned slot;
    int e;
    bool valid[4], any;
};

// GEOM = true: tracking -- only the gradient components the POSE depends on (x, y, conic, z, z^2) are produced.
// HITS: (quad, Gaussian) pairs a lane evaluates per warp iteration.
template <int CH, int NW, bool GEOM, int HITS>
__global__ void __launch_bounds__(THREADS, 8)
blend_bwd_lq_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                    const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                    const int32_t* __restrict__ last, const float* __restrict__ v_render,
                    const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    using R = Ring<NW, true>;
    R& st = reinterpret_cast<R*>(s_raw)[wid];
    // half tile h = 2 * tile_y + half; a CTA holds blockDim.x / 32 consecutive halves of one tile column
    const int hidx = blockIdx.y * (blockDim.x >> 5) + wid, ty = hidx >> 1, half = hidx & 1;
    const int tile = ty * cam.tw + blockIdx.x;
    const int qx = lane & 7, qy = lane >> 3;
    const int x0 = blockIdx.x * TILE + 2 * qx, y0 = ty * TILE + 8 * half + 2 * qy;
    const float bx0 = (float)(blockIdx.x * TILE) + 0.5f, by0 = (float)(ty * TILE + 8 * half) + 0.5f;
    const float pxa = (float)x0 + 0.5f, pxb = (float)x0 + 1.5f, pya = (float)y0 + 0.5f, pyb = (float)y0 + 1.5f;
    const int start = ranges[tile];
    const int64_t P = (int64_t)cam.width * cam.height;

    int first[4]; // first list entry (back to front) that contributed to pixel p
    float T[4], bdot[4], va_eff[4], vr[4][CH];
    int quad_last = start - 1;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int xi = x0 + (p & 1), yi = y0 + (p >> 1);
        const bool inside = xi < cam.width && yi < cam.height;
        const int64_t pix = inside ? (int64_t)yi * cam.width + xi : 0;
        const int my_last = inside ? last[pix] : start - 1;
        first[p] = my_last; // turned into an entry index below
        quad_last = max(quad_last, my_last);
        const float T_final = inside ? 1.0f - alpha[pix] : 1.0f;
        T[p] = T_final;
#pragma unroll
        for (int k = 0; k < CH; ++k) vr[p][k] = inside ? v_render[k * P + pix] : 0.f;
        const float va = (inside && v_alpha) ? v_alpha[pix] : 0.f;
        float bg_dot = 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) bg_dot += cam.bg[k] * vr[p][k];
        va_eff[p] = T_final * (va - bg_dot);
        bdot[p] = 0.f;
    }
    const int warp_last = __reduce_max_sync(FULL, quad_last);
    if (warp_last < start) return; // nothing was blended into this warp's half tile
    const int n_entries = warp_last - start + 1;
    const int n_words = (n_entries + 31) >> 5;
    const int e_first = warp_last - quad_last;
#pragma unroll
    for (int p = 0; p < 4; ++p) first[p] = warp_last - first[p];

    int staged = 0, base = 0; // warp-uniform: words [base, staged) are live in the ring
    staged = initial_fill<NW, true, -1>(st, warp_last, n_entries, n_words, rec, vals, bx0, by0, lane);
    Pf pf;
    prefetch(pf, staged, warp_last, n_entries, rec, vals, lane);
    static_assert((NW & (NW - 1)) == 0, "NW must be a power of two");
    const uint32_t* hq = st.hits + lane; // this quad's hit word of ring slot base wsl is hq[wsl]
    auto fetch = [&](int w, unsigned wsl_) -> uint32_t {
        const uint32_t m = hq[wsl_];
        const int cut = e_first - 32 * w; // entries in front of e_first lie behind this quad's last contributor
        return cut <= 0 ? m : (cut >= 32 ? 0u : (m & ~((1u << cut) - 1u)));
    };
    int jq = min(e_first >> 5, staged - 1);
    unsigned wsl = ((unsigned)jq & (NW - 1)) * 32u;
    uint32_t bits = fetch(jq, wsl);
    if (quad_last < start) { jq = FAR; bits = 0u; wsl = 0u; }
    while (true) {
        if (bits == 0u && jq + 1 < staged) {
            ++jq;
            wsl = (wsl + 32u) & (32u * NW - 1u);
            bits = fetch(jq, wsl);
        }
        const bool act = bits != 0u;
        const int wmin = __reduce_min_sync(FULL, act ? jq : jq + 1);
        if (wmin >= n_words) break;
        // up to HITS hits of this quad (from its current ring word) per warp iteration: the cursor / retire bookkeeping
        // is amortised over them and their arithmetic interleaves (the per-pixel T / bdot recurrences chain hit b
        // behind hit a, everything else is independent)
        Hit h[HITS];
#pragma unroll
        for (int k = 0; k < HITS; ++k) {
            const bool on = bits != 0u;
            const int sk = on ? __ffs(bits) - 1 : 0;
            bits &= bits - 1u;
            h[k].slot = wsl + sk;
            h[k].e = on ? 32 * jq + sk : -1;
            h[k].g0 = st.g0[h[k].slot];
            h[k].g1 = st.g1[h[k].slot];
            h[k].dx[0] = __fsub_rn(h[k].g0.x, pxa); h[k].dx[1] = __fsub_rn(h[k].g0.x, pxb);
            h[k].dy[0] = __fsub_rn(h[k].g0.y, pya); h[k].dy[1] = __fsub_rn(h[k].g0.y, pyb);
            const float zxa = __fmul_rn(h[k].g0.z, h[k].dx[0]), zxb = __fmul_rn(h[k].g0.z, h[k].dx[1]);
            const float cya = __fmul_rn(__fmul_rn(h[k].g1.x, h[k].dy[0]), h[k].dy[0]);
            const float cyb = __fmul_rn(__fmul_rn(h[k].g1.x, h[k].dy[1]), h[k].dy[1]);
            h[k].sigma[0] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[0], zxa), h[k].dx[0], cya);
            h[k].sigma[1] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[0], zxb), h[k].dx[1], cya);
            h[k].sigma[2] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[1], zxa), h[k].dx[0], cyb);
            h[k].sigma[3] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[1], zxb), h[k].dx[1], cyb);
            h[k].any = false;
            uint32_t lim = sigma_limit(h[k].g1.y, true);   // (no hit: e = -1 < first[p])
            asm volatile("" : "+r"(lim));                   // one limit per hit, not re-derived per pixel
            // all the predicate work of a hit happens here: a pixel the Gaussian does not reach (or that lies behind
            // the pixel's last contributor) gets sigma = +huge, so that below vis = a = 0 EXACTLY, ra = 1, Tn = T and
            // every term it adds is an exact zero -- the body needs no per-pixel flags
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const bool v = h[k].e >= first[p] && __float_as_uint(h[k].sigma[p]) < lim;
                h[k].sigma[p] = v ? h[k].sigma[p] : 1e30f;
                asm volatile("" : "+f"(h[k].sigma[p]));   // select HERE: the predicate dies instead of riding into the body
                h[k].any |= v;
            }
        }
        bool anyv = h[0].any;
#pragma unroll
        for (int k = 1; k < HITS; ++k) anyv |= h[k].any;
        if (__any_sync(FULL, anyv)) {
#pragma unroll
            for (int k = 0; k < HITS; ++k) {
                const Hit& hk = h[k];
                const float4 g2 = st.g2[hk.slot];
                const float opac = hk.g1.z;
                float Swx = 0.f, Swy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f, Su = 0.f;
                float col[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float dx = hk.dx[p & 1], dy = hk.dy[p >> 1];
                    const float vis = __expf(-hk.sigma[p]);
                    const float ov = opac * vis;
                    const float a = fminf(MAX_ALPHA, ov);
                    const float ra = __fdividef(1.0f, 1.0f - a);
                    const float Tn = T[p] * ra; // transmittance in front of this Gaussian
                    T[p] = Tn;
                    const float fac = a * Tn;
                    float cdot = hk.g1.w * vr[p][0];
                    cdot = fmaf(g2.x, vr[p][1], cdot);
                    cdot = fmaf(g2.y, vr[p][2], cdot);
                    cdot = fmaf(g2.z, vr[p][3], cdot);
                    if (CH == 5) cdot = fmaf(g2.w, vr[p][CH - 1], cdot);
                    const float v_a = fmaf(Tn, cdot, ra * (va_eff[p] - bdot[p]));
                    bdot[p] = fmaf(fac, cdot, bdot[p]);
                    const float u = (ov <= MAX_ALPHA) ? vis * v_a : 0.f;
                    const float v_sigma = -opac * u;
                    const float wx = v_sigma * dx, wy = v_sigma * dy;
                    Swx += wx;
                    Swy += wy;
                    Sxx = fmaf(wx, dx, Sxx);
                    Sxy = fmaf(wx, dy, Sxy);
                    Syy = fmaf(wy, dy, Syy);
                    if (!GEOM) {
                        Su += u;
#pragma unroll
                        for (int c = 0; c < 3; ++c) col[c] = fmaf(fac, vr[p][c], col[c]);
                    }
#pragma unroll
                    for (int c = 3; c < CH; ++c) col[c] = fmaf(fac, vr[p][c], col[c]);
                }
                if (hk.any) {
                    const float ca = 2.0f * hk.g0.z, cb = hk.g0.w, cc = 2.0f * hk.g1.x;
                    const float v0 = fmaf(ca, Swx, cb * Swy), v1 = fmaf(cb, Swx, cc * Swy);
                    float* dst = v_rec + (int64_t)st.id[hk.slot] * REC;
                    red_add_v4(dst, v0, v1, 0.5f * Sxx, Sxy);
                    if (GEOM) {
                        atomicAdd(dst + 4, 0.5f * Syy);
                        red_add_v4(dst + 8, 0.f, col[3], col[4], 0.f);
                    } else {
                        red_add_v4(dst + 4, 0.5f * Syy, Su, col[0], col[1]);
                        red_add_v4(dst + 8, col[2], col[3], col[4], 0.f);
                    }
                }
            }
        }
        if (wmin > base) {
            ++base;
            if (staged < n_words) {
                commit<NW, true>(st, pf, staged, bx0, by0, lane);
                ++staged;
                prefetch(pf, staged, warp_last, n_entries, rec, vals, lane);
            }
        }
    }
}

// Forward, same organisation: a lane owns a 2x2 quad (T, 5 accumulators, last index per pixel in registers), a warp a
// 16x8 half tile.  Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188.  The per-pixel body is predicated, not
// branched; the contribute / saturate decisions are those of blend.cu::blend_fwd_kernel.
template <int CH, int NW, int HITS>
__global__ void __launch_bounds__(THREADS, 8)
blend_fwd_lq_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                    const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                    int32_t* __restrict__ last) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using R = Ring<NW, false>;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    R& st = reinterpret_cast<R*>(s_raw)[wid];
    // half tile h = 2 * tile_y + half; a CTA holds blockDim.x / 32 consecutive halves of one tile column
    const int hidx = blockIdx.y * (blockDim.x >> 5) + wid, ty = hidx >> 1, half = hidx & 1;
    const int tile = ty * cam.tw + blockIdx.x;
    const int qx = lane & 7, qy = lane >> 3;
    const int x0 = blockIdx.x * TILE + 2 * qx, y0 = ty * TILE + 8 * half + 2 * qy;
    const float bx0 = (float)(blockIdx.x * TILE) + 0.5f, by0 = (float)(ty * TILE + 8 * half) + 0.5f;
    const float pxa = (float)x0 + 0.5f, pxb = (float)x0 + 1.5f, pya = (float)y0 + 0.5f, pyb = (float)y0 + 1.5f;
    const int start = ranges[tile], end = ranges[tile + 1];
    const int n_entries = end - start;
    const int n_words = (n_entries + 31) >> 5;

    // T[p] > 0: transmittance of a live pixel; T[p] < 0: the pixel is DONE (saturated, or outside the image) and -T[p]
    // is its final transmittance -- the flag rides in the sign bit, so the loop carries no separate booleans
    float T[4], acc[4][CH];
    int last_idx[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        T[p] = (x0 + (p & 1) < cam.width && y0 + (p >> 1) < cam.height) ? 1.0f : -1.0f;
        last_idx[p] = start - 1;
#pragma unroll
        for (int k = 0; k < CH; ++k) acc[p][k] = 0.f;
    }
    const bool quad_done = T[0] < 0.f && T[1] < 0.f && T[2] < 0.f && T[3] < 0.f;

    if (n_words > 0 && !__all_sync(FULL, quad_done)) {
        int staged = 0, base = 0;
        staged = initial_fill<NW, false, 1>(st, start, n_entries, n_words, rec, vals, bx0, by0, lane);
        Pf pf;
        prefetch<1>(pf, staged, start, n_entries, rec, vals, lane);
        const uint32_t* hq = st.hits + lane;
        int jq = 0;
        unsigned wsl = 0;
        uint32_t bits = hq[0];
        if (quad_done) { jq = FAR; bits = 0u; }
        while (true) {
            if (bits == 0u && jq + 1 < staged) {
                ++jq;
                wsl = (wsl + 32u) & (32u * NW - 1u);
                bits = hq[wsl];
            }
            const bool act = bits != 0u;
            const int wmin = __reduce_min_sync(FULL, act ? jq : jq + 1);
            if (wmin >= n_words) break;
            Hit h[HITS];
#pragma unroll
            for (int k = 0; k < HITS; ++k) {
                const bool on = bits != 0u;
                const int sk = on ? __ffs(bits) - 1 : 0;
                bits &= bits - 1u;
                h[k].slot = wsl + sk;
                h[k].e = 32 * jq + sk;
                h[k].g0 = st.g0[h[k].slot];
                h[k].g1 = st.g1[h[k].slot];
                h[k].dx[0] = __fsub_rn(h[k].g0.x, pxa); h[k].dx[1] = __fsub_rn(h[k].g0.x, pxb);
                h[k].dy[0] = __fsub_rn(h[k].g0.y, pya); h[k].dy[1] = __fsub_rn(h[k].g0.y, pyb);
                const float zxa = __fmul_rn(h[k].g0.z, h[k].dx[0]), zxb = __fmul_rn(h[k].g0.z, h[k].dx[1]);
                const float cya = __fmul_rn(__fmul_rn(h[k].g1.x, h[k].dy[0]), h[k].dy[0]);
                const float cyb = __fmul_rn(__fmul_rn(h[k].g1.x, h[k].dy[1]), h[k].dy[1]);
                h[k].sigma[0] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[0], zxa), h[k].dx[0], cya);
                h[k].sigma[1] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[0], zxb), h[k].dx[1], cya);
                h[k].sigma[2] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[1], zxa), h[k].dx[0], cyb);
                h[k].sigma[3] = __fmaf_rn(__fmaf_rn(h[k].g0.w, h[k].dy[1], zxb), h[k].dx[1], cyb);
                h[k].any = false;
                uint32_t lim = sigma_limit(h[k].g1.y, on);
                asm volatile("" : "+r"(lim));
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    if (HITS == 1) {
                        // one hit per iteration: the pixel's liveness cannot change before the hit is applied, so it is
                        // folded in here and a pixel that is not hit gets sigma = +huge (a = 0 exactly below)
                        const bool v = __float_as_uint(h[k].sigma[p]) < lim && T[p] > 0.f;
                        h[k].sigma[p] = v ? h[k].sigma[p] : 1e30f;
                        asm volatile("" : "+f"(h[k].sigma[p]));   // select HERE: the predicate dies instead of riding into the body
                        h[k].any |= v;
                    } else {
                        // (liveness is re-checked when the hit is applied: an earlier hit of this iteration may saturate the pixel)
                        h[k].valid[p] = __float_as_uint(h[k].sigma[p]) < lim;
                        h[k].any |= h[k].valid[p] && T[p] > 0.f;
                    }
                }
            }
            bool anyh = h[0].any;
#pragma unroll
            for (int k = 1; k < HITS; ++k) anyh |= h[k].any;
            if (__any_sync(FULL, anyh)) {
#pragma unroll
                for (int k = 0; k < HITS; ++k) {
                    const Hit& hk = h[k];
                    const float4 g2 = st.g2[hk.slot];
                    const int idx = start + hk.e;
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        if (HITS == 1) {
                            // a = 0 exactly for a pixel that is not hit (or done): nT = T, vis = 0.  A live pixel has
                            // T > T_CUTOFF, a done one T < 0 (so has nT): "live and nT <= cutoff" is ONE unsigned compare
                            const float a = fminf(MAX_ALPHA, hk.g1.z * __expf(-hk.sigma[p]));
                            const float nT = T[p] * (1.0f - a);
                            const bool sat = __float_as_uint(nT) <= __float_as_uint(T_CUTOFF);
                            const float vis = sat ? 0.f : a * T[p];
                            acc[p][0] = fmaf(hk.g1.w, vis, acc[p][0]);
                            acc[p][1] = fmaf(g2.x, vis, acc[p][1]);
                            acc[p][2] = fmaf(g2.y, vis, acc[p][2]);
                            acc[p][3] = fmaf(g2.z, vis, acc[p][3]);
                            if (CH == 5) acc[p][CH - 1] = fmaf(g2.w, vis, acc[p][CH - 1]);
                            last_idx[p] = vis > 0.f ? idx : last_idx[p];
                            T[p] = sat ? -T[p] : nT;
                            continue;
                        }
                        const bool hit = hk.valid[p] && T[p] > 0.f;
                        // (a pixel that is not hit may evaluate anything here: every use below is behind a select)
                        const float a = fminf(MAX_ALPHA, hk.g1.z * __expf(-hk.sigma[p]));
                        const float nT = T[p] * (1.0f - a);
                        const bool sat = hit && nT <= T_CUTOFF;
                        const bool contrib = hit && !sat;
                        const float vis = contrib ? a * T[p] : 0.f;
                        acc[p][0] = fmaf(hk.g1.w, vis, acc[p][0]);
                        acc[p][1] = fmaf(g2.x, vis, acc[p][1]);
                        acc[p][2] = fmaf(g2.y, vis, acc[p][2]);
                        acc[p][3] = fmaf(g2.z, vis, acc[p][3]);
                        if (CH == 5) acc[p][CH - 1] = fmaf(g2.w, vis, acc[p][CH - 1]);
                        last_idx[p] = contrib ? idx : last_idx[p];
                        T[p] = contrib ? nT : (sat ? -T[p] : T[p]);
                    }
                }
                // quad saturated: all four sign bits set
                if ((int32_t)(__float_as_uint(T[0]) & __float_as_uint(T[1]) & __float_as_uint(T[2]) & __float_as_uint(T[3])) < 0) {
                    jq = FAR; bits = 0u;
                }
            }
            if (wmin > base) {
                ++base;
                if (staged < n_words) {
                    commit<NW, false>(st, pf, staged, bx0, by0, lane);
                    ++staged;
                    prefetch<1>(pf, staged, start, n_entries, rec, vals, lane);
                }
            }
        }
    }
    const int64_t P = (int64_t)cam.width * cam.height;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int xi = x0 + (p & 1), yi = y0 + (p >> 1);
        if (xi < cam.width && yi < cam.height) {
            const int64_t pix = (int64_t)yi * cam.width + xi;
#pragma unroll
            const float Tf = fabsf(T[p]);
#pragma unroll
            for (int k = 0; k < CH; ++k) render[k * P + pix] = (k < 3) ? fmaf(Tf, cam.bg[k], acc[p][k]) : acc[p][k];
            alpha[pix] = 1.0f - Tf;
            last[pix] = last_idx[p];
        }
    }
}

template <int CH, int NW, int HITS>
int launch_fwd(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges, float* render, float* alpha,
               int32_t* last, cudaStream_t stream) {
    constexpr size_t smem = sizeof(Ring<NW, false>) * WARPS;
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        ISOGS_CUDA(cudaFuncSetAttribute(blend_fwd_lq_kernel<CH, NW, HITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_fwd_lq_kernel<CH, NW, HITS><<<grid, THREADS, smem, stream>>>(dc, (const float4*)rec, vals, ranges, render, alpha, last);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

template <int CH, int NW, bool GEOM, int HITS>
int launch(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges, const float* alpha,
           const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec, cudaStream_t stream) {
    constexpr size_t smem = sizeof(Ring<NW, true>) * WARPS;
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        ISOGS_CUDA(cudaFuncSetAttribute(blend_bwd_lq_kernel<CH, NW, GEOM, HITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_bwd_lq_kernel<CH, NW, GEOM, HITS><<<grid, THREADS, smem, stream>>>(dc, (const float4*)rec, vals, ranges, alpha, last,
                                                                                 v_render, v_alpha, v_rec);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

} // namespace lq

} // namespace isogs

using namespace isogs;

extern "C" int isogs_blend_fwd(const isogs_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                               const int32_t* tile_ranges, float* render, float* alpha, int32_t* last,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_fwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    if ((uintptr_t)rec & 15) { set_error("isogs_blend_fwd: rec must be 16-byte aligned"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    // one hit per lane per iteration in the forward (two measured 14 % slower: 0.305 vs 0.267 ms), two in the backward
    if (ch == 4) return lq::launch_fwd<4, lq::NW_DEFAULT, lq::HITS_FWD>(dc, rec, vals_sorted, tile_ranges, render, alpha, last, stream);
    return lq::launch_fwd<5, lq::NW_DEFAULT, lq::HITS_FWD>(dc, rec, vals_sorted, tile_ranges, render, alpha, last, stream);
}

static int blend_bwd_entry(const char* who, bool geom, const isogs_camera* cam, int ch, const float* rec,
                           const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha, const int32_t* last,
                           const float* v_render, const float* v_alpha, float* v_rec, cudaStream_t stream) {
    if (!cam || (ch != 4 && ch != 5)) { set_error("%s: ch must be 4 or 5", who); return ISOGS_E_ARG; }
    if (((uintptr_t)rec | (uintptr_t)v_rec) & 15) { set_error("%s: rec and v_rec must be 16-byte aligned", who); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
#define LQ(CHN, G) lq::launch<CHN, lq::NW_DEFAULT, G, lq::HITS_BWD>(dc, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream)
    if (geom) return ch == 4 ? LQ(4, true) : LQ(5, true);
    return ch == 4 ? LQ(4, false) : LQ(5, false);
#undef LQ
}

extern "C" int isogs_blend_bwd(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                               const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                               const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                               void* stream_) {
    (void)N;
    return blend_bwd_entry("isogs_blend_bwd", false, cam, ch, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha,
                           v_rec, (cudaStream_t)stream_);
}

// Tracking variant: identical traversal, but only the gradient components the POSE depends on are reduced and
// written: v_rec[0..4] (x, y, conic) and v_rec[9..10] (z, z^2); the colour / opacity slots are left untouched.
extern "C" int isogs_blend_bwd_geom(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                                    const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                                    const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                                    void* stream_) {
    (void)N;
    return blend_bwd_entry("isogs_blend_bwd_geom", true, cam, ch, rec, vals_sorted, tile_ranges, alpha, last, v_render,
                           v_alpha, v_rec, (cudaStream_t)stream_);
}
