// blend.cu -- north_star (3) renderCUDA forward: per 16x16 tile alpha-blend of colour, depth,
// depth^2 and silhouette; north_star (4) backward: reverse per-pixel traversal with quad-reduced
// no-return atomics (RED).  Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188,
// RasterizeToPixels3DGSBwd.cu:15-278.
//
// These are the four-lanes-per-quad kernels (ISOGS_FWD_LQ=0 / ISOGS_BWD_LQ=0); the default kernels live in
// blend_lq.cu and share this file's C entry points.
//
// Design: warp-autonomous, no CTA barrier.  A CTA is one tile, a warp owns an 8x4
// pixel block of it and walks the tile's depth-sorted list on its own through a warp-private RING
// of NW words x 32 Gaussians in shared memory: staging a word, lane L loads Gaussian L and tests
// its {sigma <= L} footprint against the warp's EIGHT 2x2 pixel quads; eight ballots give every
// quad (4 lanes) its own 32-bit hit mask.  The quads then consume the ring INDEPENDENTLY, each
// with its own cursor: in one warp iteration every quad blends a different Gaussian (its next
// hit).  A depth-sorted list sweeps across the tile, so at any list position only a band of the
// tile is hit; independent cursors let quads ahead of / behind the band keep working, and a word
// is retired (and its slot refilled) once every quad has passed it.  Backward: the 11 partial
// gradients are summed over the 4 lanes of
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

// Warp-private ring of NW words x 32 staged Gaussians (dynamic shared memory).
template <int NW, bool WITH_ID>
struct __align__(16) WarpRing {
    float4 g0[32 * NW]; // x, y, 0.5*conic_a, conic_b
    float4 g1[32 * NW]; // 0.5*conic_c, L (= sigma_max), opacity, c0
    float4 g2[32 * NW]; // c1, c2, c3 (depth), c4 (depth^2)
    uint32_t hits[NW][8]; // per word, per quad: which of the 32 Gaussians can touch the quad
    int id[WITH_ID ? 32 * NW : 4];
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
// conservative and works on SQUARED distances (no square root): squared axis-aligned half extents
// 2 L cov_xx, 2 L cov_yy AND the squared bounding-circle radius 2 L (max(cov_xx, cov_yy) + |cov_xy|)
// (Gershgorin bound on lambda_max), each padded for rounding, against the quad's pixel-centre square.
__device__ __forceinline__ uint32_t quad_hits(bool have, const float4& g0, const float4& g1, float bx0, float by0,
                                              int lane) {
    float hx2 = -1.f, hy2 = -1.f, R2 = -1.f;
    if (have) {
        const float a = 2.0f * g0.z, b = g0.w, c = 2.0f * g1.x;
        const float inv_det = __fdividef(1.0f, fmaf(a, c, -(b * b)));
        const float twoL = 2.0f * g1.y * inv_det;
        const float cxx = c * twoL, cyy = a * twoL, cxy = fabsf(b * twoL); // 2 L cov
        hx2 = fmaf(cxx, 1.01f, 0.05f);
        hy2 = fmaf(cyy, 1.01f, 0.05f);
        R2 = fmaf(fmaxf(cxx, cyy) + cxy, 1.01f, 0.05f);
    }
    const float X = g0.x - bx0, Y = g0.y - by0; // relative to the block's first pixel centre
    float ex2[4], ey2[2];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float e = fmaxf(fmaxf((float)(2 * c) - X, X - (float)(2 * c + 1)), 0.f);
        ex2[c] = e * e;
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const float e = fmaxf(fmaxf((float)(2 * r) - Y, Y - (float)(2 * r + 1)), 0.f);
        ey2[r] = e * e;
    }
    uint32_t bits = 0;
    const int myq = lane >> 2;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float ex = ex2[q & 3], ey = ey2[q >> 2];
        const bool hit = ex <= hx2 && ey <= hy2 && ex + ey <= R2;
        const uint32_t m = __ballot_sync(FULL, hit);
        if (myq == q) bits = m;
    }
    return bits;
}

// One list entry held in registers between its global load and its commit to the ring, so the
// vals -> rec dependent load chain of word w+1 overlaps the blending of the words before it.  The chain is
// split over TWO calls: the call for word w issues the record loads with the Gaussian id that the call for
// word w-1 requested, and requests the id of word w+1 -- otherwise the warp would sit on the long scoreboard
// inside the prefetch, waiting for `vals` before it can even issue the `rec` loads (8 % of the forward's stall
// samples in profiles/r1h_blend_fwd).
struct Prefetch {
    float4 g0, g1, g2;
    int id;
    bool have;
    int id_next;      // id of entry (32 * next_word + lane), requested one call ahead
    int next_word;    // warp-uniform; -1 before the first call
};

__device__ __forceinline__ int list_id(int word, int first, int dir, int n_entries, const int32_t* __restrict__ vals,
                                       int lane) {
    const int e = 32 * word + lane;
    return (e < n_entries) ? __ldg(vals + first + dir * e) : -1;
}

// Load list entry (32*word + lane), i.e. sorted index first + dir*(32*word + lane).  TWO_STAGE requests the id
// of word+1 as well (used by the forward; in the backward the extra registers cost more than the stall).
template <bool TWO_STAGE>
__device__ __forceinline__ void prefetch_word(Prefetch& pf, int word, int first, int dir, int n_entries,
                                              const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                                              int lane) {
    const int id = (TWO_STAGE && pf.next_word == word) ? pf.id_next : list_id(word, first, dir, n_entries, vals, lane);
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
    if (TWO_STAGE) {
        pf.id_next = list_id(word + 1, first, dir, n_entries, vals, lane);
        pf.next_word = word + 1;
    }
}

// Commit the prefetched word into ring slot word % NW and publish its per-quad hit masks.
// Called by the whole warp at a convergent point; the caller guarantees the slot is retired.
template <int NW, bool WITH_ID>
__device__ __forceinline__ void commit_word(WarpRing<NW, WITH_ID>& st, const Prefetch& pf, int word, float bx0,
                                            float by0, int lane) {
    const unsigned wsl = ((unsigned)word & (NW - 1)) * 32u;
    const unsigned slot = wsl + lane;
    __syncwarp(); // every lane has finished reading the slot being overwritten
    if (pf.have) {
        st.g0[slot] = pf.g0;
        st.g1[slot] = pf.g1;
        st.g2[slot] = pf.g2;
        if (WITH_ID) st.id[slot] = pf.id;
    }
    const uint32_t m = quad_hits(pf.have, pf.g0, pf.g1, bx0, by0, lane);
    if ((lane & 3) == 0) st.hits[0][(wsl >> 2) + (lane >> 2)] = m;
    __syncwarp();
}

template <int CH, int NW>
__global__ void __launch_bounds__(BLOCK)
blend_fwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                 int32_t* __restrict__ last) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using Ring = WarpRing<NW, false>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    Ring& st = reinterpret_cast<Ring*>(s_raw)[wid];
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(wid, lane);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const float bx0 = (float)(blockIdx.x * TILE + (wid & 1) * RW) + 0.5f;
    const float by0 = (float)(blockIdx.y * TILE + (wid >> 1) * RH) + 0.5f;
    const int start = ranges[tile], end = ranges[tile + 1];
    const int n_entries = end - start;
    const int n_words = (n_entries + 31) >> 5;
    const int qshift = lane & 28, q = lane >> 2;
    constexpr int FAR = 0x3fffffff;

    float T = 1.0f;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) acc[k] = 0.f;
    int last_idx = start - 1;
    bool done = !inside;

    if (n_words > 0 && !__all_sync(FULL, done)) {
        int staged = 0, base = 0; // warp-uniform: words [base, staged) are live in the ring
        Prefetch pf;
        pf.next_word = -1;
        prefetch_word<true>(pf, 0, start, 1, n_entries, rec, vals, lane);
        for (; staged < min(NW, n_words); ++staged) {
            commit_word<NW, false>(st, pf, staged, bx0, by0, lane);
            prefetch_word<true>(pf, staged + 1, start, 1, n_entries, rec, vals, lane);
        }
        static_assert((NW & (NW - 1)) == 0, "NW must be a power of two");
        const uint32_t* hq = &st.hits[0][q]; // this quad's hit word of ring word w is hq[8 * w]
        int jq = 0;                          // this quad's cursor (absolute word), what is left of that word,
        unsigned wsl = 0;                    // and the ring slot of its first entry: (jq % NW) * 32
        uint32_t bits = hq[0];
        if (((__ballot_sync(FULL, !done) >> qshift) & 15u) == 0u) { jq = FAR; bits = 0u; }
        while (true) {
            if (bits == 0u && jq + 1 < staged) { // step to the next staged word (one per iteration)
                ++jq;
                wsl = (wsl + 32u) & (32u * NW - 1u);
                bits = hq[wsl >> 2];
            }
            const bool act = bits != 0u;
            // progress of this quad = first word it may still read
            const int wmin = __reduce_min_sync(FULL, act ? jq : jq + 1);
            if (wmin >= n_words) break; // every quad is past the end of the list (or saturated)
            STAT(1, 1);
            STAT(2, __popc(__ballot_sync(FULL, act)));
            const int sl = act ? __ffs(bits) - 1 : 0;
            bits &= bits - 1u;
            const float4* gp = st.g0 + (wsl + sl);
            const float4 g0s = gp[0];
            const float4 g1s = gp[32 * NW];
            float dx, dy;
            const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
            bool newly = false;
            if (act && !done && sigma >= 0.f && sigma <= g1s.y) {
                const float a = fminf(MAX_ALPHA, g1s.z * __expf(-sigma));
                const float nT = T * (1.0f - a);
                if (nT <= T_CUTOFF) {
                    done = true;
                    newly = true;
                } else {
                    const float vis = a * T;
                    const float4 g2 = gp[64 * NW];
                    acc[0] = fmaf(g1s.w, vis, acc[0]);
                    acc[1] = fmaf(g2.x, vis, acc[1]);
                    acc[2] = fmaf(g2.y, vis, acc[2]);
                    acc[3] = fmaf(g2.z, vis, acc[3]);
                    if (CH == 5) acc[CH - 1] = fmaf(g2.w, vis, acc[CH - 1]);
                    last_idx = start + 32 * jq + sl;
                    T = nT;
                }
            }
            if (__any_sync(FULL, newly)) {
                if (((__ballot_sync(FULL, !done) >> qshift) & 15u) == 0u) { jq = FAR; bits = 0u; } // quad saturated
            }
            // retire the oldest word once every quad is past it; refill its slot with the next one
            if (wmin > base) {
                ++base;
                if (staged < n_words) {
                    commit_word<NW, false>(st, pf, staged, bx0, by0, lane);
                    ++staged;
                    prefetch_word<true>(pf, staged, start, 1, n_entries, rec, vals, lane);
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

// Same scheme for 8 values (6 shuffles): lane l ends with the totals of components
// 4*(l&1) + 2*((l>>1)&1) + {0,1}.
__device__ __forceinline__ void quad_reduce8(float (&v)[8], int lane) {
    {
        const bool up = lane & 1;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float send = up ? v[k] : v[k + 4];
            const float keep = up ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 1);
        }
    }
    {
        const bool up = lane & 2;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float send = up ? v[k] : v[k + 2];
            const float keep = up ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 2);
        }
    }
}

// GEOM = true: tracking.  Only the pose gets a gradient there (splatam.py:1247-1290), and it depends on the
// record through {x, y, conic, z, z^2} alone -- colour and opacity gradients are neither reduced nor written
// (7 values, 6 shuffles, 2 REDs per lane instead of 11 / 9 / 3).
template <int CH, int NW, bool GEOM>
__global__ void __launch_bounds__(BLOCK)
blend_bwd_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                 const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                 const int32_t* __restrict__ last, const float* __restrict__ v_render,
                 const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using Ring = WarpRing<NW, true>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    Ring& st = reinterpret_cast<Ring*>(s_raw)[wid];
    const int tile = blockIdx.y * cam.tw + blockIdx.x;
    const PixelMap pm = pixel_of(wid, lane);
    const bool inside = pm.px_i < cam.width && pm.py_i < cam.height;
    const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
    const float bx0 = (float)(blockIdx.x * TILE + (wid & 1) * RW) + 0.5f;
    const float by0 = (float)(blockIdx.y * TILE + (wid >> 1) * RH) + 0.5f;
    const int start = ranges[tile];
    const int64_t P = (int64_t)cam.width * cam.height;
    const int64_t pix = inside ? (int64_t)pm.py_i * cam.width + pm.px_i : 0;
    const int qshift = lane & 28, q = lane >> 2;

    const int my_last = inside ? last[pix] : start - 1;
    const int warp_last = __reduce_max_sync(FULL, my_last);
    if (warp_last < start) return; // nothing was blended into this warp's block
    int quad_last = max(my_last, __shfl_xor_sync(FULL, my_last, 1));
    quad_last = max(quad_last, __shfl_xor_sync(FULL, quad_last, 2));
    // the list is walked back to front: entry e is sorted index warp_last - e
    const int n_entries = warp_last - start + 1;
    const int n_words = (n_entries + 31) >> 5;
    const int e_first = warp_last - quad_last; // entries in front of it lie behind this quad's last contributor
    const int my_first = warp_last - my_last;  // same, for this pixel

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

    // component triple this lane owns after quad_reduce12 (pair after quad_reduce8: vector {x,y,a,b,c,z,z^2,-}
    // lives at record offsets {0,1,2,3,4,9,10})
    const int comp0 = GEOM ? 4 * (lane & 1) + 2 * ((lane >> 1) & 1) : 6 * (lane & 1) + 3 * ((lane >> 1) & 1);
    const bool third = comp0 + 2 < 6 + CH;
    const int goff0 = comp0 < 5 ? comp0 : comp0 + 4, goff1 = comp0 + 1 < 5 ? comp0 + 1 : comp0 + 5;

    int staged = 0, base = 0; // warp-uniform: words [base, staged) are live in the ring
    Prefetch pf;
    pf.next_word = -1;
    prefetch_word<false>(pf, 0, warp_last, -1, n_entries, rec, vals, lane);
    for (; staged < min(NW, n_words); ++staged) {
        commit_word<NW, true>(st, pf, staged, bx0, by0, lane);
        prefetch_word<false>(pf, staged + 1, warp_last, -1, n_entries, rec, vals, lane);
    }
    static_assert((NW & (NW - 1)) == 0, "NW must be a power of two");
    const uint32_t* hq = &st.hits[0][q]; // this quad's hit word of ring word w is hq[8 * w]
    // hit word of this quad with the entries in front of e_first masked off
    auto fetch = [&](int w, unsigned wsl_) -> uint32_t {
        const uint32_t m = hq[wsl_ >> 2];
        const int cut = e_first - 32 * w;
        return cut <= 0 ? m : (cut >= 32 ? 0u : (m & ~((1u << cut) - 1u)));
    };
    int jq = min(e_first >> 5, staged - 1);
    unsigned wsl = ((unsigned)jq & (NW - 1)) * 32u;
    uint32_t bits = fetch(jq, wsl);
    if (quad_last < start) { jq = 0x3fffffff; bits = 0u; } // nothing was blended into this quad
    while (true) {
        if (bits == 0u && jq + 1 < staged) {
            ++jq;
            wsl = (wsl + 32u) & (32u * NW - 1u);
            bits = fetch(jq, wsl);
        }
        const bool act = bits != 0u;
        const int wmin = __reduce_min_sync(FULL, act ? jq : jq + 1);
        if (wmin >= n_words) break;
        const int s = act ? __ffs(bits) - 1 : 0;
        bits &= bits - 1u;
        const unsigned slot = wsl + s;
        const float4* gp = st.g0 + slot;
        const float4 g0s = gp[0];
        const float4 g1s = gp[32 * NW];
        float dx, dy;
        const float sigma = sigma_of(g0s, g1s, px, py, dx, dy);
        const bool valid = act && (32 * jq + s >= my_first) && sigma >= 0.f && sigma <= g1s.y;
        const uint32_t vb = __ballot_sync(FULL, valid);
        STAT(9, 1);
        STAT(10, __popc(__ballot_sync(FULL, act)));
        STAT(11, __popc(vb));
        if (vb != 0u) {
            STAT(12, 1);
            // branch-free body: an invalid lane runs with sigma = 0 and contributes exact zeros
            const float4 g2 = gp[64 * NW];
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
            if constexpr (GEOM) {
                float v[8];
                v[0] = fmaf(ca, wx, cb * wy);
                v[1] = fmaf(cb, wx, cc * wy);
                v[2] = 0.5f * wx * dx;
                v[3] = wx * dy;
                v[4] = 0.5f * wy * dy;
                v[5] = fac * vr[3];
                v[6] = (CH == 5) ? fac * vr[CH - 1] : 0.f;
                v[7] = 0.f;
                quad_reduce8(v, lane);
                if ((vb >> qshift) & 15u) {
                    float* dst = v_rec + (int64_t)st.id[slot] * REC;
                    atomicAdd(dst + goff0, v[0]);
                    if (comp0 != 6) atomicAdd(dst + goff1, v[1]);
                }
            } else {
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
                quad_reduce12(v, lane);
                if ((vb >> qshift) & 15u) {
                    float* dst = v_rec + (int64_t)st.id[slot] * REC + comp0;
                    atomicAdd(dst, v[0]);
                    atomicAdd(dst + 1, v[1]);
                    if (third) atomicAdd(dst + 2, v[2]);
                }
            }
        }
        if (wmin > base) {
            ++base;
            if (staged < n_words) {
                commit_word<NW, true>(st, pf, staged, bx0, by0, lane);
                ++staged;
                prefetch_word<false>(pf, staged, warp_last, -1, n_entries, rec, vals, lane);
            }
        }
    }
}

} // namespace isogs

using namespace isogs;

namespace {
constexpr int NW_FWD = 4, NW_BWD = 4, NW_LQ = 8;

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

template <int CH, int NW>
static int launch_fwd(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges, float* render,
                      float* alpha, int32_t* last, cudaStream_t stream) {
    constexpr size_t smem = sizeof(WarpRing<NW, false>) * NWARP;
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        int rc = set_smem(blend_fwd_kernel<CH, NW>, smem);
        if (rc != 0) return rc;
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_fwd_kernel<CH, NW><<<grid, BLOCK, smem, stream>>>(dc, (const float4*)rec, vals, ranges, render, alpha, last);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

template <int CH, int NW, bool GEOM = false>
static int launch_bwd(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges,
                      const float* alpha, const int32_t* last, const float* v_render, const float* v_alpha,
                      float* v_rec, cudaStream_t stream) {
    constexpr size_t smem = sizeof(WarpRing<NW, true>) * NWARP;
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        int rc = set_smem(blend_bwd_kernel<CH, NW, GEOM>, smem);
        if (rc != 0) return rc;
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_bwd_kernel<CH, NW, GEOM><<<grid, BLOCK, smem, stream>>>(dc, (const float4*)rec, vals, ranges, alpha, last, v_render,
                                                                   v_alpha, v_rec);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

namespace isogs {
int launch_blend_bwd_lq(const DevCam& dc, int ch, int nw, bool geom, bool vec, bool idg, const float* rec, const int32_t* vals,
                        const int32_t* ranges, const float* alpha, const int32_t* last, const float* v_render,
                        const float* v_alpha, float* v_rec, cudaStream_t stream); // blend_lq.cu
int launch_blend_fwd_lq(const DevCam& dc, int ch, int nw, const float* rec, const int32_t* vals, const int32_t* ranges,
                        float* render, float* alpha, int32_t* last, cudaStream_t stream);
int launch_blend_fwd_ts(const DevCam& dc, int ch, bool bulk, const float* rec, const int32_t* vals, const int32_t* ranges, float* render,
                        float* alpha, int32_t* last, cudaStream_t stream); // blend_ts.cu
int launch_blend_bwd_ts(const DevCam& dc, int ch, bool geom, bool bulk, const float* rec, const int32_t* vals, const int32_t* ranges,
                        const float* alpha, const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                        cudaStream_t stream);
}

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

static int env_nw(const char* name, int dflt) {
    const char* e = getenv(name);
    const int v = e ? atoi(e) : dflt;
    return (v == 2 || v == 4 || v == 8) ? v : dflt;
}

extern "C" int isogs_blend_fwd(const isogs_camera* cam, int ch, const float* rec, const int32_t* vals_sorted,
                               const int32_t* tile_ranges, float* render, float* alpha, int32_t* last,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_fwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    if ((uintptr_t)rec & 15) { set_error("isogs_blend_fwd: rec must be 16-byte aligned"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    // default: the lane-per-quad kernel of blend_lq.cu with a ring of 8 words; ISOGS_FWD_LQ=0 selects the
    // quad-of-lanes kernel above (ring of 4 words), kept as the A/B baseline
    // default: the tile-shared-ring lane-per-quad kernel of blend_ts.cu; ISOGS_TS=0 falls back to the families below
    static const int cqsel = env_int("ISOGS_TS", 1);
    static const int bulk = env_int("ISOGS_TS_BULK", 0);
    if (cqsel) return launch_blend_fwd_ts(dc, ch, bulk != 0, rec, vals_sorted, tile_ranges, render, alpha, last, stream);
    static const int lq = env_int("ISOGS_FWD_LQ", 1);
    static const int nw = env_nw("ISOGS_NW_FWD", lq ? NW_LQ : NW_FWD); // ring words per warp (tuning knob)
    if (lq) return launch_blend_fwd_lq(dc, ch, nw, rec, vals_sorted, tile_ranges, render, alpha, last, stream);
#define ISOGS_FWD(CHN, NWN) launch_fwd<CHN, NWN>(dc, rec, vals_sorted, tile_ranges, render, alpha, last, stream)
    if (ch == 4) return nw == 2 ? ISOGS_FWD(4, 2) : nw == 8 ? ISOGS_FWD(4, 8) : ISOGS_FWD(4, 4);
    return nw == 2 ? ISOGS_FWD(5, 2) : nw == 8 ? ISOGS_FWD(5, 8) : ISOGS_FWD(5, 4);
#undef ISOGS_FWD
}

extern "C" int isogs_blend_bwd(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                               const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                               const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    (void)N;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_bwd: ch must be 4 or 5"); return ISOGS_E_ARG; }
    if (((uintptr_t)rec | (uintptr_t)v_rec) & 15) { set_error("isogs_blend_bwd: rec and v_rec must be 16-byte aligned"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    if (dc.tw <= 0 || dc.th <= 0) return ISOGS_OK;
    static const int cqsel = env_int("ISOGS_TS", 1);
    static const int bulk = env_int("ISOGS_TS_BULK", 0);
    if (cqsel) return launch_blend_bwd_ts(dc, ch, false, bulk != 0, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
    static const int lq = env_int("ISOGS_BWD_LQ", 1); // 1: lane-per-quad kernel (blend_lq.cu), 2: same with scalar REDs, 0: quad-of-lanes kernel
    static const int nw = env_nw("ISOGS_NW_BWD", lq ? NW_LQ : NW_BWD);
    if (lq) return launch_blend_bwd_lq(dc, ch, nw, false, lq != 2, lq == 3, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
#define ISOGS_BWD(CHN, NWN) \
    launch_bwd<CHN, NWN>(dc, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream)
    if (ch == 4) return nw == 2 ? ISOGS_BWD(4, 2) : nw == 8 ? ISOGS_BWD(4, 8) : ISOGS_BWD(4, 4);
    return nw == 2 ? ISOGS_BWD(5, 2) : nw == 8 ? ISOGS_BWD(5, 8) : ISOGS_BWD(5, 4);
#undef ISOGS_BWD
}

// Tracking variant: identical traversal, but only the gradient components the POSE depends on are reduced and
// written: v_rec[0..4] (x, y, conic) and v_rec[9..10] (z, z^2); the colour / opacity slots are left untouched.
extern "C" int isogs_blend_bwd_geom(const isogs_camera* cam, int ch, int64_t N, const float* rec,
                                    const int32_t* vals_sorted, const int32_t* tile_ranges, const float* alpha,
                                    const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                                    void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    (void)N;
    if (!cam || (ch != 4 && ch != 5)) { set_error("isogs_blend_bwd_geom: ch must be 4 or 5"); return ISOGS_E_ARG; }
    if (((uintptr_t)rec | (uintptr_t)v_rec) & 15) { set_error("isogs_blend_bwd_geom: rec and v_rec must be 16-byte aligned"); return ISOGS_E_ARG; }
    DevCam dc = make_devcam(*cam);
    static const int cqsel = env_int("ISOGS_TS", 1);
    static const int bulk = env_int("ISOGS_TS_BULK", 0);
    if (cqsel) return launch_blend_bwd_ts(dc, ch, true, bulk != 0, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
    static const int lq = env_int("ISOGS_BWD_LQ", 1);
    if (lq) return launch_blend_bwd_lq(dc, ch, env_nw("ISOGS_NW_BWD", NW_LQ), true, true, lq == 3, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
    if (ch == 4) return launch_bwd<4, 4, true>(dc, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
    return launch_bwd<5, 4, true>(dc, rec, vals_sorted, tile_ranges, alpha, last, v_render, v_alpha, v_rec, stream);
}
