// blend_ts.cu -- north_star (3) forward and (4) backward alpha-blend: lane-per-quad kernels over ONE ring per TILE,
// filled by bulk asynchronous copies that complete on mbarriers.
// Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188, RasterizeToPixels3DGSBwd.cu:15-278.
//
// Work mapping as in blend_lq.cu: a lane owns a 2x2 pixel quad (its four pixels live in registers), a warp a 16x8 half
// tile, every lane blends its quad's next hit per warp iteration; per-quad hit masks, independent cursors, no CTA
// barrier inside the walk; backward gradients leave as three red.global.add.v4.f32 per (quad, Gaussian).
//
// What changed is the staging.  A depth-sorted tile list sweeps across the tile, so the quads of a half tile work at
// list positions hundreds of entries apart; with a private ring of 8 words (256 entries) per warp the slowest quad
// pins the ring while the fastest starves.  A CPU replay of the walk on the bench workload (scripts/blend_walk_sim.py,
// which reproduces the measured 0.87 M warp iterations / 18 live lanes of blend_lq.cu) gives 0.56 M iterations / 29 live
// lanes once 16 words are resident -- practically the unbounded-window optimum.  Two private rings of 16 words do not
// fit the occupancy (57 KB per CTA); ONE ring of 16 words shared by the tile's two warps does (31 KB, 7 CTAs per SM):
//
//   * the records are staged ONCE per tile (each warp used to gather the whole tile list for itself);
//   * warp 0 is also the producer: per ring word one coalesced load of 32 Gaussian ids, then every lane issues ONE
//     48-byte cp.async.bulk.shared.global of its record, completing on the word's mbarrier (expect_tx = 48 x entries);
//     no register staging (the 13 prefetch registers of blend_lq.cu are gone), and at start all 16 words are in flight;
//   * each warp waits on a word's mbarrier when it first needs it (try_wait; blocking only when nothing else is
//     consumable) and builds its OWN per-quad hit masks for it;
//   * a slot is recycled when both warps have retired its word: each warp publishes its ring base in shared memory
//     (release fence), the producer reads the other warp's (acquire) -- no second set of barriers;
//   * a warp that finishes early (forward: all quads saturated) marks itself finished; warp 0 keeps producing for
//     warp 1 until that one is finished too, then waits for every copy it issued before the CTA's shared memory is freed.
//
// Every spin is bounded: on a timeout the warp stops (its pixels / gradients are then wrong and the parity tests fail
// loudly) and a device counter readable through isogs_debug_blend_timeouts() is bumped -- a kernel never hangs.
//
// The sigma expression, the contribute / skip / saturate decisions and the per-pixel accumulation order are those of
// blend_lq.cu, so images, `last` and gradients are unchanged.
#include "common.cuh"

namespace isogs {
namespace ts {

constexpr unsigned FULL = 0xffffffffu;
constexpr int WARPS = 2;   // half tiles (16x8 pixels, one warp each) per tile = per CTA
constexpr int THREADS = 32 * WARPS;
constexpr int FAR = 0x01000000;
constexpr int FIN = 0x7fffffff;
constexpr long long WAIT_CYCLES = 2000000000ll; // bound of every wait (about a second): a kernel never hangs

__device__ unsigned int g_timeouts = 0;

template <int NW, bool WITH_ID>
struct __align__(16) TileRing {
    float4 rec[3 * 32 * NW];       // entry slot s = global record verbatim: rec[3s] = x, y, 0.5*a, b | 0.5*c, L, opacity, c0 | c1, c2, z, z^2
    uint32_t hits[WARPS][32 * NW]; // per warp [ring word][quad]: which of the word's 32 entries can touch the quad
    int id[WITH_ID ? 32 * NW : 4]; // Gaussian ids (backward: address of the gradient record)
    unsigned long long full[NW];   // mbarriers: word landed (1 arrival + 48 bytes per entry)
    int prog[WARPS];               // ring base of each warp (FIN when the warp is finished)
    int issued;                    // words issued by the producer (warp 1 waits on a word only once it was issued)
    int warp_last[WARPS];          // backward: last contributor of each half tile
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(void* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(void* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(void* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking probe (try_wait may suspend the thread for a while; test_wait never does)
__device__ __forceinline__ bool mbar_test_wait(void* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded blocking wait of a WHOLE WARP (every lane must observe the completion itself: that is its acquire); false
// on timeout.  The phase must be the barrier's current or just-completed one (see try_publish).
__device__ __forceinline__ bool mbar_wait_warp(void* bar, uint32_t parity) {
    const long long t0 = clock64();
    while (!__all_sync(FULL, mbar_try_wait(bar, parity)))
        if (clock64() - t0 > WAIT_CYCLES) return false;
    return true;
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ int ld_volatile_shared(const int* p) {
    int v;
    asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_shared(int* p, int v) {
    asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
// one warp-uniform value of a shared word another warp updates (lane 0 reads, broadcast).  No fence: what is ordered
// against it is either an mbarrier wait (its own acquire) or a write into a slot whose last readers have CONSUMED
// the values they loaded before they published their base (see publish_base).
__device__ __forceinline__ int peek(const int* p, int lane) {
    int v = 0;
    if (lane == 0) v = ld_volatile_shared(p);
    return __shfl_sync(FULL, v, 0);
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
// the barrier receives one (pre-counted) arrival when all cp.async of this thread issued so far have landed
__device__ __forceinline__ void cp_async_arrive_noinc(void* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void note_timeout(int lane) {
    if (lane == 0) atomicAdd(&g_timeouts, 1u);
}

// ---- producer (warp 0): one ring word = 32 consecutive entries of the walk ------------------------------------
// list entry e is sorted index first + DIR * e (backward walk: DIR = -1)
template <int DIR>
__device__ __forceinline__ int load_id(int word, int first, int n_entries, const int32_t* __restrict__ vals, int lane) {
    const int e = 32 * word + lane;
    return (e < n_entries) ? __ldg(vals + first + DIR * e) : -1;
}

// Issue word `word` (its ids are in `id`, one per lane): every lane copies its 48-byte record into the ring with one
// bulk asynchronous copy that completes on the word's mbarrier; lane 0 arms the barrier with the byte count.
// BULK = true: one cp.async.bulk (48 B, bulk-copy engine, completes by transaction bytes; barrier count 1);
// BULK = false: three 16-byte cp.async per lane + cp.async.mbarrier.arrive.noinc (barrier count 32, one per lane).
template <int NW, bool WITH_ID, bool BULK>
__device__ __forceinline__ void issue_word(TileRing<NW, WITH_ID>& st, int word, int id, int n_entries,
                                           const float4* __restrict__ rec, int lane) {
    const unsigned slotw = (unsigned)word & (NW - 1);
    const unsigned slot = slotw * 32u + lane;
    if (BULK) {
        // the slot's previous word was read through the generic proxy; order those reads before the async-proxy writes
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (id >= 0) {
            if (WITH_ID) st.id[slot] = id;
            bulk_g2s(&st.rec[3 * slot], rec + 3 * (int64_t)id, 48u, &st.full[slotw]);
        }
        __syncwarp(); // the id stores of all lanes precede lane 0's (releasing) arrive
        if (lane == 0) {
            const int n_valid = min(32, n_entries - 32 * word);
            mbar_arrive_expect_tx(&st.full[slotw], 48u * (uint32_t)n_valid);
        }
    } else {
        if (id >= 0) {
            if (WITH_ID) st.id[slot] = id;
            const float4* r = rec + 3 * (int64_t)id;
            cp_async16(&st.rec[3 * slot], r);
            cp_async16(&st.rec[3 * slot + 1], r + 1);
            cp_async16(&st.rec[3 * slot + 2], r + 2);
        }
        cp_async_arrive_noinc(&st.full[slotw]); // every lane, with or without an entry: 32 arrivals per phase
        __syncwarp();
    }
    if (lane == 0) st_volatile_shared(&st.issued, word + 1);
}

// Hit masks of the 32 entries of ring word slot base `wsl` (lane L tests entry L) against the warp's 32 quads, written
// to hits[wsl + q].  Footprint test: squared half extents + Gershgorin circle, padded, conservative; per quad row r the
// three tests (x extent, y extent, circle) collapse into one threshold on ex2: ex2 <= min(hx2, R2 - ey2[r]) if
// ey2[r] <= hy2, never otherwise.
__device__ __forceinline__ void publish_hits(uint32_t* hits, bool have, const float4& g0, const float4& g1, unsigned wsl,
                                             float bx0, float by0, int lane) {
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
        if (lane == 0) hits[wsl + q] = m;
    }
}

// warp-uniform walk state of one warp
struct WS {
    int n_words;  // words of the tile's walk
    int pub;      // words this warp has hit masks for: [base, pub) are consumable
    int base;     // this warp's ring base (published to prog[wid])
    int issued;   // warp 0: words issued so far; warp 1: the last value of the producer's counter it has seen
    int next_id;  // producer only: this lane's Gaussian id of word `issued` (prefetched)
    int other;    // producer only: the last value of warp 1's ring base it has seen
    bool dead;    // a bounded wait timed out
};

// Publish this warp's hit masks for word s.pub if its records have landed; never blocks.  A word is only waited on
// once the producer has ISSUED it: the producer re-arms a slot's barrier only after some warp has seen the slot's
// previous phase complete, so the barrier is then in exactly this word's phase and the parity test is unambiguous
// (a warp that skips words, or starts in the middle of the walk, has not seen the earlier phases itself).
template <int NW, bool WITH_ID>
__device__ __forceinline__ bool try_publish(TileRing<NW, WITH_ID>& st, WS& s, int wid, int n_entries, float bx0, float by0,
                                            int lane) {
    if (s.pub >= s.n_words || s.pub - s.base >= NW) return false;
    if (s.pub >= s.issued) {
        if (wid == 0) return false;                 // the producer knows its own count
        s.issued = peek(&st.issued, lane);
        if (s.pub >= s.issued) return false;
    }
    const unsigned slotw = (unsigned)s.pub & (NW - 1);
    const uint32_t parity = ((unsigned)s.pub / NW) & 1u;
    if (!__all_sync(FULL, mbar_test_wait(&st.full[slotw], parity))) return false;
    const unsigned wsl = slotw * 32u;
    const bool have = 32 * s.pub + lane < n_entries;
    float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
    if (have) {
        g0 = st.rec[3 * (wsl + lane)];
        g1 = st.rec[3 * (wsl + lane) + 1];
    }
    publish_hits(st.hits[wid], have, g0, g1, wsl, bx0, by0, lane);
    __syncwarp();
    ++s.pub;
    return true;
}

// Producer step of warp 0: issue the next word if its slot has been retired by both warps.
template <int NW, bool WITH_ID, int DIR, bool BULK>
__device__ __forceinline__ bool try_issue(TileRing<NW, WITH_ID>& st, WS& s, int own_base, int first, int n_entries,
                                          const float4* __restrict__ rec, const int32_t* __restrict__ vals, int lane) {
    if (s.issued >= s.n_words) return false;
    if (s.issued >= NW) {
        if (s.issued - own_base >= NW) return false;
        if (s.issued - s.other >= NW) {          // the cached value blocks: look again
            s.other = peek(&st.prog[1], lane);
            if (s.issued - s.other >= NW) return false;
        }
    }
    issue_word<NW, WITH_ID, BULK>(st, s.issued, s.next_id, n_entries, rec, lane);
    ++s.issued;
    s.next_id = load_id<DIR>(s.issued, first, n_entries, vals, lane);
    return true;
}

// Retire: publish this warp's ring base (the producer recycles a slot once both warps are past it).
template <int NW, bool WITH_ID>
__device__ __forceinline__ void publish_base(TileRing<NW, WITH_ID>& st, int wid, int base, int lane) {
    // every lane has CONSUMED (in arithmetic, in program order) what it loaded from the retired word, so those
    // loads are complete; __syncwarp orders the lanes, the volatile store publishes
    __syncwarp();
    if (lane == 0) st_volatile_shared(&st.prog[wid], base);
}

// One service step of the ring, called when something can have changed (a word was retired, the warp is starved, or an
// unpublished word is inside the window): warp 0 issues at most one word, then the warp publishes its hit masks for
// the next word if that one has landed.  Returns whether more service is pending (an unpublished word in the window).
// A STARVED warp (nothing consumable) that could not publish backs off briefly; its waiting is bounded.
template <int NW, bool WITH_ID, int DIR, bool BULK>
__device__ __forceinline__ bool service(TileRing<NW, WITH_ID>& st, WS& s, int wid, bool starved, long long& t_starve,
                                        int first, int n_entries, const float4* __restrict__ rec,
                                        const int32_t* __restrict__ vals, float bx0, float by0, int lane) {
    if (wid == 0) try_issue<NW, WITH_ID, DIR, BULK>(st, s, s.base, first, n_entries, rec, vals, lane);
    const bool published = try_publish<NW, WITH_ID>(st, s, wid, n_entries, bx0, by0, lane);
    if (starved && !published) {
        const long long now = clock64();
        if (t_starve == 0) t_starve = now;
        else if (now - t_starve > WAIT_CYCLES) { s.dead = true; note_timeout(lane); }
        __nanosleep(20);
    } else {
        t_starve = 0;
    }
    return s.pub < s.n_words && s.pub - s.base < NW;
}

// After its own walk, warp 0 keeps the ring fed until warp 1 is finished, then waits for every copy it issued (the
// CTA's shared memory must not be released with bulk copies in flight).
template <int NW, bool WITH_ID, int DIR, bool BULK>
__device__ __forceinline__ void producer_tail(TileRing<NW, WITH_ID>& st, WS& s, int first, int n_entries,
                                              const float4* __restrict__ rec, const int32_t* __restrict__ vals, int lane) {
    const long long t0 = clock64();
    while (s.issued < s.n_words && !s.dead) {
        if (peek(&st.prog[1], lane) == FIN) break;
        if (try_issue<NW, WITH_ID, DIR, BULK>(st, s, FIN, first, n_entries, rec, vals, lane)) continue;
        __nanosleep(64);
        if (clock64() - t0 > WAIT_CYCLES) { s.dead = true; note_timeout(lane); }
    }
    for (int w = max(0, s.issued - NW); w < s.issued; ++w)
        if (!mbar_wait_warp(&st.full[(unsigned)w & (NW - 1)], ((unsigned)w / NW) & 1u)) { note_timeout(lane); break; }
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int NW, bool WITH_ID, bool BULK>
__device__ __forceinline__ void ring_init(TileRing<NW, WITH_ID>& st) {
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NW; ++i) mbar_init(&st.full[i], BULK ? 1u : 32u);
        st.prog[0] = 0;
        st.prog[1] = 0;
        st.issued = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
}

// GEOM = true: tracking (only x, y, conic, z, z^2 gradients are produced).
template <int CH, int NW, bool GEOM, bool BULK>
__global__ void __launch_bounds__(THREADS, 7)
blend_bwd_ts_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                    const int32_t* __restrict__ ranges, const float* __restrict__ alpha,
                    const int32_t* __restrict__ last, const float* __restrict__ v_render,
                    const float* __restrict__ v_alpha, float* __restrict__ v_rec) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using R = TileRing<NW, true>;
    R& st = *reinterpret_cast<R*>(s_raw);
    static_assert((NW & (NW - 1)) == 0, "NW must be a power of two");
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ty = blockIdx.y, half = wid;
    const int tile = ty * cam.tw + blockIdx.x;
    const int qx = lane & 7, qy = lane >> 3;
    const int x0 = blockIdx.x * TILE + 2 * qx, y0 = ty * TILE + 8 * half + 2 * qy;
    const float bx0 = (float)(blockIdx.x * TILE) + 0.5f, by0 = (float)(ty * TILE + 8 * half) + 0.5f;
    const float pxa = (float)x0 + 0.5f, pxb = (float)x0 + 1.5f, pya = (float)y0 + 0.5f, pyb = (float)y0 + 1.5f;
    const int start = ranges[tile];
    const int64_t P = (int64_t)cam.width * cam.height;
    ring_init<NW, true, BULK>(st);

    int first[4]; // first walk position (back to front from the TILE's last contributor) that contributed to pixel p
    float T[4], bdot[4], va_eff[4], vr[4][CH];
    int quad_last = start - 1;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int xi = x0 + (p & 1), yi = y0 + (p >> 1);
        const bool inside = xi < cam.width && yi < cam.height;
        const int64_t pix = inside ? (int64_t)yi * cam.width + xi : 0;
        const int my_last = inside ? last[pix] : start - 1;
        first[p] = my_last; // turned into a walk position below
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
    if (lane == 0) st.warp_last[wid] = warp_last;
    __syncthreads(); // barriers initialised, both halves' last contributors known
    const int tile_last = max(st.warp_last[0], st.warp_last[1]);
    if (tile_last < start) return; // nothing was blended into this tile
    const int n_entries = tile_last - start + 1;
    WS s;
    s.n_words = (n_entries + 31) >> 5;
    s.issued = 0;
    s.next_id = -1;
    s.other = 0;
    s.dead = false;
    // this warp's quads only need the walk from its own last contributor on: skip the words in front of it
    const bool idle = warp_last < start;
    s.base = s.pub = idle ? s.n_words : ((tile_last - warp_last) >> 5);
    if (s.base > 0 || idle) publish_base<NW, true>(st, wid, idle ? FIN : s.base, lane);
    if (wid == 0) {
        s.next_id = load_id<-1>(0, tile_last, n_entries, vals, lane);
        while (try_issue<NW, true, -1, BULK>(st, s, s.base, tile_last, n_entries, rec, vals, lane)) {}
    }
    if (!idle) {
        const int e_first = tile_last - quad_last; // entries in front of it lie behind this quad's last contributor
#pragma unroll
        for (int p = 0; p < 4; ++p) first[p] = tile_last - first[p];
        const uint32_t* hq = st.hits[wid] + lane; // this quad's hit word of ring slot base wsl is hq[wsl]
        auto fetch = [&](int w, unsigned wsl_) -> uint32_t {
            const uint32_t m = hq[wsl_];
            const int cut = e_first - 32 * w;
            return cut <= 0 ? m : (cut >= 32 ? 0u : (m & ~((1u << cut) - 1u)));
        };
        int jq = s.base;
        unsigned wsl = ((unsigned)jq & (NW - 1)) * 32u;
        uint32_t bits = 0u;
        bool fresh = true;   // the cursor's current word has not been fetched yet (it was not published)
        bool poll = true;    // an unpublished word is inside the window
        long long t_starve = 0;
        if (quad_last < start) { jq = FAR; fresh = false; wsl = 0u; }
        while (!s.dead) {
            if (fresh) {
                if (jq < s.pub) { bits = fetch(jq, wsl); fresh = false; }
            } else if (bits == 0u && jq + 1 < s.pub) {
                ++jq;
                wsl = (wsl + 32u) & (32u * NW - 1u);
                bits = fetch(jq, wsl);
            }
            const bool act = bits != 0u;
            const int wmin = __reduce_min_sync(FULL, fresh ? jq : (act ? jq : jq + 1));
            if (wmin >= s.n_words) break;
            const bool starved = wmin >= s.pub; // nothing consumable
            if (!starved) {
            const int sb = act ? __ffs(bits) - 1 : 0;
            bits &= bits - 1u;
            const unsigned slot = wsl + sb;
            const float4 g0s = st.rec[3 * slot];
            const float4 g1s = st.rec[3 * slot + 1];
            const int e = act ? 32 * jq + sb : -1;
            const float dxa = __fsub_rn(g0s.x, pxa), dxb = __fsub_rn(g0s.x, pxb);
            const float dya = __fsub_rn(g0s.y, pya), dyb = __fsub_rn(g0s.y, pyb);
            const float zxa = __fmul_rn(g0s.z, dxa), zxb = __fmul_rn(g0s.z, dxb);
            const float cya = __fmul_rn(__fmul_rn(g1s.x, dya), dya), cyb = __fmul_rn(__fmul_rn(g1s.x, dyb), dyb);
            float sigma[4];
            sigma[0] = __fmaf_rn(__fmaf_rn(g0s.w, dya, zxa), dxa, cya);
            sigma[1] = __fmaf_rn(__fmaf_rn(g0s.w, dya, zxb), dxb, cya);
            sigma[2] = __fmaf_rn(__fmaf_rn(g0s.w, dyb, zxa), dxa, cyb);
            sigma[3] = __fmaf_rn(__fmaf_rn(g0s.w, dyb, zxb), dxb, cyb);
            bool valid[4];
            bool anyv = false;
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                valid[p] = e >= first[p] && sigma[p] >= 0.f && sigma[p] <= g1s.y;
                anyv |= valid[p];
            }
            if (__any_sync(FULL, anyv)) {
                const float4 g2 = st.rec[3 * slot + 2];
                const float opac = g1s.z;
                float Swx = 0.f, Swy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f, Su = 0.f;
                float col[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float dx = (p & 1) ? dxb : dxa, dy = (p >> 1) ? dyb : dya;
                    // branch-free: an invalid pixel runs with sigma = 0 and contributes exact zeros
                    const float vis = __expf(valid[p] ? -sigma[p] : 0.f);
                    const float ov = opac * vis;
                    const float a = fminf(MAX_ALPHA, ov);
                    const float ra = __fdividef(1.0f, 1.0f - a);
                    const float Tn = T[p] * ra; // transmittance in front of this Gaussian
                    T[p] = valid[p] ? Tn : T[p];
                    const float fac = valid[p] ? a * Tn : 0.f;
                    float cdot = g1s.w * vr[p][0];
                    cdot = fmaf(g2.x, vr[p][1], cdot);
                    cdot = fmaf(g2.y, vr[p][2], cdot);
                    cdot = fmaf(g2.z, vr[p][3], cdot);
                    if (CH == 5) cdot = fmaf(g2.w, vr[p][CH - 1], cdot);
                    const float v_a = fmaf(Tn, cdot, ra * (va_eff[p] - bdot[p]));
                    bdot[p] = fmaf(fac, cdot, bdot[p]);
                    const float u = (valid[p] && ov <= MAX_ALPHA) ? vis * v_a : 0.f;
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
                        for (int k = 0; k < 3; ++k) col[k] = fmaf(fac, vr[p][k], col[k]);
                    }
#pragma unroll
                    for (int k = 3; k < CH; ++k) col[k] = fmaf(fac, vr[p][k], col[k]);
                }
                if (anyv) {
                    const float ca = 2.0f * g0s.z, cb = g0s.w, cc = 2.0f * g1s.x;
                    const float v0 = fmaf(ca, Swx, cb * Swy), v1 = fmaf(cb, Swx, cc * Swy);
                    float* dst = v_rec + (int64_t)st.id[slot] * REC;
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
            } // !starved
            bool retired = false;
            if (wmin > s.base && s.base < s.pub) { ++s.base; publish_base<NW, true>(st, wid, s.base, lane); retired = true; }
            if (starved || retired || poll)
                poll = service<NW, true, -1, BULK>(st, s, wid, starved, t_starve, tile_last, n_entries, rec, vals, bx0, by0, lane);
        }
        publish_base<NW, true>(st, wid, FIN, lane);
    }
    if (wid == 0) producer_tail<NW, true, -1, BULK>(st, s, tile_last, n_entries, rec, vals, lane);
}

// Forward, same organisation.  Reference: gsplat RasterizeToPixels3DGSFwd.cu:17-188.  The per-pixel body is
// predicated, not branched; the contribute / saturate decisions are those of blend_lq.cu::blend_fwd_lq_kernel.
template <int CH, int NW, bool BULK>
__global__ void __launch_bounds__(THREADS, 7)
blend_fwd_ts_kernel(DevCam cam, const float4* __restrict__ rec, const int32_t* __restrict__ vals,
                    const int32_t* __restrict__ ranges, float* __restrict__ render, float* __restrict__ alpha,
                    int32_t* __restrict__ last) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    using R = TileRing<NW, false>;
    R& st = *reinterpret_cast<R*>(s_raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ty = blockIdx.y, half = wid;
    const int tile = ty * cam.tw + blockIdx.x;
    const int qx = lane & 7, qy = lane >> 3;
    const int x0 = blockIdx.x * TILE + 2 * qx, y0 = ty * TILE + 8 * half + 2 * qy;
    const float bx0 = (float)(blockIdx.x * TILE) + 0.5f, by0 = (float)(ty * TILE + 8 * half) + 0.5f;
    const float pxa = (float)x0 + 0.5f, pxb = (float)x0 + 1.5f, pya = (float)y0 + 0.5f, pyb = (float)y0 + 1.5f;
    const int start = ranges[tile], end = ranges[tile + 1];
    const int n_entries = end - start;
    ring_init<NW, false, BULK>(st);

    float T[4], acc[4][CH];
    int last_idx[4];
    bool done[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        T[p] = 1.0f;
        last_idx[p] = start - 1;
        done[p] = !(x0 + (p & 1) < cam.width && y0 + (p >> 1) < cam.height);
#pragma unroll
        for (int k = 0; k < CH; ++k) acc[p][k] = 0.f;
    }
    const bool quad_done = done[0] && done[1] && done[2] && done[3];
    __syncthreads(); // barriers initialised

    if (n_entries > 0) {
        WS s;
        s.n_words = (n_entries + 31) >> 5;
        s.issued = 0;
        s.next_id = -1;
        s.other = 0;
        s.dead = false;
        s.base = s.pub = 0;
        const bool idle = __all_sync(FULL, quad_done); // half tile entirely outside the image
        if (idle) publish_base<NW, false>(st, wid, FIN, lane);
        if (wid == 0) {
            s.next_id = load_id<1>(0, start, n_entries, vals, lane);
            while (try_issue<NW, false, 1, BULK>(st, s, s.base, start, n_entries, rec, vals, lane)) {}
        }
        if (!idle) {
            const uint32_t* hq = st.hits[wid] + lane;
            int jq = 0;
            unsigned wsl = 0u;
            uint32_t bits = 0u;
            bool fresh = true;   // the cursor's current word has not been fetched yet (it was not published)
            bool poll = true;    // an unpublished word is inside the window
            long long t_starve = 0;
            if (quad_done) { jq = FAR; fresh = false; }
            while (!s.dead) {
                if (fresh) {
                    if (jq < s.pub) { bits = hq[wsl]; fresh = false; }
                } else if (bits == 0u && jq + 1 < s.pub) {
                    ++jq;
                    wsl = (wsl + 32u) & (32u * NW - 1u);
                    bits = hq[wsl];
                }
                const bool act = bits != 0u;
                const int wmin = __reduce_min_sync(FULL, fresh ? jq : (act ? jq : jq + 1));
                if (wmin >= s.n_words) break; // list exhausted, or every quad saturated (FAR)
                const bool starved = wmin >= s.pub;
                if (!starved) {
                const int sb = act ? __ffs(bits) - 1 : 0;
                bits &= bits - 1u;
                const unsigned slot = wsl + sb;
                const float4 g0s = st.rec[3 * slot];
                const float4 g1s = st.rec[3 * slot + 1];
                const float dxa = __fsub_rn(g0s.x, pxa), dxb = __fsub_rn(g0s.x, pxb);
                const float dya = __fsub_rn(g0s.y, pya), dyb = __fsub_rn(g0s.y, pyb);
                const float zxa = __fmul_rn(g0s.z, dxa), zxb = __fmul_rn(g0s.z, dxb);
                const float cya = __fmul_rn(__fmul_rn(g1s.x, dya), dya), cyb = __fmul_rn(__fmul_rn(g1s.x, dyb), dyb);
                float sigma[4];
                sigma[0] = __fmaf_rn(__fmaf_rn(g0s.w, dya, zxa), dxa, cya);
                sigma[1] = __fmaf_rn(__fmaf_rn(g0s.w, dya, zxb), dxb, cya);
                sigma[2] = __fmaf_rn(__fmaf_rn(g0s.w, dyb, zxa), dxa, cyb);
                sigma[3] = __fmaf_rn(__fmaf_rn(g0s.w, dyb, zxb), dxb, cyb);
                bool hit[4];
                bool anyh = false;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    hit[p] = act && !done[p] && sigma[p] >= 0.f && sigma[p] <= g1s.y;
                    anyh |= hit[p];
                }
                if (__any_sync(FULL, anyh)) {
                    const float4 g2 = st.rec[3 * slot + 2];
                    const int idx = start + 32 * jq + sb;
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        const float a = fminf(MAX_ALPHA, g1s.z * __expf(hit[p] ? -sigma[p] : 0.f));
                        const float nT = T[p] * (1.0f - a);
                        const bool sat = hit[p] && nT <= T_CUTOFF;
                        const bool contrib = hit[p] && !sat;
                        done[p] = done[p] || sat;
                        const float vis = contrib ? a * T[p] : 0.f;
                        acc[p][0] = fmaf(g1s.w, vis, acc[p][0]);
                        acc[p][1] = fmaf(g2.x, vis, acc[p][1]);
                        acc[p][2] = fmaf(g2.y, vis, acc[p][2]);
                        acc[p][3] = fmaf(g2.z, vis, acc[p][3]);
                        if (CH == 5) acc[p][CH - 1] = fmaf(g2.w, vis, acc[p][CH - 1]);
                        last_idx[p] = contrib ? idx : last_idx[p];
                        T[p] = contrib ? nT : T[p];
                    }
                    if (done[0] && done[1] && done[2] && done[3]) { jq = FAR; bits = 0u; } // quad saturated
                }
                } // !starved
                bool retired = false;
                if (wmin > s.base && s.base < s.pub) { ++s.base; publish_base<NW, false>(st, wid, s.base, lane); retired = true; }
                if (starved || retired || poll)
                    poll = service<NW, false, 1, BULK>(st, s, wid, starved, t_starve, start, n_entries, rec, vals, bx0, by0, lane);
            }
            publish_base<NW, false>(st, wid, FIN, lane);
        }
        if (wid == 0) producer_tail<NW, false, 1, BULK>(st, s, start, n_entries, rec, vals, lane);
    }
    const int64_t P = (int64_t)cam.width * cam.height;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int xi = x0 + (p & 1), yi = y0 + (p >> 1);
        if (xi < cam.width && yi < cam.height) {
            const int64_t pix = (int64_t)yi * cam.width + xi;
#pragma unroll
            for (int k = 0; k < CH; ++k) render[k * P + pix] = (k < 3) ? fmaf(T[p], cam.bg[k], acc[p][k]) : acc[p][k];
            alpha[pix] = 1.0f - T[p];
            last[pix] = last_idx[p];
        }
    }
}

template <int CH, int NW, bool BULK>
int launch_fwd(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges, float* render, float* alpha,
               int32_t* last, cudaStream_t stream) {
    constexpr size_t smem = sizeof(TileRing<NW, false>);
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        ISOGS_CUDA(cudaFuncSetAttribute(blend_fwd_ts_kernel<CH, NW, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_fwd_ts_kernel<CH, NW, BULK><<<grid, THREADS, smem, stream>>>(dc, (const float4*)rec, vals, ranges, render, alpha, last);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

template <int CH, int NW, bool GEOM, bool BULK>
int launch_bwd(const DevCam& dc, const float* rec, const int32_t* vals, const int32_t* ranges, const float* alpha,
               const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec, cudaStream_t stream) {
    constexpr size_t smem = sizeof(TileRing<NW, true>);
    static thread_local int attr_dev = -1;
    int dev = 0;
    ISOGS_CUDA(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        ISOGS_CUDA(cudaFuncSetAttribute(blend_bwd_ts_kernel<CH, NW, GEOM, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
        attr_dev = dev;
    }
    dim3 grid(dc.tw, dc.th);
    blend_bwd_ts_kernel<CH, NW, GEOM, BULK><<<grid, THREADS, smem, stream>>>(dc, (const float4*)rec, vals, ranges, alpha, last,
                                                                        v_render, v_alpha, v_rec);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

} // namespace ts

// called by isogs_blend_fwd / isogs_blend_bwd / isogs_blend_bwd_geom (blend.cu)
int launch_blend_fwd_ts(const DevCam& dc, int ch, bool bulk, const float* rec, const int32_t* vals, const int32_t* ranges,
                        float* render, float* alpha, int32_t* last, cudaStream_t stream) {
#define TSF(CHN, B) ts::launch_fwd<CHN, 16, B>(dc, rec, vals, ranges, render, alpha, last, stream)
    if (bulk) return ch == 4 ? TSF(4, true) : TSF(5, true);
    return ch == 4 ? TSF(4, false) : TSF(5, false);
#undef TSF
}

int launch_blend_bwd_ts(const DevCam& dc, int ch, bool geom, bool bulk, const float* rec, const int32_t* vals, const int32_t* ranges,
                        const float* alpha, const int32_t* last, const float* v_render, const float* v_alpha, float* v_rec,
                        cudaStream_t stream) {
#define TS(CHN, G, B) ts::launch_bwd<CHN, 16, G, B>(dc, rec, vals, ranges, alpha, last, v_render, v_alpha, v_rec, stream)
    if (bulk) {
        if (geom) return ch == 4 ? TS(4, true, true) : TS(5, true, true);
        return ch == 4 ? TS(4, false, true) : TS(5, false, true);
    }
    if (geom) return ch == 4 ? TS(4, true, false) : TS(5, true, false);
    return ch == 4 ? TS(4, false, false) : TS(5, false, false);
#undef TS
}

} // namespace isogs

// bounded-spin timeouts since the last call (0 in a healthy run); resets the counter
extern "C" unsigned int isogs_debug_blend_timeouts(void) {
    unsigned int v = 0, z = 0;
    cudaMemcpyFromSymbol(&v, isogs::ts::g_timeouts, sizeof(v));
    cudaMemcpyToSymbol(isogs::ts::g_timeouts, &z, sizeof(z));
    return v;
}
