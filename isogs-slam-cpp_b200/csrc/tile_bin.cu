// tile_bin.cu -- north_star (2): duplicateWithKeys emits (tile | depth) 64-bit keys, a CUB onesweep
// radix sort orders them, identifyTileRanges builds per-tile ranges.  Integer work, bit-exact
// against the oracle.  Reference: gsplat IntersectTile.cu:87-113,296-339,209-257.
#include <cub/block/block_merge_sort.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>


#include "common.cuh"

namespace isogs {

// one thread per Gaussian; emission order inside a Gaussian is row-major over its tile rect,
// so equal keys keep Gaussian-index order after the stable LSD sort.
__global__ void __launch_bounds__(256)
emit_keys_kernel(int tw, int th, int64_t N, const float4* __restrict__ rec, const int2* __restrict__ radii,
                 const int32_t* __restrict__ cum, uint64_t* __restrict__ keys, int32_t* __restrict__ vals) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int2 rad = radii[i];
    if (rad.x <= 0 || rad.y <= 0) return;
    float4 r0 = rec[3 * i];
    float depth = rec[3 * i + 2].z;
    Rect q = tile_rect(r0.x, r0.y, rad.x, rad.y, tw, th);
    int64_t cur = (i == 0) ? 0 : (int64_t)cum[i - 1];
    const uint64_t depth_bits = (uint64_t)__float_as_uint(depth);
    for (int y = q.y0; y < q.y1; ++y)
        for (int x = q.x0; x < q.x1; ++x) {
            uint64_t tile_id = (uint64_t)(y * tw + x);
            keys[cur] = (tile_id << 32) | depth_bits;
            vals[cur] = (int32_t)i;
            ++cur;
        }
}

__global__ void __launch_bounds__(256)
tile_ranges_kernel(int64_t M, int n_tiles, const uint64_t* __restrict__ keys, int32_t* __restrict__ ranges) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M) return;
    int cur = (int)(keys[idx] >> 32);
    if (idx == 0) {
        for (int t = 0; t <= cur; ++t) ranges[t] = 0;
    } else {
        int prev = (int)(keys[idx - 1] >> 32);
        for (int t = prev + 1; t <= cur; ++t) ranges[t] = (int32_t)idx;
    }
    if (idx == M - 1) {
        for (int t = cur + 1; t <= n_tiles; ++t) ranges[t] = (int32_t)M;
    }
}

// ---- fast path: counting sort by tile, then a per-tile sort of (depth bits, Gaussian id) --------
// The (tile | depth) key has only ~12 tile bits: a histogram + scan over tiles IS identifyTileRanges,
// and a scatter with per-tile cursors buckets the pairs.  Inside a tile, sorting the 64-bit value
// (depth_bits << 32 | gaussian_id) reproduces exactly the order of the stable LSD radix sort on
// (tile | depth): equal depths keep Gaussian-index (= emission) order.

__global__ void __launch_bounds__(256)
tile_count_kernel(int tw, int th, int64_t N, const float4* __restrict__ rec, const int2* __restrict__ radii,
                  int32_t* __restrict__ counts) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int2 rad = radii[i];
    if (rad.x <= 0 || rad.y <= 0) return;
    float4 r0 = rec[3 * i];
    Rect q = tile_rect(r0.x, r0.y, rad.x, rad.y, tw, th);
    for (int y = q.y0; y < q.y1; ++y)
        for (int x = q.x0; x < q.x1; ++x) atomicAdd(counts + (y * tw + x), 1);
}

// single CTA exclusive scan of counts[0..n) -> ranges[0..n]; also zeroes the cursors.
// capacity >= 0 (sync-free pipeline): `ranges` is clamped to the capacity of the intersection
// buffers, the unclamped offsets go to `true_ranges` (the scatter places by them and drops what
// does not fit) and totals[0] = n_isects, totals[1] = 1 if it exceeds the capacity.
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int n, const int32_t* __restrict__ counts, int32_t* __restrict__ ranges,
                 int32_t* __restrict__ cursors, int32_t capacity, int32_t* __restrict__ true_ranges,
                 int32_t* __restrict__ totals, int big, int32_t* __restrict__ n_big) {
    __shared__ int32_t s_warp[32];
    __shared__ int32_t s_carry;
    __shared__ int32_t s_big;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) { s_carry = 0; s_big = 0; }
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int32_t c = (i < n) ? counts[i] : 0;
        if (c > big) atomicAdd(&s_big, 1); // tiles the block merge sort does not take
        int32_t v = c;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            int32_t u = __shfl_up_sync(0xffffffffu, v, off);
            if (lane >= off) v += u;
        }
        if (lane == 31) s_warp[wid] = v;
        __syncthreads();
        if (wid == 0) {
            int32_t w = s_warp[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                int32_t u = __shfl_up_sync(0xffffffffu, w, off);
                if (lane >= off) w += u;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int32_t carry = s_carry;
        const int32_t incl = v + (wid > 0 ? s_warp[wid - 1] : 0) + carry;
        if (i < n) {
            const int32_t ex = incl - c;
            ranges[i] = capacity >= 0 ? min(ex, capacity) : ex;
            if (true_ranges) true_ranges[i] = ex;
            cursors[i] = 0;
        }
        __syncthreads();
        if (tid == 1023) s_carry = incl;
        __syncthreads();
    }
    if (tid == 0) {
        if (n_big) *n_big = s_big;
        const int32_t total = s_carry;
        ranges[n] = capacity >= 0 ? min(total, capacity) : total;
        if (true_ranges) true_ranges[n] = total;
        if (totals) {
            // the overflow word is STICKY (only the host clears it, after growing the buffers): while it is set the
            // step's Adam / best-pose kernels are no-ops, and totals[0] keeps the largest count seen
            const bool over = capacity >= 0 && total > capacity;
            if (over) { totals[0] = totals[1] ? max(totals[0], total) : total; totals[1] = 1; }
            else if (!totals[1]) totals[0] = total;
        }
    }
}

__global__ void __launch_bounds__(256)
tile_scatter_kernel(int tw, int th, int64_t N, const float4* __restrict__ rec, const int2* __restrict__ radii,
                    const int32_t* __restrict__ ranges, int32_t* __restrict__ cursors,
                    uint64_t* __restrict__ bucket, int32_t capacity) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int2 rad = radii[i];
    if (rad.x <= 0 || rad.y <= 0) return;
    float4 r0 = rec[3 * i];
    const float depth = rec[3 * i + 2].z;
    Rect q = tile_rect(r0.x, r0.y, rad.x, rad.y, tw, th);
    const uint64_t v = ((uint64_t)__float_as_uint(depth) << 32) | (uint64_t)(uint32_t)i;
    for (int y = q.y0; y < q.y1; ++y)
        for (int x = q.x0; x < q.x1; ++x) {
            const int t = y * tw + x;
            const int slot = ranges[t] + atomicAdd(cursors + t, 1);
            if (slot < capacity) bucket[slot] = v;
        }
}

// Block-privatised versions: neighbouring Gaussians hit the same few tiles, so global atomics on
// counts[tile] serialise in L2 (measured 74 + 89 us).  Each CTA first bins into shared memory and
// then issues ONE global atomic per touched tile.
constexpr int BIN_THREADS = 512;
constexpr int BIN_KEEP = 4; // cells per Gaussian ranked through shared memory; the rest go direct

__global__ void __launch_bounds__(BIN_THREADS)
tile_count_smem_kernel(int tw, int th, int n_tiles, int64_t N, const float4* __restrict__ rec,
                       const int2* __restrict__ radii, int32_t* __restrict__ counts) {
    extern __shared__ int32_t s_bins[];
    for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) s_bins[b] = 0;
    __syncthreads();
    int64_t i = (int64_t)blockIdx.x * BIN_THREADS + threadIdx.x;
    if (i < N) {
        int2 rad = radii[i];
        if (rad.x > 0 && rad.y > 0) {
            float4 r0 = rec[3 * i];
            Rect q = tile_rect(r0.x, r0.y, rad.x, rad.y, tw, th);
            for (int y = q.y0; y < q.y1; ++y)
                for (int x = q.x0; x < q.x1; ++x) atomicAdd(s_bins + (y * tw + x), 1);
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) {
        const int32_t c = s_bins[b];
        if (c) atomicAdd(counts + b, c);
    }
}

__global__ void __launch_bounds__(BIN_THREADS)
tile_scatter_smem_kernel(int tw, int th, int n_tiles, int64_t N, const float4* __restrict__ rec,
                         const int2* __restrict__ radii, const int32_t* __restrict__ ranges,
                         int32_t* __restrict__ cursors, uint64_t* __restrict__ bucket, int32_t capacity) {
    extern __shared__ int32_t s_bins[];
    for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) s_bins[b] = 0;
    __syncthreads();
    int64_t i = (int64_t)blockIdx.x * BIN_THREADS + threadIdx.x;
    Rect q = {0, 0, 0, 0};
    uint64_t v = 0;
    int rank[BIN_KEEP];
    if (i < N) {
        int2 rad = radii[i];
        if (rad.x > 0 && rad.y > 0) {
            float4 r0 = rec[3 * i];
            q = tile_rect(r0.x, r0.y, rad.x, rad.y, tw, th);
            v = ((uint64_t)__float_as_uint(rec[3 * i + 2].z) << 32) | (uint64_t)(uint32_t)i;
        }
    }
    const int w = q.x1 - q.x0, ncell = w * (q.y1 - q.y0);
#pragma unroll
    for (int c = 0; c < BIN_KEEP; ++c)
        if (c < ncell) rank[c] = atomicAdd(s_bins + ((q.y0 + c / w) * tw + q.x0 + c % w), 1);
    __syncthreads();
    for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) {
        const int32_t c = s_bins[b];
        if (c) s_bins[b] = atomicAdd(cursors + b, c); // base of this CTA's block inside the tile bucket
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < BIN_KEEP; ++c)
        if (c < ncell) {
            const int t = (q.y0 + c / w) * tw + q.x0 + c % w;
            const int slot = ranges[t] + s_bins[t] + rank[c];
            if (slot < capacity) bucket[slot] = v;
        }
    for (int c = BIN_KEEP; c < ncell; ++c) {
        const int t = (q.y0 + c / w) * tw + q.x0 + c % w;
        const int slot = ranges[t] + atomicAdd(cursors + t, 1);
        if (slot < capacity) bucket[slot] = v;
    }
}

// the same scatter fed by the 16-byte (tile rect, depth bits) rows preprocess_fwd wrote: coalesced 16 B per Gaussian
// instead of 28 B gathered out of the 56-byte (rec, radii) rows
__global__ void __launch_bounds__(BIN_THREADS)
tile_scatter_rect_kernel(int tw, int n_tiles, int64_t N, const uint4* __restrict__ rects,
                         const int32_t* __restrict__ ranges, int32_t* __restrict__ cursors, uint64_t* __restrict__ bucket,
                         int32_t capacity, int use_bins) {
    extern __shared__ int32_t s_bins[];
    if (use_bins) {
        for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) s_bins[b] = 0;
        __syncthreads();
    }
    int64_t i = (int64_t)blockIdx.x * BIN_THREADS + threadIdx.x;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    uint64_t v = 0;
    if (i < N) {
        const uint4 r = rects[i];
        x0 = (int)(r.x & 0xffffu); x1 = (int)(r.x >> 16); y0 = (int)(r.y & 0xffffu); y1 = (int)(r.y >> 16);
        v = ((uint64_t)r.z << 32) | (uint64_t)(uint32_t)i;
    }
    const int w = x1 - x0, ncell = w * (y1 - y0);
    int rank[BIN_KEEP];
    const int keep = use_bins ? BIN_KEEP : 0;
#pragma unroll
    for (int c = 0; c < BIN_KEEP; ++c)
        if (c < keep && c < ncell) rank[c] = atomicAdd(s_bins + ((y0 + c / w) * tw + x0 + c % w), 1);
    if (use_bins) {
        __syncthreads();
        for (int b = threadIdx.x; b < n_tiles; b += BIN_THREADS) {
            const int32_t c = s_bins[b];
            if (c) s_bins[b] = atomicAdd(cursors + b, c); // base of this CTA's block inside the tile bucket
        }
        __syncthreads();
    }
#pragma unroll
    for (int c = 0; c < BIN_KEEP; ++c)
        if (c < keep && c < ncell) {
            const int t = (y0 + c / w) * tw + x0 + c % w;
            const int slot = ranges[t] + s_bins[t] + rank[c];
            if (slot < capacity) bucket[slot] = v;
        }
    for (int c = keep; c < ncell; ++c) {
        const int t = (y0 + c / w) * tw + x0 + c % w;
        const int slot = ranges[t] + atomicAdd(cursors + t, 1);
        if (slot < capacity) bucket[slot] = v;
    }
}

constexpr int SORT_SMEM_MAX = 8192; // entries per tile sorted in shared memory (64 KB)

// Bitonic sorting network over n (ANY length) 64-bit values, in shared or global memory.  The
// normalised form is used (first step of every merge pairs i with i ^ (k-1), all comparators
// ascending), so slots >= n behave as +inf without being stored.  Threads iterate over compare-
// exchange PAIRS (no idle half); 32 consecutive pairs live in one aligned 64-element block for spans
// <= 32, so those steps only need a warp-level barrier.
__device__ __forceinline__ void cmpx(uint64_t* a, int i, int ixj, int n) {
    if (ixj < n) {
        const uint64_t x = a[i], y = a[ixj];
        if (x > y) { a[i] = y; a[ixj] = x; }
    }
}
__device__ __forceinline__ void bitonic_sort(uint64_t* a, int n, int tid, int nthreads) {
    int n_pad = 2;
    while (n_pad < n) n_pad <<= 1;
    const int half = n_pad >> 1;
    for (int k = 2; k <= n_pad; k <<= 1) {
        const int h = k >> 1;
        for (int p = tid; p < half; p += nthreads) {
            const int i = ((p & ~(h - 1)) << 1) | (p & (h - 1));
            cmpx(a, i, i ^ (k - 1), n);
        }
        if (h > 32 || h == 1) __syncthreads();
        else __syncwarp();
        for (int j = h >> 1; j > 0; j >>= 1) {
            for (int p = tid; p < half; p += nthreads) {
                const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
                cmpx(a, i, i | j, n);
            }
            if (j > 32 || j == 1) __syncthreads();
            else __syncwarp();
        }
    }
}

// one CTA per tile: sort the tile's bucket by (depth bits, gaussian id); emit ids and (optionally)
// the reference-format 64-bit keys.  Segments longer than SORT_SMEM_MAX are sorted in place in
// global memory (padded slots are not available there, so an odd-even transposition is used).
// Launched twice: (lo=0, hi=SORT_SMEM_SMALL] with 16 KB of shared memory (the common case, high
// occupancy) and (SORT_SMEM_SMALL, inf) with 64 KB; a CTA whose tile is outside its range exits.
constexpr int SORT_SMEM_SMALL = 2048;
__global__ void __launch_bounds__(256)
tile_sort_kernel(int lo, int hi, int n_tiles, const int32_t* __restrict__ n_big, const int32_t* __restrict__ ranges,
                 uint64_t* __restrict__ bucket, uint64_t* __restrict__ keys_sorted, int32_t* __restrict__ vals_sorted) {
    extern __shared__ uint64_t s_a[];
    if (n_big != nullptr && *n_big == 0) return; // no tile is longer than `lo` (counted by tile_scan_kernel)
    const int tid = threadIdx.x;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int start = ranges[tile], n = ranges[tile + 1] - start;
    if (n <= lo || n > hi) continue;
    uint64_t* seg = bucket + start;
    const bool in_smem = n <= SORT_SMEM_MAX;
    uint64_t* a = in_smem ? s_a : seg; // rare: longer segments are sorted in place in global memory
    if (in_smem) {
        for (int i = tid; i < n; i += 256) s_a[i] = seg[i];
    }
    __syncthreads();
    if (n > 1) bitonic_sort(a, n, tid, 256);
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        const uint64_t v = a[i];
        vals_sorted[start + i] = (int32_t)(uint32_t)v;
        if (keys_sorted) keys_sorted[start + i] = ((uint64_t)tile << 32) | (v >> 32);
    }
    __syncthreads();
    } // tile loop
}

constexpr int SORT_REG_MAX = 2048;

// ---- per-tile sort for tiles with <= 2048 intersections (the common case): a block-wide merge sort
// (cub::BlockMergeSort: thread t sorts its E consecutive keys, then log2(128) merge-path rounds through shared
// memory).  Measured against a hand-written 55-stage register / shuffle bitonic network on the same buckets (round
// 2, replica shape): 65 us against 76 us -- the network was bound by the ALU / shuffle pipes (46 M warp
// instructions), the merge sort executes about half as many.  Block size (same shape, whole binning stage): 256
// threads 0.096 ms, 128 threads 0.074 ms, 64 threads 0.085 ms -- the merge rounds are barrier / shared-memory latency
// bound, 128 threads x E <= 8 keys halves the threads waiting at each of one round fewer.  A finer ladder of E, or
// separate launches per shared-memory size, changed nothing. -------------------------------------------------
struct Less64 {
    __device__ __forceinline__ bool operator()(const uint64_t& a, const uint64_t& b) const { return a < b; }
};

template <int T, int E>
__device__ __forceinline__ void tile_sort_merge(int tile, int start, int n, const uint64_t* __restrict__ seg, void* smem,
                                                uint64_t* __restrict__ keys_sorted, int32_t* __restrict__ vals_sorted, int tid) {
    using Sort = cub::BlockMergeSort<uint64_t, T, E>;
    uint64_t v[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const int i = tid * E + r;
        v[r] = i < n ? seg[i] : ~0ull;
    }
    Sort(*reinterpret_cast<typename Sort::TempStorage*>(smem)).Sort(v, Less64(), n, ~0ull);
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const int i = tid * E + r;
        if (i < n) {
            vals_sorted[start + i] = (int32_t)(uint32_t)v[r];
            if (keys_sorted) keys_sorted[start + i] = ((uint64_t)tile << 32) | (v[r] >> 32);
        }
    }
}

constexpr int SORT_THREADS = 128;

__global__ void __launch_bounds__(SORT_THREADS)
tile_sort_merge_kernel(const int32_t* __restrict__ ranges, const uint64_t* __restrict__ bucket,
                       uint64_t* __restrict__ keys_sorted, int32_t* __restrict__ vals_sorted) {
    constexpr int T = SORT_THREADS;
    __shared__ __align__(16) unsigned char s_tmp[sizeof(typename cub::BlockMergeSort<uint64_t, T, SORT_REG_MAX / T>::TempStorage)];
    const int tile = blockIdx.x, tid = threadIdx.x;
    const int start = ranges[tile], n = ranges[tile + 1] - start;
    if (n <= 0 || n > SORT_REG_MAX) return;
    const uint64_t* seg = bucket + start;
#define TSM(E) tile_sort_merge<T, E>(tile, start, n, seg, s_tmp, keys_sorted, vals_sorted, tid)
    if (n <= 2 * T) TSM(2);
    else if (n <= 4 * T) TSM(4);
    else if (n <= 5 * T) TSM(5);
    else if (n <= 6 * T) TSM(6);
    else if (n <= 7 * T) TSM(7);
    else if (n <= 8 * T) TSM(8);
    else if (n <= 12 * T) TSM(12);
    else TSM(16);
#undef TSM
}

inline int sort_end_bit(int n_tiles) {
    int bits = 0;
    while ((1ll << bits) <= (long long)n_tiles) ++bits; // floor(log2(n))+1, IntersectTile.cu:151
    return 32 + bits;
}

} // namespace isogs

using namespace isogs;

// counting sort by tile + per-tile sort (see the kernels above); scratch = counts | cursors | true ranges.
// clamp=false: `capacity` is the exact intersection count M.  clamp=true (sync-free pipeline):
// `capacity` is the size of the intersection buffers, M is not known on the host; totals (DEVICE
// int32[2]) receives {M, overflow flag}.
static int counting_bin(int tw, int th, int64_t N, int32_t capacity, bool clamp, const float* rec,
                        const int32_t* radii, uint64_t* bucket, uint64_t* keys_sorted, int32_t* vals_sorted,
                        int32_t* tile_ranges, void* sort_temp, size_t sort_temp_bytes, int32_t* totals,
                        cudaStream_t stream, const uint32_t* tile_rects = nullptr, bool counts_ready = false) {
    const int n_tiles = tw * th;
    const size_t need_scratch = sizeof(int32_t) * (3 * (size_t)(n_tiles + 1) + 1);
    if (need_scratch > sort_temp_bytes) { set_error("sort temp too small for %d tiles", n_tiles); return ISOGS_E_ARG; }
    int32_t* counts = (int32_t*)sort_temp;
    int32_t* cursors = counts + (n_tiles + 1);
    int32_t* true_ranges = cursors + (n_tiles + 1);
    int32_t* n_big = true_ranges + (n_tiles + 1);
    const int block = 256;
    const unsigned gN = (unsigned)((N + block - 1) / block);
    const bool smem_bins = n_tiles <= 11264; // 44 KB of bins fits the default 48 KB dynamic limit
    const unsigned gB = (unsigned)((N + BIN_THREADS - 1) / BIN_THREADS);
    const size_t bin_bytes = sizeof(int32_t) * (size_t)n_tiles;
    if (!counts_ready) { // else: preprocess_fwd already accumulated the per-tile counts into sort_temp (fused step)
        ISOGS_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * (n_tiles + 1), stream));
        if (smem_bins)
            tile_count_smem_kernel<<<gB, BIN_THREADS, bin_bytes, stream>>>(tw, th, n_tiles, N, (const float4*)rec,
                                                                           (const int2*)radii, counts);
        else
            tile_count_kernel<<<gN, block, 0, stream>>>(tw, th, N, (const float4*)rec, (const int2*)radii, counts);
    }
    tile_scan_kernel<<<1, 1024, 0, stream>>>(n_tiles, counts, tile_ranges, cursors, clamp ? capacity : -1,
                                             clamp ? true_ranges : nullptr, totals, SORT_REG_MAX, n_big);
    const int32_t* place = clamp ? true_ranges : tile_ranges;
    if (tile_rects != nullptr)
        tile_scatter_rect_kernel<<<gB, BIN_THREADS, smem_bins ? bin_bytes : 0, stream>>>(
            tw, n_tiles, N, (const uint4*)tile_rects, place, cursors, bucket, capacity, smem_bins ? 1 : 0);
    else if (smem_bins)
        tile_scatter_smem_kernel<<<gB, BIN_THREADS, bin_bytes, stream>>>(tw, th, n_tiles, N, (const float4*)rec,
                                                                         (const int2*)radii, place, cursors, bucket,
                                                                         capacity);
    else
        tile_scatter_kernel<<<gN, block, 0, stream>>>(tw, th, N, (const float4*)rec, (const int2*)radii, place, cursors,
                                                      bucket, capacity);
    static bool attr_set = false;
    if (!attr_set) {
        ISOGS_CUDA(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        SORT_SMEM_MAX * (int)sizeof(uint64_t)));
        attr_set = true;
    }
    static_assert(SORT_REG_MAX == SORT_SMEM_SMALL, "the two sort kernels split the tiles at the same length");
    tile_sort_merge_kernel<<<n_tiles, SORT_THREADS, 0, stream>>>(tile_ranges, bucket, keys_sorted, vals_sorted);
    // the rare long tiles: a small persistent grid that exits at once when tile_scan_kernel counted none
    tile_sort_kernel<<<std::min(n_tiles, 148), 256, SORT_SMEM_MAX * sizeof(uint64_t), stream>>>(
        SORT_SMEM_SMALL, 0x7fffffff, n_tiles, n_big, tile_ranges, bucket, keys_sorted, vals_sorted);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_temp_bytes(int64_t N, int64_t cap, size_t* scan_bytes, size_t* sort_bytes) {
    if (N < 0 || cap < 0 || N >= (1ll << 31) || cap >= (1ll << 31)) { set_error("isogs_temp_bytes: bad sizes"); return ISOGS_E_ARG; }
    if (scan_bytes) {
        size_t need = 0;
        ISOGS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, (const int32_t*)nullptr, (int32_t*)nullptr,
                                                 (int)(N > 0 ? N : 1), (cudaStream_t)0));
        *scan_bytes = need + 256;
    }
    if (sort_bytes) {
        size_t need = 0;
        ISOGS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                                   (const int32_t*)nullptr, (int32_t*)nullptr,
                                                   (int)(cap > 0 ? cap : 1), 0, 64, (cudaStream_t)0));
        *sort_bytes = need + 256 + (1u << 20); /* + counts/cursors of the counting-sort path */
    }
    return ISOGS_OK;
}

extern "C" int isogs_tile_bin(const isogs_camera* cam, int64_t N, int64_t M, const float* rec, const int32_t* radii,
                              const int32_t* cum, uint64_t* keys_unsorted, int32_t* vals_unsorted,
                              uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges, void* sort_temp,
                              size_t sort_temp_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || N < 0 || M < 0 || M >= (1ll << 31)) { set_error("isogs_tile_bin: bad argument"); return ISOGS_E_ARG; }
    const int tw = (cam->width + TILE - 1) / TILE, th = (cam->height + TILE - 1) / TILE;
    const int n_tiles = tw * th;
    if (M == 0 || N == 0) {
        ISOGS_CUDA(cudaMemsetAsync(tile_ranges, 0, sizeof(int32_t) * (n_tiles + 1), stream));
        return ISOGS_OK;
    }
    return counting_bin(tw, th, N, (int32_t)M, /*clamp=*/false, rec, radii, keys_unsorted, keys_sorted, vals_sorted,
                        tile_ranges, sort_temp, sort_temp_bytes, nullptr, stream);
}

// The LITERAL pipeline north_star (2) names -- duplicateWithKeys, a CUB radix sort of the (tile | depth) keys,
// identifyTileRanges -- kept as an independently written cross-check of the counting-sort path (same outputs, bit
// for bit; tests/test_gpu_parity.py) and for callers that want the reference's exact sequence of kernels.
extern "C" int isogs_tile_bin_cub(const isogs_camera* cam, int64_t N, int64_t M, const float* rec, const int32_t* radii,
                                  const int32_t* cum, uint64_t* keys_unsorted, int32_t* vals_unsorted,
                                  uint64_t* keys_sorted, int32_t* vals_sorted, int32_t* tile_ranges, void* sort_temp,
                                  size_t sort_temp_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || N < 0 || M < 0 || M >= (1ll << 31) || !cum) { set_error("isogs_tile_bin_cub: bad argument"); return ISOGS_E_ARG; }
    const int tw = (cam->width + TILE - 1) / TILE, th = (cam->height + TILE - 1) / TILE;
    const int n_tiles = tw * th;
    if (M == 0 || N == 0) {
        ISOGS_CUDA(cudaMemsetAsync(tile_ranges, 0, sizeof(int32_t) * (n_tiles + 1), stream));
        return ISOGS_OK;
    }
    const int block = 256;
    const unsigned gN = (unsigned)((N + block - 1) / block);
    emit_keys_kernel<<<gN, block, 0, stream>>>(
        tw, th, N, (const float4*)rec, (const int2*)radii, cum, keys_unsorted, vals_unsorted);
    ISOGS_CUDA(cudaGetLastError());
    size_t need = 0;
    const int end_bit = sort_end_bit(n_tiles);
    ISOGS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, (const uint64_t*)keys_unsorted, keys_sorted,
                                               (const int32_t*)vals_unsorted, vals_sorted, (int)M, 0, end_bit, stream));
    if (need > sort_temp_bytes) { set_error("sort temp too small: need %zu have %zu", need, sort_temp_bytes); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceRadixSort::SortPairs(sort_temp, sort_temp_bytes, (const uint64_t*)keys_unsorted, keys_sorted,
                                               (const int32_t*)vals_unsorted, vals_sorted, (int)M, 0, end_bit, stream));
    tile_ranges_kernel<<<(unsigned)((M + block - 1) / block), block, 0, stream>>>(M, n_tiles, keys_sorted, tile_ranges);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_tile_bin_async(const isogs_camera* cam, int64_t N, int64_t capacity, const float* rec,
                                    const int32_t* radii, uint64_t* bucket, uint64_t* keys_sorted,
                                    int32_t* vals_sorted, int32_t* tile_ranges, int32_t* totals, void* sort_temp,
                                    size_t sort_temp_bytes, const uint32_t* tile_rects, int counts_ready, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!cam || N < 0 || capacity < 0 || capacity >= (1ll << 31) || !totals) {
        set_error("isogs_tile_bin_async: bad argument");
        return ISOGS_E_ARG;
    }
    const int tw = (cam->width + TILE - 1) / TILE, th = (cam->height + TILE - 1) / TILE;
    if (N == 0) {
        ISOGS_CUDA(cudaMemsetAsync(tile_ranges, 0, sizeof(int32_t) * (tw * th + 1), stream));
        ISOGS_CUDA(cudaMemsetAsync(totals, 0, sizeof(int32_t), stream)); // the sticky overflow word stays
        return ISOGS_OK;
    }
    if (counts_ready && !tile_rects) { set_error("isogs_tile_bin_async: counts_ready needs tile_rects"); return ISOGS_E_ARG; }
    if ((uintptr_t)tile_rects & 15) { set_error("isogs_tile_bin_async: tile_rects must be 16-byte aligned"); return ISOGS_E_ARG; }
    return counting_bin(tw, th, N, (int32_t)capacity, /*clamp=*/true, rec, radii, bucket, keys_sorted, vals_sorted,
                        tile_ranges, sort_temp, sort_temp_bytes, totals, stream, tile_rects, counts_ready != 0);
}
