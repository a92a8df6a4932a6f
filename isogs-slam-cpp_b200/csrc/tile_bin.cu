// tile_bin.cu -- north_star (2): duplicateWithKeys emits (tile | depth) 64-bit keys, a CUB onesweep
// radix sort orders them, identifyTileRanges builds per-tile ranges.  Integer work, bit-exact
// against the oracle.  Reference: gsplat IntersectTile.cu:87-113,296-339,209-257.
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

inline int sort_end_bit(int n_tiles) {
    int bits = 0;
    while ((1ll << bits) <= (long long)n_tiles) ++bits; // floor(log2(n))+1, IntersectTile.cu:151
    return 32 + bits;
}

} // namespace isogs

using namespace isogs;

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
        *sort_bytes = need + 256;
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
    const int block = 256;
    emit_keys_kernel<<<(unsigned)((N + block - 1) / block), block, 0, stream>>>(
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
