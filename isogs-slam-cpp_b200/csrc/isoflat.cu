// isoflat.cu -- SURVEY 8(f)-2: the IsoGS geometric terms of the mapping loss (python_ordinary/scripts/splatam.py:
// compute_flat_loss :42-54, compute_iso_surface_loss_sampled :162-262, their use in get_loss :572-675, weights
// flat 50 / iso 2 :733-739; cpp/cuda/kernels/{flat_loss,iso_loss,spatial_hash}.cu).
//
//   flat  = mean_i min_k s_ik                                    s = clamp(exp(log_scales), 1e-5)
//   iso   = mean_q (D_q - 1)^2,  D_q = sum_{j in KNN16(x_q)} sigmoid(o_j) exp(-1/2 d^T R_j S_j^-2 R_j^T d),
//           d = x_q - mu_j, S^-2 = 1 / (s^2 + 1e-8), queries = 8192 random Gaussian centres
//
// The reference finds the 16 nearest centres with torch.cdist + topk in chunks of 128 queries (8192 x N distances
// per iteration; "dominates mapping time / memory" per the authors' log).  Here the search is EXACT but local:
// a uniform grid over the bounding box of the centres is rebuilt every iteration with a counting sort (bbox
// reduction -> cell histogram -> scan -> scatter of (x, y, z, id)), and one warp per query walks Chebyshev shells
// of cells until the 16th best distance is provably inside the searched cube.  Everything stays on the device
// (grid parameters included), so the step keeps running without a host sync.
#include <cub/device/device_scan.cuh>

#include <cfloat>
#include <climits>

#include "common.cuh"

namespace isogs {

constexpr int KNN_K = 16;

struct GridParams { float ox, oy, oz, h, inv_h; int nx, ny, nz; };   // lives in DEVICE memory (8 words)

__device__ __forceinline__ int f2o(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float o2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void bbox_init_kernel(int* bb) {
    if (threadIdx.x < 3) bb[threadIdx.x] = INT_MAX;
    else if (threadIdx.x < 6) bb[threadIdx.x] = INT_MIN;
}

__global__ void __launch_bounds__(256) bbox_kernel(int64_t N, const float* __restrict__ means, int* __restrict__ bb) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float v = means[3 * i + k]; lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); }
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(bb + k, f2o(lo[k])); atomicMax(bb + 3 + k, f2o(hi[k])); }
}

// h = largest extent / res, so every axis has at most res + 1 cells (the tables are sized for (res + 1)^3)
__global__ void grid_params_kernel(const int* __restrict__ bb, int res, GridParams* gp) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float lo[3], ext[3], m = 0.f;
    for (int k = 0; k < 3; ++k) { lo[k] = o2f(bb[k]); ext[k] = fmaxf(o2f(bb[3 + k]) - lo[k], 0.f); m = fmaxf(m, ext[k]); }
    const float h = (m > 0.f) ? m / (float)res : 1.0f;
    gp->ox = lo[0]; gp->oy = lo[1]; gp->oz = lo[2]; gp->h = h; gp->inv_h = 1.0f / h;
    gp->nx = min(res + 1, (int)(ext[0] / h) + 1);
    gp->ny = min(res + 1, (int)(ext[1] / h) + 1);
    gp->nz = min(res + 1, (int)(ext[2] / h) + 1);
}

__device__ __forceinline__ void cell_coords(const GridParams& g, float x, float y, float z, int& cx, int& cy, int& cz) {
    cx = min(max((int)((x - g.ox) * g.inv_h), 0), g.nx - 1);
    cy = min(max((int)((y - g.oy) * g.inv_h), 0), g.ny - 1);
    cz = min(max((int)((z - g.oz) * g.inv_h), 0), g.nz - 1);
}

__global__ void __launch_bounds__(256)
cell_count_kernel(int64_t N, const float* __restrict__ means, const GridParams* __restrict__ gp, int32_t* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const GridParams g = *gp;
    int cx, cy, cz;
    cell_coords(g, means[3 * i], means[3 * i + 1], means[3 * i + 2], cx, cy, cz);
    atomicAdd(counts + ((int64_t)cz * g.ny + cy) * g.nx + cx, 1);
}

__global__ void __launch_bounds__(256)
cell_scatter_kernel(int64_t N, const float* __restrict__ means, const GridParams* __restrict__ gp,
                    const int32_t* __restrict__ starts, int32_t* __restrict__ fill, float4* __restrict__ sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const GridParams g = *gp;
    const float x = means[3 * i], y = means[3 * i + 1], z = means[3 * i + 2];
    int cx, cy, cz;
    cell_coords(g, x, y, z, cx, cy, cz);
    const int64_t c = ((int64_t)cz * g.ny + cy) * g.nx + cx;
    const int pos = starts[c] + atomicAdd(fill + c, 1);
    sorted[pos] = make_float4(x, y, z, __int_as_float((int)i));
}

// ---- exact KNN: one warp per query, per-lane sorted candidate lists in shared memory ----------------------
struct LaneLists { float d[KNN_K][32]; int id[KNN_K][32]; };

__device__ __forceinline__ void lane_insert(LaneLists& L, int lane, float d2, int id) {
    int k = KNN_K - 1;
    while (k > 0 && L.d[k - 1][lane] > d2) { L.d[k][lane] = L.d[k - 1][lane]; L.id[k][lane] = L.id[k - 1][lane]; --k; }
    L.d[k][lane] = d2; L.id[k][lane] = id;
}

__device__ __forceinline__ void scan_range(LaneLists& L, int lane, const float4* __restrict__ sorted, int b, int e,
                                           float qx, float qy, float qz) {
    for (int p = b + lane; p < e; p += 32) {
        const float4 v = sorted[p];
        const float dx = v.x - qx, dy = v.y - qy, dz = v.z - qz;
        const float d2 = dx * dx + dy * dy + dz * dz;
        if (d2 < L.d[KNN_K - 1][lane]) lane_insert(L, lane, d2, __float_as_int(v.w));
    }
}

// K-way merge of the 32 sorted lane lists: lane t (< 16) ends up holding the t-th nearest (distance, id);
// returns the 16th smallest distance (warp-uniform)
__device__ __forceinline__ float merge_lists(const LaneLists& L, int lane, float& out_d, int& out_id) {
    int hp = 0;
    float kth = FLT_MAX;
    for (int t = 0; t < KNN_K; ++t) {
        const float v = (hp < KNN_K) ? L.d[hp][lane] : FLT_MAX;
        const unsigned m = __reduce_min_sync(0xffffffffu, __float_as_uint(v));
        const unsigned who = __ballot_sync(0xffffffffu, __float_as_uint(v) == m);
        const int src = __ffs(who) - 1;
        const int cand = (hp < KNN_K) ? L.id[hp][lane] : -1;
        const int id = __shfl_sync(0xffffffffu, cand, src);
        if (lane == t) { out_d = __uint_as_float(m); out_id = (m == __float_as_uint(FLT_MAX)) ? -1 : id; }
        if (lane == src) ++hp;
        kth = __uint_as_float(m);
    }
    return kth;
}

__global__ void __launch_bounds__(128)
knn_kernel(int Q, const int32_t* __restrict__ query_idx, const float* __restrict__ means,
           const GridParams* __restrict__ gp, const int32_t* __restrict__ starts, const float4* __restrict__ sorted,
           int32_t* __restrict__ nbr) {
    __shared__ LaneLists lists[4];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    if (q >= Q) return;
    LaneLists& L = lists[wib];
#pragma unroll
    for (int k = 0; k < KNN_K; ++k) { L.d[k][lane] = FLT_MAX; L.id[k][lane] = -1; }
    const GridParams g = *gp;
    const int64_t qi = query_idx ? query_idx[q] : q;
    const float qx = means[3 * qi], qy = means[3 * qi + 1], qz = means[3 * qi + 2];
    int cx, cy, cz;
    cell_coords(g, qx, qy, qz, cx, cy, cz);
    // distance from the query to the nearest face of its own cell: everything outside the cube of Chebyshev
    // radius r (in cells) is farther than r * h + face
    const float fx = (qx - g.ox) - cx * g.h, fy = (qy - g.oy) - cy * g.h, fz = (qz - g.oz) - cz * g.h;
    const float face = fmaxf(0.f, fminf(fminf(fminf(fx, g.h - fx), fminf(fy, g.h - fy)), fminf(fz, g.h - fz)));
    const int rmax = max(max(max(cx, g.nx - 1 - cx), max(cy, g.ny - 1 - cy)), max(cz, g.nz - 1 - cz));
    float out_d = FLT_MAX;
    int out_id = -1;
    for (int r = 0; r <= rmax; ++r) {
        for (int dz = -r; dz <= r; ++dz) {
            const int z = cz + dz;
            if (z < 0 || z >= g.nz) continue;
            for (int dy = -r; dy <= r; ++dy) {
                const int y = cy + dy;
                if (y < 0 || y >= g.ny) continue;
                const int64_t row = ((int64_t)z * g.ny + y) * g.nx;
                if (abs(dz) == r || abs(dy) == r) {                     // a full x-run of the shell: one range
                    const int xa = max(cx - r, 0), xb = min(cx + r, g.nx - 1);
                    scan_range(L, lane, sorted, starts[row + xa], starts[row + xb + 1], qx, qy, qz);
                } else {                                                // only the two end cells belong to the shell
                    if (cx - r >= 0) scan_range(L, lane, sorted, starts[row + cx - r], starts[row + cx - r + 1], qx, qy, qz);
                    if (cx + r < g.nx) scan_range(L, lane, sorted, starts[row + cx + r], starts[row + cx + r + 1], qx, qy, qz);
                }
            }
        }
        __syncwarp();
        const float kth = merge_lists(L, lane, out_d, out_id);
        const float bound = (float)r * g.h + face;
        if (kth <= bound * bound) break;
    }
    if (lane < KNN_K) nbr[(int64_t)q * KNN_K + lane] = out_id;
}

// ---- iso-surface density loss: 16 lanes per query, one neighbour per lane -------------------------------
__global__ void __launch_bounds__(128)
iso_loss_kernel(int Q, const int32_t* __restrict__ query_idx, const int32_t* __restrict__ nbr,
                const float* __restrict__ means, const float* __restrict__ log_scales, const float* __restrict__ quats,
                const float* __restrict__ logit_opac, float w_iso, float target, float* __restrict__ g_means,
                float* __restrict__ g_ls, float* __restrict__ g_quats, float* __restrict__ g_logit,
                float* __restrict__ losses, float* __restrict__ out) {
    const int l16 = threadIdx.x & 15;
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const bool live = q < Q;
    const int64_t qi = live ? (query_idx ? query_idx[q] : q) : 0;
    const int j = live ? nbr[(int64_t)q * KNN_K + l16] : -1;
    const bool ok = j >= 0;
    float d[3] = {0.f, 0.f, 0.f}, R[9], y[3], s[3], w[3], ls[3] = {0.f, 0.f, 0.f}, qj[4] = {1.f, 0.f, 0.f, 0.f};
    float alpha = 0.f, e = 0.f;
    if (ok) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { d[k] = means[3 * qi + k] - means[3 * (int64_t)j + k]; ls[k] = log_scales[3 * (int64_t)j + k]; }
#pragma unroll
        for (int k = 0; k < 4; ++k) qj[k] = quats[4 * (int64_t)j + k];
        alpha = 1.0f / (1.0f + expf(-logit_opac[j]));
    }
    quat_to_R(qj, R, nullptr, nullptr);
    float quad = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        y[k] = R[k] * d[0] + R[3 + k] * d[1] + R[6 + k] * d[2];                 // (R^T d)_k
        s[k] = fmaxf(expf(ls[k]), 1e-5f);                                       // clamp(min=1e-5), :582
        w[k] = 1.0f / (s[k] * s[k] + 1e-8f);                                    // :626
        quad += w[k] * y[k] * y[k];
    }
    if (ok) e = expf(-0.5f * quad);
    float D = alpha * e;
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) D += __shfl_xor_sync(0xffffffffu, D, off);   // stays inside the 16 lanes
    const float diff = D - target;
    float vq_self[3] = {0.f, 0.f, 0.f};
    if (ok && g_means) {
        const float c = w_iso * 2.0f * diff / (float)Q;
        const float g = c * alpha * e * -0.5f;                                  // dL/dquad
        const float u[3] = {w[0] * y[0], w[1] * y[1], w[2] * y[2]};
        float vR[9], vq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float dd = 2.0f * g * (R[3 * a] * u[0] + R[3 * a + 1] * u[1] + R[3 * a + 2] * u[2]);   // dL/dd_a
            vq_self[a] = dd;
            atomicAdd(g_means + 3 * (int64_t)j + a, -dd);
#pragma unroll
            for (int k = 0; k < 3; ++k) vR[3 * a + k] = 2.0f * g * d[a] * u[k];
        }
        quat_to_R_vjp(qj, vR, vq);
#pragma unroll
        for (int k = 0; k < 4; ++k) atomicAdd(g_quats + 4 * (int64_t)j + k, vq[k]);
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (expf(ls[k]) >= 1e-5f)                                            // clamp passes the gradient inside
                atomicAdd(g_ls + 3 * (int64_t)j + k, g * y[k] * y[k] * (-2.0f * s[k] * w[k] * w[k]) * s[k]);
        atomicAdd(g_logit + j, c * e * alpha * (1.0f - alpha));
    }
    // the query's own centre receives the sum over its 16 neighbours
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) vq_self[a] += __shfl_xor_sync(0xffffffffu, vq_self[a], off);
    if (live && l16 == 0 && g_means)
#pragma unroll
        for (int a = 0; a < 3; ++a) atomicAdd(g_means + 3 * qi + a, vq_self[a]);
    // block reduction of (D - target)^2 and D over the 8 queries of this block
    __shared__ float s_l[8], s_d[8];
    if (l16 == 0) { s_l[threadIdx.x >> 4] = live ? diff * diff : 0.f; s_d[threadIdx.x >> 4] = live ? D : 0.f; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < 8; ++k) { a += s_l[k]; b += s_d[k]; }
        if (losses) atomicAdd(losses, w_iso * a / (float)Q);
        atomicAdd(out + 1, a / (float)Q);      // iso loss (unweighted)
        atomicAdd(out + 2, b / (float)Q);      // mean density (monitoring, :667-668)
    }
}

// flat = mean_i min_k clamp(exp(ls_ik), 1e-5)  (compute_flat_loss :42-54, :580-583)
__global__ void __launch_bounds__(256)
flat_loss_kernel(int64_t N, const float* __restrict__ log_scales, float w_flat, float* __restrict__ g_ls,
                 float* __restrict__ losses, float* __restrict__ out) {
    float acc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        float best = FLT_MAX, raw = 0.f;
        int arg = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float ex = expf(log_scales[3 * i + k]);
            const float sc = fmaxf(ex, 1e-5f);
            if (sc < best) { best = sc; arg = k; raw = ex; }                     // first minimum wins
        }
        acc += best;
        if (g_ls && raw >= 1e-5f) g_ls[3 * i + arg] += w_flat * raw / (float)N;  // one thread per Gaussian: no atomics
    }
    __shared__ float red[8];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f;
        for (int k = 0; k < 8; ++k) a += red[k];
        if (losses) atomicAdd(losses, w_flat * a / (float)N);
        atomicAdd(out, a / (float)N);
    }
}

// query centres: hash(seed, s) % N (with replacement; the reference uses torch.randperm(N)[:Q], :648)
__global__ void __launch_bounds__(256) sample_queries_kernel(int Q, int64_t N, uint64_t seed, int32_t* __restrict__ idx) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= Q) return;
    uint64_t x = seed * 0x9E3779B97F4A7C15ull + (uint64_t)s * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull;
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    idx[s] = (int32_t)(x % (uint64_t)N);
}

} // namespace isogs

using namespace isogs;

static inline int64_t grid_cells(int res) { return (int64_t)(res + 1) * (res + 1) * (res + 1); }

extern "C" int isogs_knn_workspace(int64_t N, int res, size_t* grid_bytes, size_t* sorted_bytes) {
    if (N < 0 || res < 1 || res > 512) { set_error("isogs_knn_workspace: bad argument"); return ISOGS_E_ARG; }
    // [params 8 words | bbox 8 words | starts cells+1 | fill cells]
    if (grid_bytes) *grid_bytes = sizeof(int32_t) * (size_t)(16 + 2 * grid_cells(res) + 1);
    if (sorted_bytes) *sorted_bytes = sizeof(float) * 4 * (size_t)N;
    return ISOGS_OK;
}

extern "C" int isogs_knn16(int64_t N, const float* means, int res, int Q, const int32_t* query_idx, uint64_t seed,
                           int32_t* sampled_idx, int32_t* grid, float* sorted, void* scan_temp, size_t scan_temp_bytes,
                           int32_t* nbr, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N <= 0 || N >= (1ll << 31) || !means || res < 1 || res > 512 || Q <= 0 || !grid || !sorted || !nbr ||
        ((uintptr_t)sorted & 15) || (!query_idx && Q != N && !sampled_idx)) {
        set_error("isogs_knn16: bad argument");
        return ISOGS_E_ARG;
    }
    const int64_t cells = grid_cells(res);
    GridParams* gp = reinterpret_cast<GridParams*>(grid);
    int* bb = grid + 8;
    int32_t* starts = grid + 16;
    int32_t* fill = starts + cells + 1;
    ISOGS_CUDA(cudaMemsetAsync(starts, 0, sizeof(int32_t) * (size_t)(2 * cells + 1), stream));
    bbox_init_kernel<<<1, 32, 0, stream>>>(bb);
    const unsigned gN = (unsigned)((N + 255) / 256);
    bbox_kernel<<<min(gN, 148u * 8u), 256, 0, stream>>>(N, means, bb);
    grid_params_kernel<<<1, 32, 0, stream>>>(bb, res, gp);
    cell_count_kernel<<<gN, 256, 0, stream>>>(N, means, gp, starts);
    size_t need = 0;
    ISOGS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, starts, starts, (int)(cells + 1), stream));
    if (need > scan_temp_bytes) { set_error("isogs_knn16: scan temp too small (%zu needed)", need); return ISOGS_E_ARG; }
    ISOGS_CUDA(cub::DeviceScan::ExclusiveSum(scan_temp, scan_temp_bytes, starts, starts, (int)(cells + 1), stream));
    cell_scatter_kernel<<<gN, 256, 0, stream>>>(N, means, gp, starts, fill, reinterpret_cast<float4*>(sorted));
    const int32_t* qidx = query_idx;
    if (!qidx && sampled_idx) {
        sample_queries_kernel<<<(unsigned)((Q + 255) / 256), 256, 0, stream>>>(Q, N, seed, sampled_idx);
        qidx = sampled_idx;
    }
    knn_kernel<<<(unsigned)((Q + 3) / 4), 128, 0, stream>>>(Q, qidx, means, gp, starts,
                                                           reinterpret_cast<const float4*>(sorted), nbr);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_iso_flat_loss(int64_t N, const float* means, const float* log_scales, const float* quats,
                                   const float* logit_opac, int Q, const int32_t* query_idx, const int32_t* nbr,
                                   float w_iso, float w_flat, float target, float* g_means, float* g_log_scales,
                                   float* g_quats, float* g_logit_opac, float* losses, float* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool any_g = g_means || g_log_scales || g_quats || g_logit_opac;
    if (N <= 0 || !means || !log_scales || !quats || !logit_opac || Q < 0 || (Q > 0 && !nbr) || !out ||
        (any_g && !(g_means && g_log_scales && g_quats && g_logit_opac))) {
        set_error("isogs_iso_flat_loss: bad argument (three-axis scales and all four gradients, or none)");
        return ISOGS_E_ARG;
    }
    if (Q > 0)
        iso_loss_kernel<<<(unsigned)((Q + 7) / 8), 128, 0, stream>>>(Q, query_idx, nbr, means, log_scales, quats, logit_opac,
                                                                    w_iso, target, g_means, g_log_scales, g_quats,
                                                                    g_logit_opac, losses, out);
    const unsigned gN = (unsigned)std::min<int64_t>((N + 255) / 256, 148 * 8);
    flat_loss_kernel<<<gN, 256, 0, stream>>>(N, log_scales, w_flat, g_log_scales, losses, out);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
