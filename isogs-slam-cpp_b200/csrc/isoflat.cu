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
    // one set of 6 atomics per CTA (the 6 addresses are shared by the whole grid)
    __shared__ float s_lo[3][8], s_hi[3][8];
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int k = 0; k < 3; ++k) { s_lo[k][threadIdx.x >> 5] = lo[k]; s_hi[k][threadIdx.x >> 5] = hi[k]; }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int k = threadIdx.x;
        float a = s_lo[k][0], b = s_hi[k][0];
        for (int w = 1; w < 8; ++w) { a = fminf(a, s_lo[k][w]); b = fmaxf(b, s_hi[k][w]); }
        atomicMin(bb + k, f2o(a));
        atomicMax(bb + 3 + k, f2o(b));
    }
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

// ---- exact KNN: one warp per query, ONE shared candidate buffer per warp ---------------------------------
// buf[0..15] = the best 16 so far (ascending after a compaction), buf[16..] = pending candidates that beat the
// current 16th distance tau.  A round of 32 candidates costs one compare per lane plus a ballot; only when more
// than 16 are pending does the warp re-select the best 16 (16 rounds of a warp-wide minimum over two registers per
// lane).  tau drops fast, so later candidates are rejected by the single compare: ~60 accepted out of ~600.
struct WarpBuf { float d[64]; int id[64]; };

__device__ __forceinline__ float compact_top16(WarpBuf& B, int lane, int n_total) {
    __syncwarp();
    float a = (lane < n_total) ? B.d[lane] : FLT_MAX;
    int ia = (lane < n_total) ? B.id[lane] : -1;
    float b = (lane + 32 < n_total) ? B.d[lane + 32] : FLT_MAX;
    int ib = (lane + 32 < n_total) ? B.id[lane + 32] : -1;
    __syncwarp();
    float kth = FLT_MAX;
#pragma unroll 4
    for (int t = 0; t < KNN_K; ++t) {
        const bool useb = b < a;
        const float v = useb ? b : a;
        const unsigned m = __reduce_min_sync(0xffffffffu, __float_as_uint(v));
        const unsigned who = __ballot_sync(0xffffffffu, __float_as_uint(v) == m);
        if (lane == __ffs(who) - 1) {
            B.d[t] = v;
            B.id[t] = (m == __float_as_uint(FLT_MAX)) ? -1 : (useb ? ib : ia);
            if (useb) b = FLT_MAX; else a = FLT_MAX;
        }
        kth = __uint_as_float(m);
    }
    __syncwarp();
    return kth;
}

// Scan a batch of up to 32 cell ranges (lane l holds range l: [rb, rb + rl)) as ONE virtual list: the bounds were
// loaded by different lanes in parallel (one memory latency for the whole batch), an inclusive warp scan of the
// lengths maps a virtual index to (range, offset), and every lane strides over the concatenation, so the point
// loads of a shell are independent of each other.
__device__ __forceinline__ void scan_batch(WarpBuf& B, int lane, const float4* __restrict__ sorted, int rb, int rl,
                                           float qx, float qy, float qz, float& tau, int& n_pending) {
    int incl = rl;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    for (int v = lane; v < ((total + 31) & ~31); v += 32) {
        // range of virtual index v = number of lanes whose inclusive prefix is <= v (prefixes are monotone:
        // 5-step binary search through shuffles)
        int src = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
            if (__shfl_sync(0xffffffffu, incl, src + step - 1) <= v) src += step;
        const int r_incl = __shfl_sync(0xffffffffu, incl, src);
        const int r_len = __shfl_sync(0xffffffffu, rl, src);
        const int r_beg = __shfl_sync(0xffffffffu, rb, src);
        float d2 = FLT_MAX;
        int id = -1;
        if (v < total) {
            const float4 p = sorted[r_beg + (v - (r_incl - r_len))];
            const float dx = p.x - qx, dy = p.y - qy, dz = p.z - qz;
            d2 = dx * dx + dy * dy + dz * dz;
            id = __float_as_int(p.w);
        }
        const bool acc = d2 < tau;
        const unsigned mask = __ballot_sync(0xffffffffu, acc);
        if (mask) {
            if (acc) {
                const int pos = KNN_K + n_pending + __popc(mask & ((1u << lane) - 1u));
                B.d[pos] = d2; B.id[pos] = id;
            }
            n_pending += __popc(mask);
            if (n_pending > KNN_K) { tau = compact_top16(B, lane, KNN_K + n_pending); n_pending = 0; }
        }
    }
}

__global__ void __launch_bounds__(128)
knn_kernel(int Q, const int32_t* __restrict__ query_idx, const float* __restrict__ means,
           const GridParams* __restrict__ gp, const int32_t* __restrict__ starts, const float4* __restrict__ sorted,
           int32_t* __restrict__ nbr) {
    __shared__ WarpBuf bufs[4];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + wib;
    if (q >= Q) return;
    WarpBuf& B = bufs[wib];
    if (lane < KNN_K) { B.d[lane] = FLT_MAX; B.id[lane] = -1; }
    __syncwarp();
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
    float tau = FLT_MAX;
    int n_pending = 0;
    for (int r = 0; r <= rmax; ++r) {
        // enumerate the shell's ranges: item t -> (dz, dy, which): a row with |dz| == r or |dy| == r is one full
        // x-run, an inner row contributes its two end cells; lanes take items t = lane, lane + 32, ...
        const int side = 2 * r + 1;
        const int n_items = (r == 0) ? 1 : 2 * side * side;          // two slots per row; full rows use slot 0 only
        for (int base = 0; base < n_items; base += 32) {
            const int t = base + lane;
            int rb = 0, rl = 0;
            if (t < n_items) {
                const int row_id = t >> 1, which = t & 1;
                const int dz = row_id / side - r, dy = row_id % side - r;
                const int z = cz + dz, y = cy + dy;
                if (z >= 0 && z < g.nz && y >= 0 && y < g.ny) {
                    const int64_t row = ((int64_t)z * g.ny + y) * g.nx;
                    const bool full = (abs(dz) == r) || (abs(dy) == r);
                    int xa = -1, xb = -2;
                    if (r == 0) { if (which == 0) { xa = cx; xb = cx; } }
                    else if (full) { if (which == 0) { xa = max(cx - r, 0); xb = min(cx + r, g.nx - 1); } }
                    else if (which == 0) { if (cx - r >= 0) { xa = cx - r; xb = cx - r; } }
                    else { if (cx + r < g.nx) { xa = cx + r; xb = cx + r; } }
                    if (xb >= xa) { rb = starts[row + xa]; rl = starts[row + xb + 1] - rb; }
                }
            }
            scan_batch(B, lane, sorted, rb, rl, qx, qy, qz, tau, n_pending);
        }
        if (n_pending > 0) { tau = compact_top16(B, lane, KNN_K + n_pending); n_pending = 0; }
        const float bound = (float)r * g.h + face;
        if (tau <= bound * bound) break;
    }
    __syncwarp();
    if (lane < KNN_K) nbr[(int64_t)q * KNN_K + lane] = B.id[lane];
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
    bbox_kernel<<<min(gN, 148u * 4u), 256, 0, stream>>>(N, means, bb);
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
