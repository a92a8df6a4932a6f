// peer.cu -- SURVEY 8(e): the one exchange step of keyframe-sharded mapping, the average of the flat
// gradient buffer over the ranks, written directly over NVLink / NVSwitch peer memory instead of through
// NCCL.  (The reference has no distributed code; its sequential analogue is one keyframe per
// iteration, python_ordinary/scripts/splatam.py:1415-1426.)
//
// Every rank allocates its gradient buffer + a small flag block with cudaMalloc, exports it with CUDA IPC and
// maps the buffers of all peers.  Per iteration one kernel per rank does a two-shot all-reduce:
//   1. tell every peer "my gradients are complete" and wait for the same message from each of them;
//   2. for the 1/world slice this rank owns: load the slice from every rank (fixed rank order, so the sum is
//      deterministic and every replica receives the same bits), scale by 1/world, store the result into the
//      buffer of every rank;
//   3. the last CTA tells every peer "my slice has landed in your buffer".
// isogs_peer_wait (one warp, in stream order before Adam) then waits for message 3 from every rank.
// NVLink bytes per GPU and direction: 2 * (world-1)/world * numel * 4 -- the same as a ring all-reduce, but with
// one launch, no staging copies and no protocol overhead.  Every spin is bounded (2 s of globaltimer): on a
// timeout status[0] is set and the kernel carries on, so a lost peer cannot hang the GPU.
#include <cstring>

#include "common.cuh"

namespace isogs {

constexpr int PEER_MAX_WORLD = 16;
// flag block layout (uint32): [0, 16) "gradients ready" epoch per source rank, [16, 32) "slice landed" epoch
// per source rank, [32] CTA arrival counter of the local kernel; the caller provides 64 words

struct PeerPtrs {
    float4* buf[PEER_MAX_WORLD];
    uint32_t* flags[PEER_MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_f4(float4* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// spin until *flag has reached `epoch` (wrap-safe); false on timeout
__device__ __forceinline__ bool wait_epoch(const uint32_t* flag, uint32_t epoch) {
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
        if (global_ns() - t0 > 2000000000ull) return false;
        __nanosleep(64);
    }
    return true;
}

template <int WORLD>
__device__ __forceinline__ void reduce_slice(const PeerPtrs& pp, int world, int64_t begin, int64_t end, float inv_world) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
        float4 v[WORLD > 0 ? WORLD : 1];
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (WORLD > 0) {
#pragma unroll
            for (int p = 0; p < WORLD; ++p) v[p] = ld_relaxed_sys_f4(pp.buf[p] + i);     // all loads in flight first
#pragma unroll
            for (int p = 0; p < WORLD; ++p) { acc.x += v[p].x; acc.y += v[p].y; acc.z += v[p].z; acc.w += v[p].w; }
        } else {
            for (int p = 0; p < world; ++p) {
                const float4 t = ld_relaxed_sys_f4(pp.buf[p] + i);
                acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
            }
        }
        acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
        const int n = WORLD > 0 ? WORLD : world;
#pragma unroll
        for (int p = 0; p < n; ++p) st_relaxed_sys_f4(pp.buf[p] + i, acc);
    }
}

__global__ void __launch_bounds__(256)
peer_allreduce_kernel(PeerPtrs pp, int world, int rank, int64_t n4, uint32_t epoch, int32_t* __restrict__ status) {
    const int tid = threadIdx.x;
    uint32_t* mine = pp.flags[rank];
    // 1. "my gradients are complete" (stream order put this kernel after the backward pass)
    if (blockIdx.x == 0 && tid < world) st_release_sys(pp.flags[tid] + rank, epoch);
    if (tid < world && !wait_epoch(mine + tid, epoch)) atomicExch(status, 1);
    __syncthreads();
    // 2. reduce the slice this rank owns and broadcast it
    const int64_t per = (n4 + world - 1) / world;
    const int64_t begin = min(n4, per * rank), end = min(n4, begin + per);
    const float inv_world = 1.0f / (float)world;
    switch (world) {
        case 1: reduce_slice<1>(pp, world, begin, end, inv_world); break;
        case 2: reduce_slice<2>(pp, world, begin, end, inv_world); break;
        case 4: reduce_slice<4>(pp, world, begin, end, inv_world); break;
        case 8: reduce_slice<8>(pp, world, begin, end, inv_world); break;
        default: reduce_slice<0>(pp, world, begin, end, inv_world); break;
    }
    // 3. last CTA: "my slice has landed"
    __threadfence_system();
    __syncthreads();
    __shared__ bool s_last;
    if (tid == 0) {
        const uint32_t prev = atomicAdd(mine + 32, 1u);
        s_last = prev == gridDim.x - 1;
        if (s_last) mine[32] = 0;
    }
    __syncthreads();
    if (s_last && tid < world) {
        __threadfence_system();
        st_release_sys(pp.flags[tid] + 16 + rank, epoch);
    }
}

// ---- fused exchange + optimizer step ------------------------------------------------------------------------------
// The slice owner does not hand the averaged GRADIENTS back: it applies Adam to its 1/world slice in registers (its
// slice of the moments is the only copy that is kept up to date -- optimizer state sharded like ZeRO-1) and stores the
// updated PARAMETERS into every rank's parameter tensors.  Same NVLink bytes as the all-reduce (slice loads from every
// peer + slice stores to every peer), but Adam shrinks from all N Gaussians to N / world per rank, one launch and the
// 32 MB re-read of averaged gradients disappear, and the replicas stay bit-identical by construction (one rank computes
// each parameter, everybody receives the same bits).
constexpr int PEER_MAX_GROUPS = 5;
struct PeerAdamArgs {
    int64_t off4[PEER_MAX_GROUPS], n4[PEER_MAX_GROUPS]; // group position in the flat gradient buffer (float4 units)
    float4* param[PEER_MAX_GROUPS][PEER_MAX_WORLD];     // parameter tensor of the group on every rank
    float4* m[PEER_MAX_GROUPS];
    float4* v[PEER_MAX_GROUPS];                         // LOCAL moments
    float lr[PEER_MAX_GROUPS];
    int n_groups;
    float b2, one_minus_b1, one_minus_b2, eps;
    const isogs_adam_state* state;
    // NVLS (multicast) addresses of the same allocations, or null: the gradient slice is then reduced INSIDE the
    // NVSwitch by one multimem.ld_reduce and the parameters are broadcast by one multimem.st
    const float4* mc_grad;
    float4* mc_param[PEER_MAX_GROUPS];
};

__device__ __forceinline__ float4 multimem_ld_reduce_add_f4(const float4* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_f4(float4* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// the arithmetic of adam_kernel (step.cu), operation for operation
__device__ __forceinline__ void peer_adam_one(float& p, float g, float& m, float& v, const PeerAdamArgs& a, float step_size,
                                              float inv_bc2_sqrt) {
    adam_update(p, g, m, v, a.one_minus_b1, a.b2, a.one_minus_b2, inv_bc2_sqrt, a.eps, step_size);
}

template <int WORLD, bool MC = false>
__device__ __forceinline__ void reduce_adam_slice(const PeerPtrs& pp, const PeerAdamArgs& a, int world, int rank,
                                                  int64_t begin, int64_t end, float inv_world) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float ib2 = (float)a.state->inv_bc2_sqrt;
    const double ibc1 = a.state->inv_bc1;
    for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (MC) {
            acc = multimem_ld_reduce_add_f4(a.mc_grad + i);
        } else if (WORLD > 0) {
            float4 t[WORLD > 0 ? WORLD : 1];
#pragma unroll
            for (int p = 0; p < WORLD; ++p) t[p] = ld_relaxed_sys_f4(pp.buf[p] + i);     // all loads in flight first
#pragma unroll
            for (int p = 0; p < WORLD; ++p) { acc.x += t[p].x; acc.y += t[p].y; acc.z += t[p].z; acc.w += t[p].w; }
        } else {
            for (int p = 0; p < world; ++p) {
                const float4 t = ld_relaxed_sys_f4(pp.buf[p] + i);
                acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
            }
        }
        acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
        int g = -1;
#pragma unroll
        for (int k = 0; k < PEER_MAX_GROUPS; ++k)
            if (k < a.n_groups && i >= a.off4[k] && i < a.off4[k] + a.n4[k]) g = k;
        if (g < 0) continue;                           // padding between groups
        const int64_t j = i - a.off4[g];
        const float ss = (float)((double)a.lr[g] * ibc1);
        float4 P4 = a.param[g][rank][j], M4 = a.m[g][j], V4 = a.v[g][j];
        peer_adam_one(P4.x, acc.x, M4.x, V4.x, a, ss, ib2);
        peer_adam_one(P4.y, acc.y, M4.y, V4.y, a, ss, ib2);
        peer_adam_one(P4.z, acc.z, M4.z, V4.z, a, ss, ib2);
        peer_adam_one(P4.w, acc.w, M4.w, V4.w, a, ss, ib2);
        a.m[g][j] = M4;
        a.v[g][j] = V4;
        if (MC) {
            multimem_st_f4(a.mc_param[g] + j, P4);
        } else {
            const int n = WORLD > 0 ? WORLD : world;
#pragma unroll
            for (int p = 0; p < n; ++p) st_relaxed_sys_f4(a.param[g][p] + j, P4);
        }
    }
}

__global__ void __launch_bounds__(256)
peer_reduce_adam_kernel(PeerPtrs pp, PeerAdamArgs a, int world, int rank, int64_t n4, uint32_t epoch,
                        int32_t* __restrict__ status) {
    const int tid = threadIdx.x;
    uint32_t* mine = pp.flags[rank];
    // 1. "my gradients are complete AND I have finished reading my parameters for this iteration" (stream order put
    //    this kernel after the backward pass), so peers may overwrite them
    if (blockIdx.x == 0 && tid < world) st_release_sys(pp.flags[tid] + rank, epoch);
    if (tid < world && !wait_epoch(mine + tid, epoch)) atomicExch(status, 1);
    __syncthreads();
    const int64_t per = (n4 + world - 1) / world;
    const int64_t begin = min(n4, per * rank), end = min(n4, begin + per);
    const float inv_world = 1.0f / (float)world;
    if (a.mc_grad != nullptr) {
        reduce_adam_slice<0, true>(pp, a, world, rank, begin, end, inv_world);
    } else {
        switch (world) {
            case 1: reduce_adam_slice<1>(pp, a, world, rank, begin, end, inv_world); break;
            case 2: reduce_adam_slice<2>(pp, a, world, rank, begin, end, inv_world); break;
            case 4: reduce_adam_slice<4>(pp, a, world, rank, begin, end, inv_world); break;
            case 8: reduce_adam_slice<8>(pp, a, world, rank, begin, end, inv_world); break;
            default: reduce_adam_slice<0>(pp, a, world, rank, begin, end, inv_world); break;
        }
    }
    // 3. last CTA: "my slice of the parameters has landed"
    __threadfence_system();
    __syncthreads();
    __shared__ bool s_last;
    if (tid == 0) {
        const uint32_t prev = atomicAdd(mine + 32, 1u);
        s_last = prev == gridDim.x - 1;
        if (s_last) mine[32] = 0;
    }
    __syncthreads();
    if (s_last && tid < world) {
        __threadfence_system();
        st_release_sys(pp.flags[tid] + 16 + rank, epoch);
    }
}

__global__ void peer_wait_kernel(const uint32_t* __restrict__ mine, int world, uint32_t epoch, int32_t* __restrict__ status) {
    if (threadIdx.x < world && !wait_epoch(mine + 16 + threadIdx.x, epoch)) atomicExch(status, 2);
}

} // namespace isogs

using namespace isogs;

extern "C" int isogs_peer_alloc(size_t bytes, void** ptr, unsigned char* handle) {
    if (!ptr || !handle || bytes == 0) { set_error("isogs_peer_alloc: bad argument"); return ISOGS_E_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void* p = nullptr;
    ISOGS_CUDA(cudaMalloc(&p, bytes));
    ISOGS_CUDA(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    const int rc = check_cuda(cudaIpcGetMemHandle(&h, p), "cudaIpcGetMemHandle");
    if (rc != 0) { cudaFree(p); return rc; }
    memcpy(handle, &h, sizeof(h));
    *ptr = p;
    return ISOGS_OK;
}

extern "C" int isogs_peer_open(const unsigned char* handle, void** ptr) {
    if (!ptr || !handle) { set_error("isogs_peer_open: bad argument"); return ISOGS_E_ARG; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    ISOGS_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return ISOGS_OK;
}

extern "C" int isogs_peer_close(void* ptr) {
    if (ptr) ISOGS_CUDA(cudaIpcCloseMemHandle(ptr));
    return ISOGS_OK;
}

extern "C" int isogs_peer_free(void* ptr) {
    if (ptr) ISOGS_CUDA(cudaFree(ptr));
    return ISOGS_OK;
}

extern "C" int isogs_allreduce_avg_p2p(int world, int rank, float* const* bufs, uint32_t* const* flags, int64_t numel,
                                       uint32_t epoch, int n_ctas, int32_t* status, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || !bufs || !flags || numel < 0 || (numel & 3) ||
        !status) {
        set_error("isogs_allreduce_avg_p2p: bad argument (world <= 16, numel a multiple of 4)");
        return ISOGS_E_ARG;
    }
    PeerPtrs pp{};
    for (int p = 0; p < world; ++p) {
        if (!bufs[p] || !flags[p] || ((uintptr_t)bufs[p] & 15)) { set_error("isogs_allreduce_avg_p2p: null or misaligned peer buffer"); return ISOGS_E_ARG; }
        pp.buf[p] = (float4*)bufs[p];
        pp.flags[p] = flags[p];
    }
    const int64_t n4 = numel / 4, per = (n4 + world - 1) / world;
    int grid = n_ctas > 0 ? n_ctas : 148 * 4;
    const int64_t want = (per + 255) / 256;
    if (want < grid) grid = (int)(want > 0 ? want : 1);
    peer_allreduce_kernel<<<grid, 256, 0, stream>>>(pp, world, rank, n4, epoch, status);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

static int reduce_adam_launch(int world, int rank, float* const* grad_bufs, uint32_t* const* flags, int64_t numel,
                              const isogs_peer_adam_group* groups, int n_groups, float beta1, float beta2, float eps,
                              const isogs_adam_state* state, uint32_t epoch, int n_ctas, int32_t* status,
                              const float* mc_grad, float* const* mc_params, cudaStream_t stream) {
    if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || !grad_bufs || !flags || numel < 0 || (numel & 3) ||
        !status || !groups || n_groups < 1 || n_groups > PEER_MAX_GROUPS || !state) {
        set_error("isogs_reduce_adam_p2p: bad argument (world <= 16, numel a multiple of 4, <= 5 groups, device Adam state)");
        return ISOGS_E_ARG;
    }
    PeerPtrs pp{};
    for (int p = 0; p < world; ++p) {
        if (!grad_bufs[p] || !flags[p] || ((uintptr_t)grad_bufs[p] & 15)) { set_error("isogs_reduce_adam_p2p: null or misaligned peer buffer"); return ISOGS_E_ARG; }
        pp.buf[p] = (float4*)grad_bufs[p];
        pp.flags[p] = flags[p];
    }
    PeerAdamArgs a{};
    a.mc_grad = (const float4*)mc_grad;
    if (mc_grad != nullptr) {
        if (((uintptr_t)mc_grad & 15) || !mc_params) { set_error("isogs_reduce_adam_nvls: null or misaligned multicast pointer"); return ISOGS_E_ARG; }
        for (int g = 0; g < n_groups; ++g) {
            if (!mc_params[g] || ((uintptr_t)mc_params[g] & 15)) { set_error("isogs_reduce_adam_nvls: group %d: null or misaligned multicast parameter pointer", g); return ISOGS_E_ARG; }
            a.mc_param[g] = (float4*)mc_params[g];
        }
    }
    a.n_groups = n_groups;
    a.b2 = beta2; a.one_minus_b1 = 1.0f - beta1; a.one_minus_b2 = 1.0f - beta2; a.eps = eps;
    a.state = state;
    for (int g = 0; g < n_groups; ++g) {
        const isogs_peer_adam_group& G = groups[g];
        if ((G.offset & 3) || G.numel < 0 || !G.exp_avg || !G.exp_avg_sq || (((uintptr_t)G.exp_avg | (uintptr_t)G.exp_avg_sq) & 15)) {
            set_error("isogs_reduce_adam_p2p: group %d: offset must be a multiple of 4 floats, moments 16-byte aligned", g);
            return ISOGS_E_ARG;
        }
        a.off4[g] = G.offset / 4;
        a.n4[g] = (G.numel + 3) / 4;                  // the tail float4 may carry up to 3 padding floats (kept zero / unused)
        a.m[g] = (float4*)G.exp_avg;
        a.v[g] = (float4*)G.exp_avg_sq;
        a.lr[g] = G.lr;
        for (int p = 0; p < world; ++p) {
            if (!G.param[p] || ((uintptr_t)G.param[p] & 15)) { set_error("isogs_reduce_adam_p2p: group %d: null or misaligned parameter pointer of rank %d", g, p); return ISOGS_E_ARG; }
            a.param[g][p] = (float4*)G.param[p];
        }
    }
    const int64_t n4 = numel / 4, per = (n4 + world - 1) / world;
    int grid = n_ctas > 0 ? n_ctas : 148 * 4;
    const int64_t want = (per + 255) / 256;
    if (want < grid) grid = (int)(want > 0 ? want : 1);
    peer_reduce_adam_kernel<<<grid, 256, 0, stream>>>(pp, a, world, rank, n4, epoch, status);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}

extern "C" int isogs_reduce_adam_p2p(int world, int rank, float* const* grad_bufs, uint32_t* const* flags, int64_t numel,
                                     const isogs_peer_adam_group* groups, int n_groups, float beta1, float beta2, float eps,
                                     const isogs_adam_state* state, uint32_t epoch, int n_ctas, int32_t* status,
                                     void* stream_) {
    return reduce_adam_launch(world, rank, grad_bufs, flags, numel, groups, n_groups, beta1, beta2, eps, state, epoch,
                              n_ctas, status, nullptr, nullptr, (cudaStream_t)stream_);
}

extern "C" int isogs_reduce_adam_nvls(int world, int rank, float* const* grad_bufs, uint32_t* const* flags, int64_t numel,
                                      const isogs_peer_adam_group* groups, int n_groups, float beta1, float beta2, float eps,
                                      const isogs_adam_state* state, uint32_t epoch, int n_ctas, int32_t* status,
                                      const float* mc_grad, float* const* mc_params, void* stream_) {
    if (!mc_grad || !mc_params) { set_error("isogs_reduce_adam_nvls: multicast pointers required"); return ISOGS_E_ARG; }
    return reduce_adam_launch(world, rank, grad_bufs, flags, numel, groups, n_groups, beta1, beta2, eps, state, epoch,
                              n_ctas, status, mc_grad, mc_params, (cudaStream_t)stream_);
}

extern "C" int isogs_peer_wait(int world, const uint32_t* flags_local, uint32_t epoch, int32_t* status, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (world < 1 || world > PEER_MAX_WORLD || !flags_local || !status) { set_error("isogs_peer_wait: bad argument"); return ISOGS_E_ARG; }
    peer_wait_kernel<<<1, 32, 0, stream>>>(flags_local, world, epoch, status);
    ISOGS_CUDA(cudaGetLastError());
    return ISOGS_OK;
}
