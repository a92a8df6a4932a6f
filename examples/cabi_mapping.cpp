// cabi_mapping.cpp -- a torch-free C++ host on top of the C ABI (include/isogs_rast.h): what
// cpp/src/rendering/renderer.cpp + cpp/src/slam/mapper.cpp of the reference would call once they link
// libisogs_rast.so instead of gsplat's ATen ops.  Only cudart + the C ABI.
//
// It back-projects a synthetic RGB-D frame into isotropic Gaussians (Mapper::addNewGaussians through
// isogs_backproject_append on an empty map), perturbs their colours, captures ONE mapping iteration
//     isogs_map_step = preprocess -> tile_bin_async -> blend_fwd -> loss_l1 + SSIM -> blend_bwd -> preprocess_bwd -> adam
// in a CUDA graph (isogs_graph_begin / _end) and replays it, printing the loss, which must fall.  Exit code 0 on success.
//
// Build:  g++ -O2 -std=c++17 -I include -I /usr/local/cuda/include examples/cabi_mapping.cpp \
//             -L isogs-slam-cpp_b200 -lisogs_rast -L /usr/local/cuda/lib64 -lcudart -Wl,-rpath,'$ORIGIN/../isogs-slam-cpp_b200' \
//             -o examples/cabi_mapping
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "isogs_rast.h"

#define CK(call)                                                                          \
    do {                                                                                  \
        int rc_ = (call);                                                                 \
        if (rc_ != 0) { std::fprintf(stderr, "%s -> %d: %s\n", #call, rc_, isogs_last_error()); return 2; } \
    } while (0)
#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 3; } \
    } while (0)

template <class T> static T* dalloc(size_t n) {
    void* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) { std::fprintf(stderr, "cudaMalloc failed\n"); std::exit(4); }
    cudaMemset(p, 0, n * sizeof(T));
    return static_cast<T*>(p);
}

int main(int argc, char** argv) {
    const int W = 320, H = 240, iters = argc > 1 ? std::atoi(argv[1]) : 30;
    const int64_t P = (int64_t)W * H;
    isogs_camera cam{};
    cam.width = W; cam.height = H; cam.fx = cam.fy = 260.f; cam.cx = 159.5f; cam.cy = 119.5f;
    cam.R[0] = cam.R[4] = cam.R[8] = 1.f;
    cam.near_plane = 0.01f; cam.far_plane = 100.f; cam.eps2d = 0.3f;
    isogs_pose pose{{1.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};

    // synthetic frame: a wavy wall with a colour pattern
    std::vector<float> rgb(3 * P), depth(P);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int64_t i = (int64_t)y * W + x;
            depth[i] = 2.0f + 0.3f * std::sin(x / 40.f) + 0.2f * std::cos(y / 25.f);
            rgb[i] = 0.5f + 0.4f * std::sin(x / 9.f);
            rgb[P + i] = 0.5f + 0.4f * std::cos(y / 7.f);
            rgb[2 * P + i] = 0.5f + 0.4f * std::sin((x + y) / 13.f);
        }
    cudaStream_t st;
    CU(cudaStreamCreate(&st));
    float *d_rgb = dalloc<float>(3 * P), *d_depth = dalloc<float>(P);
    CU(cudaMemcpy(d_rgb, rgb.data(), 3 * P * sizeof(float), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_depth, depth.data(), P * sizeof(float), cudaMemcpyHostToDevice));

    // pool (GaussianModel, cpp/include/core/gaussian_model.hpp:18-124): parameters, moments, gradients
    const int64_t cap = P;
    float *means = dalloc<float>(3 * cap), *cols = dalloc<float>(3 * cap), *quats = dalloc<float>(4 * cap),
          *logit = dalloc<float>(cap), *ls = dalloc<float>(cap);
    float* par[4] = {means, cols, logit, ls};
    const int width[4] = {3, 3, 1, 1};
    const float lr[4] = {1e-4f, 2.5e-3f, 0.05f, 1e-3f};     // configs/replica/splatam.py:109-117
    float *mom1[4], *mom2[4], *grad[4];
    for (int g = 0; g < 4; ++g) { mom1[g] = dalloc<float>(width[g] * cap); mom2[g] = dalloc<float>(width[g] * cap); grad[g] = dalloc<float>(width[g] * cap); }

    // densify an EMPTY map from the frame: silhouette 0 everywhere -> every valid pixel becomes a Gaussian
    size_t scan_b = 0, sort_b = 0;
    CK(isogs_temp_bytes(P, 4 * P, &scan_b, &sort_b));
    void *scan_tmp = dalloc<unsigned char>(scan_b), *sort_tmp = dalloc<unsigned char>(sort_b);
    int32_t *scan = dalloc<int32_t>(2 * P), *count = dalloc<int32_t>(1);
    float *zero_depth = dalloc<float>(P), *zero_alpha = dalloc<float>(P);
    CK(isogs_backproject_append(&cam, &pose, d_rgb, d_depth, zero_depth, zero_alpha, 0.5f, 1e30f, means, cols, quats,
                                logit, ls, 1, 0.f, 0, cap, scan, scan_tmp, scan_b, count, st));
    int32_t n_host = 0;
    CU(cudaMemcpyAsync(&n_host, count, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int64_t N = n_host;
    std::printf("map initialised from the frame: %lld Gaussians\n", (long long)N);
    if (N != P) { std::fprintf(stderr, "expected one Gaussian per pixel\n"); return 5; }
    // spoil the colours so that there is something to optimise
    std::vector<float> c0(3 * N);
    CU(cudaMemcpy(c0.data(), cols, 3 * N * sizeof(float), cudaMemcpyDeviceToHost));
    for (auto& v : c0) v = 0.5f * v + 0.2f;
    CU(cudaMemcpy(cols, c0.data(), 3 * N * sizeof(float), cudaMemcpyHostToDevice));

    // per-step scratch
    const int64_t isect_cap = 6 * N;
    const int n_tiles = ((W + ISOGS_TILE - 1) / ISOGS_TILE) * ((H + ISOGS_TILE - 1) / ISOGS_TILE);
    CK(isogs_temp_bytes(N, isect_cap, &scan_b, &sort_b));
    cudaFree(sort_tmp);
    sort_tmp = dalloc<unsigned char>(sort_b);
    float *rec = dalloc<float>(ISOGS_REC * N), *v_rec = dalloc<float>(ISOGS_REC * N);
    int32_t *radii = dalloc<int32_t>(2 * N), *vals = dalloc<int32_t>(isect_cap), *ranges = dalloc<int32_t>(n_tiles + 1),
            *totals = dalloc<int32_t>(2), *last = dalloc<int32_t>(P);
    uint64_t* bucket = dalloc<uint64_t>(isect_cap);
    float *render = dalloc<float>(5 * P), *v_render = dalloc<float>(5 * P), *alpha = dalloc<float>(P),
          *dmaps = dalloc<float>(9 * P);
    float* small = dalloc<float>(32);
    isogs_adam_state* adam_state = reinterpret_cast<isogs_adam_state*>(dalloc<unsigned char>(sizeof(isogs_adam_state)));
    // the device-side frame slot: ground truth, keyframe pose, 1 / #valid depth (all pixels are valid here)
    isogs_frame fr_h{};
    fr_h.gt_rgb = d_rgb; fr_h.gt_depth = d_depth;
    for (int k = 0; k < 4; ++k) fr_h.pose[k] = pose.q[k];
    for (int k = 0; k < 3; ++k) fr_h.pose[4 + k] = pose.t[k];
    fr_h.inv_n_depth = 1.0f / (float)P;
    isogs_frame* fr_d = reinterpret_cast<isogs_frame*>(dalloc<unsigned char>(sizeof(isogs_frame)));
    CU(cudaMemcpy(fr_d, &fr_h, sizeof(fr_h), cudaMemcpyHostToDevice));
    isogs_workspace ws{};
    ws.rec = rec; ws.v_rec = v_rec; ws.radii = radii; ws.bucket = bucket; ws.vals = vals; ws.ranges = ranges;
    ws.totals = totals; ws.sort_temp = sort_tmp; ws.sort_temp_bytes = sort_b; ws.isect_cap = isect_cap;
    ws.render = render; ws.v_render = v_render; ws.alpha = alpha; ws.last = last; ws.small = small;
    ws.adam_state = adam_state; ws.dmaps = dmaps;
    ws.tile_rects = dalloc<uint32_t>(4 * N);
    isogs_gaussians gs{};
    gs.N = N; gs.scale_dim = 1;
    gs.p[0] = means; gs.p[1] = cols; gs.p[2] = quats; gs.p[3] = logit; gs.p[4] = ls;
    const int slot[4] = {0, 1, 3, 4};                       // par / grad / moments above: means, colours, opacity, scale
    for (int g = 0; g < 4; ++g) { gs.g[slot[g]] = grad[g]; gs.m[slot[g]] = mom1[g]; gs.v[slot[g]] = mom2[g]; }
    isogs_step_params prm{};
    prm.flags = ISOGS_STEP_SSIM; prm.phases = ISOGS_PHASE_RENDER | ISOGS_PHASE_ADAM;
    prm.w_im = 0.5f; prm.w_depth = 1.0f; prm.sil_thres = 0.99f;
    for (int g = 0; g < 4; ++g) prm.lr[slot[g]] = lr[g];
    prm.beta1 = 0.9f; prm.beta2 = 0.999f; prm.eps = 1e-15f;
    // capture ONE mapping iteration (Mapper::optimize's loop body) and replay it: a single launch per iteration
    void* graph = nullptr;
    CK(isogs_graph_begin(st));
    CK(isogs_map_step(&cam, fr_d, &ws, &gs, &prm, nullptr, st));
    CK(isogs_graph_end(st, &graph));
    float* losses = small;
    float first = 0.f, lastv = 0.f;
    for (int it = 0; it < iters; ++it) {
        CK(isogs_graph_launch(graph, st));
        float h[32];
        int32_t tot[2];
        CU(cudaMemcpyAsync(h, losses, sizeof(h), cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(tot, totals, sizeof(tot), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (tot[1]) { std::fprintf(stderr, "intersection buffers overflowed (%d > %lld)\n", tot[0], (long long)isect_cap); return 6; }
        if (it == 0) first = h[0];
        lastv = h[0];
        if (it % 5 == 0 || it == iters - 1)
            std::printf("iter %3d  loss %.6f  (colour L1 %.5f, depth L1 %.5f, SSIM %.4f)  intersections %d\n", it, h[0], h[1],
                        h[2], h[27], tot[0]);
    }
    if (!(lastv < 0.7f * first) || !std::isfinite(lastv)) { std::fprintf(stderr, "loss did not fall: %f -> %f\n", first, lastv); return 7; }
    CK(isogs_graph_destroy(graph));
    std::printf("ok: loss %.6f -> %.6f over %d replays of one captured graph, no torch involved\n", first, lastv, iters);
    return 0;
}
