// fused_step.cu -- SURVEY 8(f)-3: one host call enqueues a whole mapping / tracking iteration, and a small
// CUDA-graph API replays it.  Replaces the Python iteration body of python_ordinary/scripts/splatam.py:1247-1290
// (tracking) and :1412-1476 (mapping) -- get_loss + loss.backward() + optimizer.step(), ~60 torch launches and >= 4
// host syncs per iteration in the reference -- and Tracker::optimizePose / Mapper::optimize
// (cpp/src/slam/tracker.cpp:15-115, cpp/src/slam/mapper.cpp:34-151).
//
// Nothing here computes: the stages are the kernels of preprocess.cu / tile_bin.cu / blend_lq.cu / step.cu / ssim.cu,
// enqueued back to back on the caller's stream with no host sync and no allocation, so the call is capturable.  What
// makes a capture REPLAYABLE: the frame (ground truth, keyframe pose, 1/n_valid) is read through a device-side slot
// (isogs_frame), and Adam's step count / bias corrections live on the device (isogs_adam_state, advanced by the
// step's first kernel).
#include <cstddef>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace isogs {
int launch_loss_l1(const isogs_camera* cam, int mode, const float* render, const float* alpha, const float* gt_rgb,
                   const float* gt_depth, const isogs_frame* frame_dev, float w_im, float w_depth, float sil_thres,
                   int use_sil, float inv_n_depth, const float* outlier_thresh, float* v_render, float* losses,
                   cudaStream_t stream, const float* inv_n_dev = nullptr);
int launch_ssim_loss(const isogs_camera* cam, const float* render, const float* gt_rgb, const isogs_frame* frame_dev,
                     float weight, float* dmaps, float* v_render, float* losses, float* ssim_out, cudaStream_t stream);
int launch_outlier_threshold(const isogs_camera* cam, const float* render_depth, const float* gt_depth,
                             const isogs_frame* frame_dev, float factor, float* scratch, void* sort_temp,
                             size_t sort_temp_bytes, float* thresh_out, cudaStream_t stream, float* inv_n_out = nullptr);

namespace {
constexpr int SMALL = 32; // losses[4] | v_pose[7] | pose_acc[16] | ssim | outlier thr | geo[3]
constexpr int FLAGS_RAW = ISOGS_F_LOG_SCALES | ISOGS_F_LOGIT_OPAC;

inline int mark(void* const* events, int k, cudaStream_t s) {
    if (events != nullptr && events[k] != nullptr) ISOGS_CUDA(cudaEventRecord((cudaEvent_t)events[k], s));
    return ISOGS_OK;
}
#define ISOGS_TRY(call)            \
    do {                           \
        const int _rc = (call);    \
        if (_rc != ISOGS_OK) return _rc; \
    } while (0)

int check_common(const isogs_camera* cam, const isogs_frame* frame, const isogs_workspace* ws, const isogs_gaussians* gs,
                 const isogs_step_params* prm, const char* who) {
    if (!cam || !frame || !ws || !gs || !prm) { set_error("%s: null argument", who); return ISOGS_E_ARG; }
    if (gs->N < 0 || (gs->scale_dim != 1 && gs->scale_dim != 3)) { set_error("%s: bad gaussians", who); return ISOGS_E_ARG; }
    if (!ws->rec || !ws->v_rec || !ws->radii || !ws->bucket || !ws->vals || !ws->ranges || !ws->totals || !ws->render ||
        !ws->v_render || !ws->alpha || !ws->last || !ws->small || !ws->sort_temp) {
        set_error("%s: workspace has a null buffer", who);
        return ISOGS_E_ARG;
    }
    return ISOGS_OK;
}

// preprocess (raw parameters, frame transform fused) -> sync-free binning -> 5-channel blend
int forward(const isogs_camera* cam, const float* pose_dev, const isogs_workspace* ws, const isogs_gaussians* gs,
            void* const* ev, cudaStream_t st) {
    const float* quats = gs->scale_dim == 3 ? gs->p[2] : nullptr;
    // with the rect rows available the per-tile counts are accumulated by the projection kernel itself: the counters
    // (the head of sort_temp, where the binning keeps them) are cleared first
    const bool fused_count = ws->tile_rects != nullptr && gs->N > 0;
    const int n_tiles = ((cam->width + ISOGS_TILE - 1) / ISOGS_TILE) * ((cam->height + ISOGS_TILE - 1) / ISOGS_TILE);
    if (fused_count) {
        if (sizeof(int32_t) * (size_t)(n_tiles + 1) > ws->sort_temp_bytes) { set_error("sort temp too small for %d tiles", n_tiles); return ISOGS_E_ARG; }
        ISOGS_CUDA(cudaMemsetAsync(ws->sort_temp, 0, sizeof(int32_t) * (size_t)(n_tiles + 1), st));
    }
    ISOGS_TRY(isogs_preprocess_fwd(cam, nullptr, pose_dev, gs->N, gs->p[0], gs->p[4], gs->scale_dim, quats, gs->p[3],
                                   gs->p[1], 5, FLAGS_RAW, ws->rec, ws->radii, nullptr, nullptr, nullptr, 0, nullptr,
                                   ws->v_rec, ws->tile_rects, fused_count ? (int32_t*)ws->sort_temp : nullptr, st));
    ISOGS_TRY(mark(ev, 1, st));
    ISOGS_TRY(isogs_tile_bin_async(cam, gs->N, ws->isect_cap, ws->rec, ws->radii, ws->bucket, nullptr, ws->vals,
                                   ws->ranges, ws->totals, ws->sort_temp, ws->sort_temp_bytes, ws->tile_rects,
                                   fused_count ? 1 : 0, st));
    ISOGS_TRY(mark(ev, 2, st));
    ISOGS_TRY(isogs_blend_fwd(cam, 5, ws->rec, ws->vals, ws->ranges, ws->render, ws->alpha, ws->last, st));
    ISOGS_TRY(mark(ev, 3, st));
    return ISOGS_OK;
}
} // namespace
} // namespace isogs

using namespace isogs;

// ba != nullptr: get_loss(mapping=True, do_ba=True) (splatam.py:505-510) -- the keyframe's pose is a parameter too
struct BaPose { float *pose, *m, *v; isogs_adam_state* state; };

static int map_step_impl(const isogs_camera* cam, const isogs_frame* frame, const isogs_workspace* ws,
                         const isogs_gaussians* gs, const isogs_step_params* prm, const BaPose* ba, void* const* ev,
                         cudaStream_t st) {
    ISOGS_TRY(check_common(cam, frame, ws, gs, prm, "isogs_map_step"));
    const bool ssim = (prm->flags & ISOGS_STEP_SSIM) != 0;
    if (ssim && !ws->dmaps) { set_error("isogs_map_step: ISOGS_STEP_SSIM needs ws->dmaps"); return ISOGS_E_ARG; }
    const bool outlier = (prm->flags & ISOGS_STEP_OUTLIER) != 0;
    if (outlier && (!ws->median_scratch || !ws->median_temp)) {
        set_error("isogs_map_step: ISOGS_STEP_OUTLIER needs the median scratch");
        return ISOGS_E_ARG;
    }
    if (ba && (!ba->pose || !ba->m || !ba->v || !ba->state)) { set_error("isogs_map_step_ba: null pose state"); return ISOGS_E_ARG; }
    const float* pose_dev = ba ? ba->pose
                               : reinterpret_cast<const float*>(reinterpret_cast<const char*>(frame) + offsetof(isogs_frame, pose));
    float* losses = ws->small;
    float* v_pose = ws->small + 4;
    float* pose_acc = ws->small + 11;
    float* thr = ws->small + 28;
    if (prm->phases & ISOGS_PHASE_RENDER) {
        ISOGS_TRY(mark(ev, 0, st));
        ISOGS_TRY(isogs_step_begin(ws->small, SMALL, ws->adam_state, prm->beta1, prm->beta2, st));
        if (ba) ISOGS_TRY(isogs_step_begin(nullptr, 0, ba->state, prm->beta1, prm->beta2, st));
        ISOGS_TRY(forward(cam, pose_dev, ws, gs, ev, st));
        // ignore_outlier_depth_loss in mapping mode (splatam.py:545-549): the depth term is the MEAN over the masked
        // pixels, so the normaliser 1/n(mask) is data dependent; it comes out of the sorted errors next to the median
        // (median_scratch[0], free once the sort has run)
        if (outlier)
            ISOGS_TRY(launch_outlier_threshold(cam, ws->render + 3 * (size_t)cam->width * cam->height, nullptr, frame, 10.0f,
                                               ws->median_scratch, ws->median_temp, ws->median_temp_bytes, thr, st,
                                               ws->median_scratch));
        ISOGS_TRY(launch_loss_l1(cam, 1, ws->render, ws->alpha, nullptr, nullptr, frame, ssim ? 0.8f * prm->w_im : prm->w_im,
                                 prm->w_depth, prm->sil_thres, 1, 1.0f, outlier ? thr : nullptr, ws->v_render, losses, st,
                                 outlier ? ws->median_scratch : nullptr));
        ISOGS_TRY(mark(ev, 4, st));
        if (ssim)
            ISOGS_TRY(launch_ssim_loss(cam, ws->render, nullptr, frame, 0.2f * prm->w_im, ws->dmaps, ws->v_render, losses,
                                       ws->small + 27, st));
        ISOGS_TRY(mark(ev, 5, st));
        if (gs->N > 0)
            ISOGS_TRY(isogs_blend_bwd(cam, 5, gs->N, ws->rec, ws->vals, ws->ranges, ws->alpha, ws->last, ws->v_render,
                                      nullptr, ws->v_rec, st));
        ISOGS_TRY(mark(ev, 6, st));
        const float* quats = gs->scale_dim == 3 ? gs->p[2] : nullptr;
        ISOGS_TRY(isogs_preprocess_bwd(cam, nullptr, pose_dev, gs->N, gs->p[0], gs->p[4], gs->scale_dim, quats, gs->p[3], 5,
                                       FLAGS_RAW, ws->radii, ws->v_rec, gs->g[0], gs->g[4], gs->g[2], gs->g[3], gs->g[1],
                                       nullptr, ba ? v_pose : nullptr, ba ? pose_acc : nullptr, st));
        ISOGS_TRY(mark(ev, 7, st));
    }
    if (prm->phases & ISOGS_PHASE_ADAM) {
        isogs_adam_group grp[5];
        int n = 0;
        for (int k = 0; k < 5; ++k) {
            if (!gs->g[k] || !gs->m[k] || !gs->v[k]) continue;
            const int width = (k == 0 || k == 1) ? 3 : (k == 2 ? 4 : (k == 3 ? 1 : gs->scale_dim));
            grp[n].param = gs->p[k]; grp[n].grad = gs->g[k]; grp[n].exp_avg = gs->m[k]; grp[n].exp_avg_sq = gs->v[k];
            grp[n].numel = gs->N * width;
            grp[n].lr = prm->lr[k];
            ++n;
        }
        ISOGS_TRY(isogs_adam_fused(grp, n, prm->beta1, prm->beta2, prm->eps, 1, ws->adam_state, ws->totals + 1, st));
        if (ba) {
            isogs_adam_group pg[2];
            pg[0].param = ba->pose; pg[0].grad = v_pose; pg[0].exp_avg = ba->m; pg[0].exp_avg_sq = ba->v;
            pg[0].numel = 4; pg[0].lr = prm->lr_pose[0];
            pg[1].param = ba->pose + 4; pg[1].grad = v_pose + 4; pg[1].exp_avg = ba->m + 4; pg[1].exp_avg_sq = ba->v + 4;
            pg[1].numel = 3; pg[1].lr = prm->lr_pose[1];
            ISOGS_TRY(isogs_adam_fused(pg, 2, prm->beta1, prm->beta2, prm->eps, 1, ba->state, ws->totals + 1, st));
        }
        ISOGS_TRY(mark(ev, 8, st));
    }
    return ISOGS_OK;
}

extern "C" int isogs_map_step(const isogs_camera* cam, const isogs_frame* frame, const isogs_workspace* ws,
                              const isogs_gaussians* gs, const isogs_step_params* prm, void* const* ev, void* stream_) {
    return map_step_impl(cam, frame, ws, gs, prm, nullptr, ev, (cudaStream_t)stream_);
}

extern "C" int isogs_map_step_ba(const isogs_camera* cam, const isogs_frame* frame, const isogs_workspace* ws,
                                 const isogs_gaussians* gs, float* pose, float* pose_m, float* pose_v,
                                 isogs_adam_state* pose_state, const isogs_step_params* prm, void* const* ev,
                                 void* stream_) {
    const BaPose ba{pose, pose_m, pose_v, pose_state};
    if (prm && prm->phases != (ISOGS_PHASE_RENDER | ISOGS_PHASE_ADAM)) {
        set_error("isogs_map_step_ba: both phases in one call");
        return ISOGS_E_ARG;
    }
    return map_step_impl(cam, frame, ws, gs, prm, &ba, ev, (cudaStream_t)stream_);
}

extern "C" int isogs_track_step(const isogs_camera* cam, const isogs_frame* frame, const isogs_workspace* ws,
                                const isogs_gaussians* gs, float* pose, float* pose_m, float* pose_v, float* best,
                                isogs_adam_state* pose_state, const isogs_step_params* prm, void* const* ev, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    ISOGS_TRY(check_common(cam, frame, ws, gs, prm, "isogs_track_step"));
    if (!pose || !pose_m || !pose_v || !best) { set_error("isogs_track_step: null pose state"); return ISOGS_E_ARG; }
    const bool outlier = (prm->flags & ISOGS_STEP_OUTLIER) != 0;
    if (outlier && (!ws->median_scratch || !ws->median_temp)) {
        set_error("isogs_track_step: ISOGS_STEP_OUTLIER needs the median scratch");
        return ISOGS_E_ARG;
    }
    float* losses = ws->small;
    float* v_pose = ws->small + 4;
    float* pose_acc = ws->small + 11;
    float* thr = ws->small + 28;
    ISOGS_TRY(mark(ev, 0, st));
    ISOGS_TRY(isogs_step_begin(ws->small, SMALL, pose_state, prm->beta1, prm->beta2, st));
    ISOGS_TRY(forward(cam, pose, ws, gs, ev, st));
    if (outlier)   // 10 x median depth error (splatam.py:552-553), device-side, ground truth through the frame slot
        ISOGS_TRY(launch_outlier_threshold(cam, ws->render + 3 * (size_t)cam->width * cam->height, nullptr, frame, 10.0f,
                                           ws->median_scratch, ws->median_temp, ws->median_temp_bytes, thr, st));
    ISOGS_TRY(launch_loss_l1(cam, 0, ws->render, ws->alpha, nullptr, nullptr, frame, prm->w_im, prm->w_depth,
                             prm->sil_thres, (prm->flags & ISOGS_STEP_NO_SIL) ? 0 : 1, 1.0f, outlier ? thr : nullptr,
                             ws->v_render, losses, st));
    ISOGS_TRY(mark(ev, 4, st));
    ISOGS_TRY(mark(ev, 5, st));
    if (gs->N > 0)
        ISOGS_TRY(isogs_blend_bwd_geom(cam, 5, gs->N, ws->rec, ws->vals, ws->ranges, ws->alpha, ws->last, ws->v_render,
                                       nullptr, ws->v_rec, st));
    ISOGS_TRY(mark(ev, 6, st));
    const float* quats = gs->scale_dim == 3 ? gs->p[2] : nullptr;
    ISOGS_TRY(isogs_preprocess_bwd(cam, nullptr, pose, gs->N, gs->p[0], gs->p[4], gs->scale_dim, quats, gs->p[3], 5,
                                   FLAGS_RAW, ws->radii, ws->v_rec, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                   v_pose, pose_acc, st));
    ISOGS_TRY(mark(ev, 7, st));
    isogs_adam_group grp[2];
    grp[0].param = pose; grp[0].grad = v_pose; grp[0].exp_avg = pose_m; grp[0].exp_avg_sq = pose_v;
    grp[0].numel = 4; grp[0].lr = prm->lr[0];
    grp[1].param = pose + 4; grp[1].grad = v_pose + 4; grp[1].exp_avg = pose_m + 4; grp[1].exp_avg_sq = pose_v + 4;
    grp[1].numel = 3; grp[1].lr = prm->lr[1];
    ISOGS_TRY(isogs_adam_fused(grp, 2, prm->beta1, prm->beta2, prm->eps, 1, pose_state, ws->totals + 1, st));
    ISOGS_TRY(isogs_pose_keep_best(losses, pose, best, ws->totals + 1, st));
    ISOGS_TRY(mark(ev, 8, st));
    return ISOGS_OK;
}

// ---- CUDA-graph capture / replay ------------------------------------------------------------------------------
static std::mutex g_graph_mu;
static std::unordered_map<void*, std::pair<int32_t, int32_t>> g_graph_info;
extern "C" int isogs_graph_begin(void* stream_) {
    ISOGS_CUDA(cudaStreamBeginCapture((cudaStream_t)stream_, cudaStreamCaptureModeThreadLocal));
    return ISOGS_OK;
}

extern "C" int isogs_graph_end(void* stream_, void** graph_exec) {
    if (!graph_exec) { set_error("isogs_graph_end: null output"); return ISOGS_E_ARG; }
    cudaGraph_t g = nullptr;
    ISOGS_CUDA(cudaStreamEndCapture((cudaStream_t)stream_, &g));
    // what was captured, counted from the graph itself (bench.py reports it as gpu_launches)
    int32_t n_kernel = 0, n_other = 0;
    size_t n_nodes = 0;
    if (cudaGraphGetNodes(g, nullptr, &n_nodes) == cudaSuccess && n_nodes > 0) {
        std::vector<cudaGraphNode_t> nodes(n_nodes);
        if (cudaGraphGetNodes(g, nodes.data(), &n_nodes) == cudaSuccess)
            for (size_t i = 0; i < n_nodes; ++i) {
                cudaGraphNodeType t;
                if (cudaGraphNodeGetType(nodes[i], &t) != cudaSuccess) continue;
                if (t == cudaGraphNodeTypeKernel) ++n_kernel; else ++n_other;
            }
    }
    cudaGraphExec_t ge = nullptr;
    const int rc = check_cuda(cudaGraphInstantiate(&ge, g, 0), "cudaGraphInstantiate");
    cudaGraphDestroy(g);
    if (rc != ISOGS_OK) return rc;
    {
        std::lock_guard<std::mutex> lock(g_graph_mu);
        g_graph_info[(void*)ge] = {n_kernel, n_other};
    }
    *graph_exec = (void*)ge;
    return ISOGS_OK;
}

extern "C" int isogs_graph_info(void* graph_exec, int32_t* n_kernel_nodes, int32_t* n_other_nodes) {
    std::lock_guard<std::mutex> lock(g_graph_mu);
    const auto it = g_graph_info.find(graph_exec);
    if (it == g_graph_info.end()) { set_error("isogs_graph_info: unknown graph"); return ISOGS_E_ARG; }
    if (n_kernel_nodes) *n_kernel_nodes = it->second.first;
    if (n_other_nodes) *n_other_nodes = it->second.second;
    return ISOGS_OK;
}

extern "C" int isogs_graph_launch(void* graph_exec, void* stream_) {
    if (!graph_exec) { set_error("isogs_graph_launch: null graph"); return ISOGS_E_ARG; }
    ISOGS_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream_));
    return ISOGS_OK;
}

extern "C" int isogs_graph_destroy(void* graph_exec) {
    if (graph_exec) {
        {
            std::lock_guard<std::mutex> lock(g_graph_mu);
            g_graph_info.erase(graph_exec);
        }
        ISOGS_CUDA(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
    }
    return ISOGS_OK;
}
