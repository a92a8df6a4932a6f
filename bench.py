"""bench.py -- the driver's measurement contract for the IsoGS-SLAM hot path on B200.

A "step" is ONE MAPPING ITERATION on one keyframe: transform_to_frame + activations + fused
5-channel render (rgb, depth, depth^2; silhouette = alpha) + masked-L1 loss + backward
(blend bwd + projection bwd) + fused Adam on every Gaussian parameter
(python_ordinary/scripts/splatam.py:1412-1476 with the L1 terms; the SSIM / flat / iso variants are reported beside it).
Workload at N=1: BASELINE.json configs[2] shape -- Replica-shaped 1200x680, 1 M isotropic Gaussians
(ISOGS_BENCH_SHAPE=tum|scannetpp selects configs[1] / configs[3]).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Timing: after W warm-up steps, blocks of EXACTLY K steps are timed with CUDA events (barrier + synchronize on both
sides, max over ranks) and repeated until >= 1 s of device time is covered; `ms_per_step` / `value` are the MEDIAN
block (the first block alone is `first_block_ms_per_step`).  Per-stage times come from a separate instrumented
block (events between the kernels of the same step).

N > 1 (torchrun): every rank renders a DIFFERENT keyframe of the 8-keyframe window against the replicated map; one
average of the flat gradient buffer per iteration (own NVLink peer-memory kernel, --allreduce nccl for
ncclAllReduce); weak scaling, value = keyframe-renders (iterations x ranks) per second.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SHAPE = os.environ.get("ISOGS_BENCH_SHAPE", "replica")
N_GAUSS = int(os.environ.get("ISOGS_BENCH_N", {"tum": "307200", "scannetpp": "3000000"}.get(SHAPE, "1000000")))
WINDOW = 8
MIN_TIMED_S = float(os.environ.get("ISOGS_BENCH_MIN_S", "1.0"))
# png_depth_scale of the reference's datasets (cpp/src/datasets/tum_dataset.cpp:112)
DEPTH_SCALE = {"tum": 5000.0, "replica": 6553.5, "scannetpp": 1000.0, "tiny": 1000.0}
SMS, SCHEDULERS_PER_SM = 148, 4


def _count(n):
    return f"{n / 1e6:g}M" if n >= 1_000_000 else f"{n / 1e3:g}K"


def metric_string(W, H, N):
    return f"mapping iters/s (render fwd+bwd + L1 loss + Adam) @{W}x{H}, {_count(N)} iso Gaussians"


def workload_string(shape, W, H, N, window):
    """Identical in the GPU arm and the reference arm (the driver compares the strings)."""
    return (f"{shape}-shaped {W}x{H}, N={N} isotropic Gaussians, {window}-keyframe window, one keyframe per rank per "
            f"iteration, uint8 colour + uint16 depth frames")


LOSS_STRING = "masked L1 colour+depth (north_star (5))"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), float(j.get("sm_max_mhz", 1965.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock / power / throttle reasons sampled DURING the timed region: NVML in-process (a query takes well under
    a millisecond); nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
    BITS = (0x8, 0x40, 0x20, 0x4)   # nvmlClocksEventReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.nvml = None
        self.t0 = time.perf_counter()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        mhz = int(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        mask = int(get(self.handle))
        try:
            watts = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
        except Exception:
            watts = float("nan")
        return [str(mhz), str(self.max_mhz)] + ["Active" if mask & b else "Not Active" for b in self.BITS] + [f"{watts:.1f}"]

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    self.rows.append([time.perf_counter() - self.t0] + self._sample_nvml())
                    time.sleep(0.002)
                    continue
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([time.perf_counter() - self.t0] + [c.strip() for c in out.strip().split(",")])
            except Exception:
                self.nvml = None
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        rows = [r for r in self.rows if len(r) >= 7 and str(r[1]).isdigit()]
        sm = [int(r[1]) for r in rows]
        mx = [int(r[2]) for r in rows if str(r[2]).isdigit()]
        reasons = set()
        for r in rows:
            for name, v in zip(self.NAMES, r[3:7]):
                if str(v).lower().startswith("active"):
                    reasons.add(name)

        def watts(r):
            try:
                return float(r[7])
            except Exception:
                return float("nan")
        pw = [w for w in (watts(r) for r in rows) if w == w]
        stride = max(1, len(rows) // 24)
        trace = [[round(1e3 * r[0], 1), int(r[1]), round(watts(r), 0)] for r in rows[::stride]][:32]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_median": float(np.median(pw)) if pw else None, "power_w_max": max(pw) if pw else None,
                "source": "nvml" if self.nvml is not None else "nvidia-smi",
                "trace_ms_mhz_w": trace}


def build_workload(shape, n_gauss, window):
    """Map + keyframe window.  Frames are quantised to the reference's SOURCE formats (uint8 colour, uint16 depth with
    the dataset's png_depth_scale, basedataset.py:299-340) and every arm -- resident, end-to-end, CPU oracle -- sees
    the same de-quantised fp32 values."""
    import importlib

    import isogs_b200  # noqa: F401
    syn = importlib.import_module("isogs_slam_cpp_b200.synthetic")
    params, frames, K = syn.make_map(shape, n_gauss, isotropic=True, seed=0, max_frames=max(window, 8))
    traj = syn.trajectory(64, step=4.0)
    while len(frames) < window:
        k = len(frames)
        c, d, _ = syn.render_frame(shape, traj[k], seed=k)
        frames.append((c, d, traj[k]))
    frames = frames[:window]
    scale = np.float32(DEPTH_SCALE.get(shape, 1000.0))
    inv_scale = np.float32(1.0 / float(scale))
    keyframes = []
    for c, d, w2c in frames:
        q, t = syn.w2c_to_pose(w2c)
        rgb8 = np.ascontiguousarray(np.clip(np.rint(c * 255.0), 0, 255).astype(np.uint8).transpose(1, 2, 0))   # HWC
        d16 = np.ascontiguousarray(np.clip(np.rint(d[0] * float(scale)), 0, 65535).astype(np.uint16))
        rgb = np.ascontiguousarray((rgb8.astype(np.float32) / np.float32(255.0)).transpose(2, 0, 1))
        depth = np.ascontiguousarray(d16.astype(np.float32) * inv_scale)
        keyframes.append(dict(rgb=rgb, depth=depth, rgb8=rgb8, d16=d16, inv_scale=float(inv_scale), q=q, t=t,
                              inv_nd=1.0 / max(1, int((depth > 0).sum()))))
    return syn, params, keyframes


def algorithmic_bytes(N, V, M, P, n_tiles):
    """SURVEY.md 8(d) per-stage algorithmic bytes (isotropic: g_in 20 B, g_col 12 B, rec 32 B)."""
    g_in, g_col, rec = 20, 12, 32
    return {
        "preprocess_fwd": N * (g_in + g_col) + V * (rec + 4),
        "tile_bin": 12 * M + 24 * M + 4 * M + 8 * n_tiles,
        "blend_fwd": M * (4 + rec + 12) + P * 9 * 4,
        "loss": P * (7 + 4) * 4 + P * 7 * 4,
        "blend_bwd": M * (4 + rec + 12) + P * 9 * 4 + 2 * V * 40,
        "preprocess_bwd": V * (40 + g_in) + N * (g_in + g_col),
        "adam": (N * (g_in + g_col) // 4) * 28,
    }


def launches_per_step(world, ssim=False):
    """Own kernels enqueued by one isogs_map_step (csrc/fused_step.cu): step_begin, preprocess_fwd (tile counts fused),
    tile_scan, tile_scatter, tile_sort_merge, tile_sort (long tiles), blend_fwd, loss_l1, blend_bwd,
    preprocess_bwd, adam = 11; + ssim fwd / bwd; at N > 1 the exchange kernel (reduce + Adam, replacing the Adam launch)
    and its wait.  Checked against the ncu launch list under profiles/."""
    return 11 + (2 if ssim else 0) + (1 if world > 1 else 0)


_JSON_FD = 1


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD, (line + "\n").encode())


def run_ours(args):
    import importlib

    import torch
    import torch.distributed as dist

    import isogs_b200
    pkg = isogs_b200.pkg
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    syn, params, kfs = build_workload(SHAPE, N_GAUSS, WINDOW)
    W, H, fx, fy, cx, cy = syn.SHAPES[SHAPE][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    L = pkg.lib()
    # the one exchange step of the sharded path: the gradient average.  Default: our own two-shot kernel over
    # NVLink peer memory (csrc/peer.cu); NCCL when asked for, or when CUDA IPC is unavailable on the box.
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    ar, ar_kind, grad_alloc = None, "none", None
    fused = False
    if world > 1:
        ar, ar_kind = dp.allreduce_mean_, "nccl all-reduce (ncclAvg)"
        kind = args.allreduce
        if kind == "nvls":                             # in-switch reduction; falls back to the peer-memory kernel
            peer = dp.NvlsReduceAdam(dev)
            try:
                peer.alloc(1024)
                ar, grad_alloc, fused = peer, peer.alloc, True
                ar_kind = ("reduce (multimem.ld_reduce inside the NVSwitch) + Adam + parameter broadcast (multimem.st) "
                           "fused in ONE kernel (isogs_reduce_adam_nvls, optimizer state sharded)")
            except RuntimeError as e:
                kind = "fused"
                if rank == 0:
                    print(f"[bench] NVLS unavailable, using the peer-memory kernel: {e}", file=sys.stderr)
        if kind in ("p2p", "fused"):
            fused = kind == "fused"
            peer = dp.PeerReduceAdam(dev) if fused else dp.PeerAllReduce(dev)
            try:
                peer.alloc(1024)                       # probe: cudaMalloc + IPC export + peer mapping on every rank
                ar, grad_alloc = peer, peer.alloc
                ar_kind = ("reduce + Adam + parameter broadcast fused in ONE kernel over NVLink peer memory "
                           "(isogs_reduce_adam_p2p, optimizer state sharded)" if fused else
                           "own two-shot kernel over NVLink peer memory (isogs_allreduce_avg_p2p)")
            except RuntimeError as e:
                fused = False
                ar_kind += f"; p2p unavailable: {e}"
                if rank == 0:
                    print(f"[bench] peer all-reduce unavailable, using NCCL: {e}", file=sys.stderr)
    gmap = ss.GaussianMap(params, device=dev, grad_alloc=grad_alloc)
    if fused:
        ar.adopt(gmap)                             # parameter tensors into IPC-shared memory (peers store into them)
    ws = ss.StepWorkspace(cam, gmap.count, dev)
    P = W * H
    # keyframes resident in HBM (for `value`) and, as uint8 / uint16 source frames, in pinned host memory (for `e2e`)
    for k in kfs:
        k["pose"] = pkg.make_pose(k["q"], k["t"])
        k["rgb_d"] = torch.from_numpy(k["rgb"]).to(dev)
        k["depth_d"] = torch.from_numpy(k["depth"]).to(dev)
        k["rgb8_h"] = torch.from_numpy(k["rgb8"]).pin_memory()
        k["d16_h"] = torch.from_numpy(k["d16"].view(np.int16)).pin_memory()
    # e2e arm: every step's keyframe comes from PINNED HOST memory in the reference's source formats (5 bytes per
    # pixel).  The copy of step i+1 runs on a copy stream while step i computes (two staging slots) and isogs_unpack_rgbd
    # expands it into the fp32 planes the step reads on that stream as well; the loss scalars of step i are read on the host while step i+1
    # is in flight (one step late), so no step waits on the PCIe round trip.
    copy_stream = torch.cuda.Stream(device=dev)
    slots = [dict(rgb8=torch.empty((H, W, 3), dtype=torch.uint8, device=dev),
                  d16=torch.empty((H, W), dtype=torch.int16, device=dev),
                  rgb=torch.empty((3, H, W), dtype=torch.float32, device=dev),
                  depth=torch.empty((H, W), dtype=torch.float32, device=dev),
                  ready=torch.cuda.Event(), free=torch.cuda.Event(),
                  loss=torch.empty((4,), dtype=torch.float32).pin_memory(), loss_ev=torch.cuda.Event())
             for _ in range(2)]
    e2e_state = dict(prefetched=-1, last_loss=0.0)
    use_graph = not args.no_graph

    def step_resident(i, marks=None):
        k = kfs[dp.keyframe_for(i, rank, world, WINDOW)]
        return ss.map_step(ws, gmap, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"], allreduce=ar, marks=marks,
                           graph=use_graph)

    def e2e_prefetch(i):
        s, k = slots[i % 2], kfs[dp.keyframe_for(i, rank, world, WINDOW)]
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(s["free"])          # the step that last used this slot has finished
            s["rgb8"].copy_(k["rgb8_h"], non_blocking=True)
            s["d16"].copy_(k["d16_h"], non_blocking=True)
            # the unpack into the fp32 planes rides on the copy stream too, behind its copy: it overlaps the step in flight
            pkg._lib.check(L.isogs_unpack_rgbd(C.byref(cam), C.c_void_p(s["rgb8"].data_ptr()), C.c_void_p(s["d16"].data_ptr()),
                                               k["inv_scale"], C.c_void_p(s["rgb"].data_ptr()), C.c_void_p(s["depth"].data_ptr()),
                                               C.c_void_p(copy_stream.cuda_stream)))
            s["ready"].record(copy_stream)
        e2e_state["prefetched"] = i

    def step_e2e(i):
        cur = torch.cuda.current_stream()
        if e2e_state["prefetched"] < i:
            e2e_prefetch(i)
        e2e_prefetch(i + 1)
        s, k = slots[i % 2], kfs[dp.keyframe_for(i, rank, world, WINDOW)]
        cur.wait_event(s["ready"])                     # copy + unpack of THIS step's frame have finished
        losses = ss.map_step(ws, gmap, k["pose"], s["rgb"], s["depth"], k["inv_nd"], allreduce=ar, graph=use_graph)
        s["free"].record(cur)
        s["loss"].copy_(losses, non_blocking=True)
        s["loss_ev"].record(cur)
        p = slots[(i + 1) % 2]                          # the previous step's loss: read it now
        if i > 0:
            p["loss_ev"].synchronize()
            e2e_state["last_loss"] = float(p["loss"][0])
        return e2e_state["last_loss"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    init_p = {k: gmap.p[k][:gmap.count].clone() for k in ss.PARAM_ORDER}

    def restore(g=None, init=None):
        """Every timed block starts from the SAME map (the initial one) with a fresh optimizer, as every mapping call
        of the reference does (splatam.py:1406): thousands of back-to-back Adam steps would otherwise let the scene
        drift (larger splats, more tile intersections) and later blocks / legs would time a different workload."""
        g, init = g or gmap, init or init_p
        for k in ss.PARAM_ORDER:
            g.p[k][:g.count].copy_(init[k])
        g.reset_optimizer()

    def timed(fn, steps, marks_list=None, first=0, reset=None):
        (reset or restore)()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(first, first + steps):
            if marks_list is not None:
                m = []
                fn(i, m)
                marks_list.append(m)
            else:
                fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed_blocks(fn, steps, min_s=MIN_TIMED_S, max_blocks=4000, reset=None):
        """Blocks of exactly `steps` steps until min_s of device time is covered; every rank derives the same block
        count from the max-over-ranks time of the first block.  Every block starts from the initial map."""
        first = timed(fn, steps, reset=reset)
        # 5 % head-room: the first block (cold clocks / caches) is the slowest, so later blocks are shorter than it
        n_more = int(min(max_blocks, max(0, np.ceil(1.05 * min_s * 1e3 / max(first, 1e-3)) - 1)))
        blocks = [first] + [timed(fn, steps, first=(b + 1) * steps, reset=reset) for b in range(n_more)]
        return blocks

    barrier()                                      # ranks leave the (CPU-side) workload build seconds apart
    for i in range(max(3, args.warmup)):
        step_resident(i)
    if world > 1 and hasattr(ar, "check_status"):
        # a peer spin that timed out during warm-up (bounded at 2 s in csrc/peer.cu) leaves garbage: every rank
        # then drops to NCCL together, on a freshly initialised map
        torch.cuda.synchronize()
        bad = (ar.status != 0).to(torch.int32)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX)
        if int(bad.item()):
            ar.status.zero_()
            ar_kind = "nccl all-reduce (ncclAvg); p2p timed out during warm-up"
            if rank == 0:
                print("[bench] peer all-reduce timed out during warm-up, using NCCL", file=sys.stderr)
            ws.drop_graphs()
            gmap.grad_flat, gmap.g = None, {}
            if fused:
                gmap.p = {}
                ar.drop_params()
                fused = False
            ar.release()
            ar = dp.allreduce_mean_
            gmap = ss.GaussianMap(params, device=dev)
            init_p = {k: gmap.p[k][:gmap.count].clone() for k in ss.PARAM_ORDER}
            barrier()
            for i in range(max(3, args.warmup)):
                step_resident(i)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    blocks = timed_blocks(step_resident, args.steps)
    clocks = sampler.summary() if sampler else None
    ms_block = float(np.median(blocks))
    # own kernel launches per step: COUNTED from the captured graph of the headline step (kernel nodes), plus the
    # exchange kernels enqueued outside the graph at N > 1; the constant of launches_per_step() only in eager mode
    launches_step, launches_note = launches_per_step(world), "eager mode: launches_per_step() (checked against the ncu launch list)"
    if use_graph:
        gc = ss.graph_node_counts(ws)
        head = [v for k, v in gc.items() if k[0] == "map" and k[1] == 0]
        adam_g = [v for k, v in gc.items() if k[0] == "adam"]
        if head:
            outside = 0 if world == 1 else (2 if hasattr(ar, "fused_adam") or hasattr(ar, "check_status") else 0)
            launches_step = head[0][0] + (adam_g[0][0] if adam_g and not hasattr(ar, "fused_adam") else 0) + outside
            launches_note = (f"own kernels per timed block of `steps` iterations, counted: {head[0][0]} kernel nodes (+ {head[0][1]} "
                             f"memset node) in the captured graph of the step"
                             + (f", {adam_g[0][0]} in the Adam graph" if adam_g and not hasattr(ar, "fused_adam") else "")
                             + (f", {outside} exchange kernels enqueued per step outside the graph" if outside else "")
                             + "; x steps")
    # per-stage device time: a separate INSTRUMENTED block (events recorded between the kernels of each step)
    marks_list = []
    ms_marked = timed(step_resident, args.steps, marks_list)
    M = ws.M
    V = int((ws.radii[:gmap.count, 0] > 0).sum().item())
    stage_ms = {}
    for m in marks_list:
        for (n0, ev0), (n1, ev1) in zip(m[:-1], m[1:]):
            stage_ms[n1] = stage_ms.get(n1, 0.0) + ev0.elapsed_time(ev1)
    stage_ms = {k: v / len(marks_list) for k, v in stage_ms.items()}
    # per-rank view of the exchange (N > 1): when each rank finished its backward relative to its own step start, and
    # how long its all-reduce stage took -- separates rank skew (different M per keyframe) from wire time
    rank_stats = None
    if world > 1:
        def span(m, a, b):
            d = dict(m)
            return d[a].elapsed_time(d[b]) if a in d and b in d else 0.0
        mine = torch.tensor([np.mean([span(m, "begin", "preprocess_bwd") for m in marks_list]),
                             np.mean([span(m, "preprocess_bwd", "allreduce") for m in marks_list])], device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        a = torch.stack(allr).cpu().numpy()
        rank_stats = {"backward_done_ms": {"min": float(a[:, 0].min()), "median": float(np.median(a[:, 0])),
                                           "max": float(a[:, 0].max())},
                      "allreduce_stage_ms": {"min": float(a[:, 1].min()), "median": float(np.median(a[:, 1])),
                                             "max": float(a[:, 1].max())},
                      "note": "allreduce stage = wait for the slowest rank's gradients + wire time; its minimum over "
                              "ranks (the rank that arrives last waits for nobody) approximates the wire time alone"}
    # end-to-end arm (host buffers, H2D + D2H inside the timed region)
    for i in range(3):
        step_e2e(i)
    torch.cuda.synchronize()
    e2e_state["prefetched"] = -1
    e2e_blocks = timed_blocks(lambda i: step_e2e(i), args.steps)
    ms_e2e = float(np.median(e2e_blocks))
    extras = {}
    if world == 1:
        # tracking iterations (single GPU by nature)
        k = kfs[1]
        ps = ss.PoseState(kfs[0]["q"], kfs[0]["t"], dev)
        for i in range(3):
            ss.track_step(ws, gmap, ps, k["rgb_d"], k["depth_d"], graph=use_graph)
        tb = timed_blocks(lambda i: ss.track_step(ws, gmap, ps, k["rgb_d"], k["depth_d"], graph=use_graph), args.steps, 0.3)
        extras["tracking_iters_per_s"] = args.steps / (float(np.median(tb)) * 1e-3)

        # mapping with the reference's full colour term 0.8*L1 + 0.2*(1-SSIM)
        def step_ssim(i, marks=None):
            k = kfs[i % WINDOW]
            return ss.map_step(ws, gmap, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"], marks=marks, ssim=True,
                               graph=use_graph)
        for i in range(3):
            step_ssim(i)
        sb = timed_blocks(step_ssim, args.steps, 0.3)
        extras["mapping_with_ssim_iters_per_s"] = args.steps / (float(np.median(sb)) * 1e-3)
        ssim_marks = []
        timed(step_ssim, args.steps, ssim_marks)
        tot = sum(dict(m)["loss"].elapsed_time(dict(m)["ssim"]) for m in ssim_marks)
        extras["ssim_fwd_bwd_ms"] = tot / max(1, len(ssim_marks))
        # the reference's COMPLETE mapping loss (L1 + SSIM + 50*flat + 2*iso, SURVEY 8(f)-1,2) needs three-axis scales
        # (splatam.py:361-368 turns the isotropic init into three axes + N(0, 0.01))
        p3 = dict(params)
        rs = np.random.RandomState(0)
        p3["log_scales"] = (np.tile(np.asarray(params["log_scales"]).reshape(-1, 1), (1, 3)) +
                            0.01 * rs.randn(gmap.count, 3)).astype(np.float32)
        gmap3 = ss.GaussianMap(p3, device=dev)
        init3 = {k: gmap3.p[k][:gmap3.count].clone() for k in ss.PARAM_ORDER}
        reset3 = lambda: restore(gmap3, init3)

        def step_full(i, marks=None):
            k = kfs[i % WINDOW]
            return ss.map_step(ws, gmap3, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"], marks=marks, ssim=True,
                               isoflat=True, iso_seed=i, graph=use_graph)
        for i in range(3):
            step_full(i)
        fb = timed_blocks(step_full, args.steps, 0.3, reset=reset3)
        extras["mapping_full_loss_3axis_iters_per_s"] = args.steps / (float(np.median(fb)) * 1e-3)
        full_marks = []
        timed(step_full, args.steps, full_marks, reset=reset3)
        tot = 0.0
        for m in full_marks:
            names = [n for n, _ in m]
            i1 = names.index("isoflat")
            tot += m[i1 - 1][1].elapsed_time(m[i1][1])
        extras["iso_flat_knn_ms"] = tot / max(1, len(full_marks))
        ws.drop_graphs()
        del gmap3
    # parity sample: k identical steps from the initial map on this GPU; the cpu_baseline leg runs the same steps on
    # the oracle and compares (outside every timed region)
    parity_gpu = None
    if world == 1 and not args.no_cpu:
        g2 = ss.GaussianMap(params, device=dev)
        pl = []
        for i in range(2):
            k = kfs[i]
            pl.append(float(ss.map_step(ws, g2, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"])[0].item()))
        parity_gpu = dict(losses=pl, params={n: g2.view(n).cpu().numpy().copy() for n in
                                             ("means3D", "rgb_colors", "logit_opacities", "log_scales")})
        del g2
    replicas_ok = None
    if world > 1:
        if hasattr(ar, "check_status"):
            ar.check_status()                      # a peer spin that timed out would have left garbage
        replicas_ok = dp.replicas_identical([gmap.p[k][:gmap.count] for k in ("means3D", "rgb_colors", "logit_opacities")])
        if hasattr(ar, "release"):
            ws.drop_graphs()
            gmap.grad_flat, gmap.g = None, {}
            if fused:
                init_p.clear()
                gmap.p = {}
                ar.drop_params()
            ar.release()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, sm_max_mhz, peak_src = peaks()
    ab = algorithmic_bytes(gmap.count, V, M, P, cam.n_tiles)
    dom = max((s for s in stage_ms if s in ab), key=lambda s: stage_ms[s])
    achieved = ab[dom] / (stage_ms[dom] * 1e-3) / 1e9
    # DRAM bytes / issued warp instructions of the dominant kernel from the committed ncu capture OF THIS SHAPE
    traffic, issue = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp)).get("shapes", {}).get(SHAPE, {}).get(dom)
        if tj:
            traffic = tj.get("dram_bytes_per_launch")
            wi = tj.get("warp_inst_per_launch")
            if wi:
                clock_hz = 1e6 * float((clocks or {}).get("sm_mhz") or sm_max_mhz)
                floor_ms = wi / (SMS * SCHEDULERS_PER_SM * clock_hz) * 1e3
                issue = {"warp_inst": wi, "issue_slots_per_s": SMS * SCHEDULERS_PER_SM * clock_hz, "floor_ms": floor_ms,
                         "frac": floor_ms / stage_ms[dom], "issue_active_pct_ncu": tj.get("issue_active_pct"),
                         "source": tj.get("source")}
    raster_ms = sum(stage_ms.get(s, 0.0) for s in ("preprocess_fwd", "tile_bin", "blend_fwd", "blend_bwd", "preprocess_bwd"))
    step_bytes = sum(ab.values())
    ms_step = ms_block / args.steps
    out = {
        "metric": metric_string(W, H, gmap.count), "value": args.steps * world / (ms_block * 1e-3), "unit": "iters/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(SHAPE, W, H, gmap.count, WINDOW), "loss": LOSS_STRING,
                   "n_isects": M,
                   "l2": "working set per step ~%.0f MB > 126 MB L2; keyframes rotate every step" % (step_bytes / 1e6 + 300),
                   "parallelism": f"keyframe-dp{world}" if world > 1 else "single",
                   "host_path": ("one CUDA-graph launch per iteration (isogs_map_step captured once, frame through the "
                                 "device-side slot)" if use_graph else "one isogs_map_step call per iteration (eager)"),
                   "allreduce": ar_kind, "replicas_identical_after_run": replicas_ok},
        "timing": {"blocks": len(blocks), "steps_per_block": args.steps, "timed_s": sum(blocks) * 1e-3,
                   "first_block_ms_per_step": blocks[0] / args.steps,
                   "min_block_ms_per_step": min(blocks) / args.steps, "max_block_ms_per_step": max(blocks) / args.steps,
                   "instrumented_block_ms_per_step": ms_marked / args.steps,
                   "note": "ms_per_step / value = median block; stage_ms from the instrumented block"},
        "raster_fwd_bwd_ms": raster_ms,
        "stage_ms": {k: round(v, 4) for k, v in stage_ms.items() if k != "begin"},
        "step_hbm_frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak,
        "e2e": {"value": args.steps * world / (ms_e2e * 1e-3), "unit": "iters/s",
                "h2d_bytes_per_step": 5 * P, "d2h_bytes_per_step": 16 + 8, "blocks": len(e2e_blocks),
                "note": "keyframe uint8 rgb + uint16 depth (the reference's source formats) copied from pinned host memory "
                        "EVERY step (prefetched one step ahead on a copy stream, two staging slots) and unpacked on the "
                        "device; loss scalars + intersection count read back every step, consumed one step late"},
        "gpu_launches": launches_step * args.steps,
        "gpu_launches_note": launches_note,
        "roofline": {"kernel": dom, "bound": "issue" if dom.startswith("blend") else "hbm", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes": ab[dom], "kernel_ms": stage_ms[dom], "issue": issue,
                     "note": "achieved / peak / frac are the HBM roofline on SURVEY 8(d)'s algorithmic bytes; the blend kernels "
                             "are instruction-issue bound (bound = issue): issue.frac = warp instructions of one launch / "
                             "(148 SMs x 4 schedulers x SM clock) / measured kernel time; traffic / warp_inst from the "
                             "committed ncu capture of this shape (profiles/traffic.json), null when none exists"},
        "clocks": clocks,
    }
    out.update(extras)
    if rank_stats:
        out["rank_stats"] = rank_stats
    if world == 1 and not args.no_cpu:
        cb, parity, ref_gpu = cpu_baseline(syn, params, kfs, parity_gpu=parity_gpu)
        out["cpu_baseline"] = cb
        out["parity"] = parity
        if ref_gpu:
            out["reference_gpu_kernels_ms"] = ref_gpu
    emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def oracle_map_iteration(orc, ocam, state, kf):
    """One mapping iteration of the CPU oracle (checker used as the timed CPU baseline)."""
    flags = orc.F_LOG_SCALES | orc.F_LOGIT_OPAC
    pose = orc.make_pose(kf["q"], kf["t"])
    p = state["p"]
    lo = p["logit_opacities"].reshape(-1)
    o = orc.render_full(ocam, p["means3D"], p["log_scales"], None, lo, p["rgb_colors"], ch=5, flags=flags, pose=pose)
    v_render, losses = orc.loss_l1(ocam, 1, o["render"], o["alpha"], kf["rgb"], kf["depth"])
    v_rec = orc.blend_bwd(ocam, 5, o["rec"], o["vals"], o["ranges"], o["alpha"], o["last"], v_render)
    g = orc.preprocess_bwd(ocam, p["means3D"], p["log_scales"], None, lo, p["rgb_colors"], o["radii"], o["rec"],
                           v_rec, ch=5, flags=flags, pose=pose)
    state["step"] += 1
    lrs = dict(means3D=1e-4, rgb_colors=2.5e-3, logit_opacities=0.05, log_scales=1e-3)
    for k, gk in (("means3D", g["means"]), ("rgb_colors", g["colors"]), ("logit_opacities", g["opacities"]),
                  ("log_scales", g["scales"])):
        orc.adam(p[k].reshape(-1), gk.reshape(-1), state["m"][k].reshape(-1), state["v"][k].reshape(-1), lrs[k],
                 eps=1e-15, step=state["step"])
    return float(losses[0])


def oracle_state(params):
    p = {k: np.ascontiguousarray(v).copy() for k, v in params.items()}
    return dict(p=p, m={k: np.zeros_like(v) for k, v in p.items()}, v={k: np.zeros_like(v) for k, v in p.items()},
                step=0)


def host_threads():
    """All the host threads this process may use (launchers such as torchrun export OMP_NUM_THREADS=1)."""
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, n)


def reference_gpu_kernels(syn, params, kfs):
    """The 'kernel to beat' (SURVEY 8(d)-4): the reference's own gsplat rasterize_to_pixels_3dgs_{fwd,bwd}_kernel<5>
    (RasterizeToPixels3DGSFwd.cu:17-188, Bwd.cu:15-278; built for sm_100a with the reference's --use_fast_math by
    oracle/build_ref.sh) timed on this box on the bench workload's own records / offsets / sorted ids, next to ours.
    Baseline leg only, outside every timed region."""
    so = os.path.join(ROOT, "oracle", "_ref", "gsplat_ref.so")
    if not os.path.exists(so):
        return None
    try:
        import importlib.util

        import torch

        import isogs_b200
        pkg = isogs_b200.pkg
        spec = importlib.util.spec_from_file_location("gsplat_ref", so)
        gsref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(gsref)
        W, H, fx, fy, cx, cy = syn.SHAPES[SHAPE][:6]
        cam = pkg.make_camera(W, H, fx, fy, cx, cy)
        dev = "cuda:0"
        flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
        tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        k = kfs[1]
        means, ls = tt(params["means3D"]), tt(params["log_scales"])
        lo, rgb = tt(params["logit_opacities"].reshape(-1)), tt(params["rgb_colors"])
        render, alpha, ctx = pkg.raster.forward(cam, means, ls, None, lo, rgb, ch=5, flags=flags,
                                                pose=pkg.make_pose(k["q"], k["t"]))
        rec = ctx.rec
        m2d = rec[:, 0:2].contiguous()[None]
        conics = torch.stack([2.0 * rec[:, 2], rec[:, 3], 2.0 * rec[:, 4]], -1).contiguous()[None]
        opac = rec[:, 6].contiguous()[None]
        colors = rec[:, 7:12].contiguous()[None]
        th, tw = (H + 15) // 16, (W + 15) // 16
        offsets = ctx.ranges[:-1].reshape(1, th, tw).contiguous()
        ids = ctx.vals.contiguous()
        v_render = torch.randn((5, H, W), device=dev)
        v_render[4] = 0
        v_hwc = v_render.permute(1, 2, 0).contiguous()[None]
        v_alpha = torch.zeros((1, H, W, 1), device=dev)

        def timeit(fn, n=20, warm=5):
            for _ in range(warm):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / n
        r_ref, a_ref, l_ref = gsref.rast_fwd(m2d, conics, colors, opac, W, H, offsets, ids)
        v_rec = torch.zeros_like(rec)

        def ours_bwd():
            v_rec.zero_()
            pkg._lib.check(pkg.lib().isogs_blend_bwd(C.byref(cam), 5, ctx.N, C.c_void_p(rec.data_ptr()),
                                                     C.c_void_p(ctx.vals.data_ptr()), C.c_void_p(ctx.ranges.data_ptr()),
                                                     C.c_void_p(ctx.alpha.data_ptr()), C.c_void_p(ctx.last.data_ptr()),
                                                     C.c_void_p(v_render.data_ptr()), None, C.c_void_p(v_rec.data_ptr()),
                                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return {"gsplat_fwd": timeit(lambda: gsref.rast_fwd(m2d, conics, colors, opac, W, H, offsets, ids)),
                "gsplat_bwd": timeit(lambda: gsref.rast_bwd(m2d, conics, colors, opac, W, H, offsets, ids, a_ref, l_ref,
                                                            v_hwc, v_alpha)),
                "ours_fwd": timeit(lambda: pkg.raster.blend_forward(cam, 5, ctx.rec, ctx.vals, ctx.ranges)),
                "ours_bwd": timeit(ours_bwd),
                "max_abs_image_diff": float((render - r_ref[0].permute(2, 0, 1)).abs().max()),
                "note": "reference = rasterize_to_pixels_3dgs_{fwd,bwd}_kernel<5> of the vendored gsplat 1.5.3 compiled for "
                        "sm_100a with --use_fast_math, same records / offsets / ids, 5 channels; each timing includes the "
                        "output allocation / zero-fill its op performs"}
    except Exception as e:  # the baseline leg must never take the bench line down
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def cpu_baseline(syn, params, kfs, budget_s=12.0, max_steps=24, parity_gpu=None):
    """Bounded sample of the same workload on the CPU oracle: the first two full-size iterations double as thread-pool
    warm-up AND as the parity sample (same steps as the GPU's parity run), then full-size mapping iterations until
    ~budget_s of CPU work has been timed.  Returns (cpu_baseline, parity, reference_gpu_kernels_ms)."""
    from oracle import oracle as orc
    orc.set_num_threads(host_threads())
    W, H, fx, fy, cx, cy = syn.SHAPES[SHAPE][:6]
    ocam = orc.make_camera(W, H, fx, fy, cx, cy)
    state = oracle_state(params)
    pl = [oracle_map_iteration(orc, ocam, state, kfs[i]) for i in range(2)]
    parity = None
    if parity_gpu is not None:
        loss_rel = max(abs(a - b) / max(abs(b), 1e-30) for a, b in zip(parity_gpu["losses"], pl))
        step_rel, flips = 0.0, 0.0
        for n, got in parity_gpu["params"].items():
            init = np.asarray(params[n], np.float32).reshape(got.shape).astype(np.float64)
            d_ref = state["p"][n].reshape(got.shape).astype(np.float64) - init
            d_got = got.astype(np.float64) - init
            step_rel = max(step_rel, float(np.linalg.norm(d_got - d_ref) / max(np.linalg.norm(d_ref), 1e-30)))
            unit = np.abs(d_ref).max()
            flips = max(flips, float((np.abs(d_got - d_ref) > 0.05 * unit).mean()))
        parity = {"steps": 2, "loss_rel": loss_rel, "param_step_rel": step_rel, "param_elements_off_by_more_than_5pct_of_a_step": flips,
                  "against": "CPU oracle (oracle/isogs_oracle.cpp) running the identical two mapping iterations from the same "
                             "initial map on the same keyframes; param_step_rel = ||dP_gpu - dP_cpu||_2 / ||dP_cpu||_2, worst group"}
    n, t0 = 0, time.perf_counter()
    while True:
        oracle_map_iteration(orc, ocam, state, kfs[(n + 2) % len(kfs)])
        n += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or n >= max_steps:
            break
    cb = {"value": n / dt, "unit": "iters/s", "cores": orc.lib().orc_num_threads(), "kind": "port",
          "sample": f"{n} full-size mapping iterations of the same workload on the CPU oracle after two warm-up "
                    f"iterations (C++/OpenMP -O3, {dt:.1f} s)"}
    return cb, parity, reference_gpu_kernels(syn, params, kfs)


def run_reference(args):
    """--impl reference: the reference's CPU path for the hot path, timed on the host cores.  The
    reference's own rasterizer is CUDA-only and its Python rasterizer module is not vendored,
    so this is the oracle port (kind "port"), all host threads, same config/metric/unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from oracle import oracle as orc
    orc.build()
    orc.set_num_threads(host_threads())           # torchrun exports OMP_NUM_THREADS=1; rank 0 alone runs this arm
    syn, params, kfs = build_workload(SHAPE, N_GAUSS, WINDOW)
    W, H, fx, fy, cx, cy = syn.SHAPES[SHAPE][:6]
    ocam = orc.make_camera(W, H, fx, fy, cx, cy)
    state = oracle_state(params)
    N = params["means3D"].shape[0]
    steps = max(1, min(args.steps, 20))   # bounded sample: ~0.7 s per full-size iteration on 16 host threads
    warm = 1 if args.warmup > 0 else 0
    for i in range(warm):
        oracle_map_iteration(orc, ocam, state, kfs[i % len(kfs)])
    t0 = time.perf_counter()
    for i in range(steps):
        oracle_map_iteration(orc, ocam, state, kfs[(i + warm) % len(kfs)])
    dt = time.perf_counter() - t0
    val = steps / dt
    cores = orc.lib().orc_num_threads()
    emit(json.dumps({
        "impl": "reference", "metric": metric_string(W, H, N), "value": val, "unit": "iters/s", "n_gpus": world,
        "steps": steps, "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(SHAPE, W, H, N, WINDOW), "loss": LOSS_STRING},
        "cpu_baseline": {"value": val, "unit": "iters/s", "cores": cores, "kind": "port",
                         "sample": f"bounded: {steps} full-size mapping iterations on the CPU oracle (C++/OpenMP -O3), {dt:.1f} s"},
        "e2e": {"value": val, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity leg")
    ap.add_argument("--no-graph", action="store_true", help="enqueue every step eagerly instead of replaying the captured graph")
    ap.add_argument("--allreduce", default="fused", choices=["nvls", "fused", "p2p", "nccl"],
                    help="exchange step at N > 1: reduce + Adam fused in one NVLink peer-memory kernel (default), the "
                         "peer-memory all-reduce followed by the ordinary Adam launch, or NCCL")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner under
    # NCCL_DEBUG=VERSION, torchrun notices) goes to stderr instead -- file descriptor 1 is pointed at stderr for
    # the run and the line is written to the saved descriptor
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
