"""Eager vs graph replay of the fused steps at a bench shape: device time (CUDA events) and host enqueue time per
step.  MEASUREMENT script.   python scripts/graph_probe.py [shape]"""
import importlib
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import isogs_b200  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "replica"
n = {"tum": 307200, "scannetpp": 3000000}.get(shape, 1000000)
pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn, params, kfs = bench.build_workload(shape, n, 8)
W, H, fx, fy, cx, cy = syn.SHAPES[shape][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
dev = torch.device("cuda:0")
gmap = ss.GaussianMap(params, device=dev)
ws = ss.StepWorkspace(cam, gmap.count, dev)
for k in kfs:
    k["pose"] = pkg.make_pose(k["q"], k["t"])
    k["rgb_d"] = torch.from_numpy(k["rgb"]).to(dev)
    k["depth_d"] = torch.from_numpy(k["depth"]).to(dev)
ps = ss.PoseState(kfs[0]["q"], kfs[0]["t"], dev)


def run(fn, steps=200):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    host = (time.perf_counter() - t0) / steps
    torch.cuda.synchronize()
    return {"device_ms": e0.elapsed_time(e1) / steps, "host_enqueue_ms": host * 1e3}


out = {"shape": shape}
for graph in (False, True):
    tag = "graph" if graph else "eager"
    out[f"map_{tag}"] = run(lambda i: ss.map_step(ws, gmap, kfs[i % 8]["pose"], kfs[i % 8]["rgb_d"], kfs[i % 8]["depth_d"],
                                                  kfs[i % 8]["inv_nd"], graph=graph))
    out[f"map_ssim_{tag}"] = run(lambda i: ss.map_step(ws, gmap, kfs[i % 8]["pose"], kfs[i % 8]["rgb_d"], kfs[i % 8]["depth_d"],
                                                       kfs[i % 8]["inv_nd"], graph=graph, ssim=True))
    out[f"track_{tag}"] = run(lambda i: ss.track_step(ws, gmap, ps, kfs[1]["rgb_d"], kfs[1]["depth_d"], graph=graph))
    out[f"graphs_cached_{tag}"] = len(ws._graphs)
print(json.dumps(out, indent=1))
