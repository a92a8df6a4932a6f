"""Is the fused step launch-bound at small sizes?  Issues K track_steps / map_steps without synchronising and
compares the host time to enqueue them with the device time to run them (CUDA events)."""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isogs_b200  # noqa: E402

pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn = importlib.import_module("isogs_slam_cpp_b200.synthetic")


def probe(shape, K=200):
    dev = torch.device("cuda", 0)
    W, H, fx, fy, cx, cy = syn.SHAPES[shape][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(4, step=1.0)
    c0, d0, Kc = syn.render_frame(shape, traj[0], seed=0)
    pts, cols, msd = syn.backproject(c0, d0, Kc, traj[0])
    gmap = ss.GaussianMap(syn.init_params(pts, cols, msd, isotropic=True, seed=0), device=dev)
    ws = ss.StepWorkspace(cam, gmap.count, dev)
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    c1, d1, _ = syn.render_frame(shape, traj[1], seed=1)
    rgb, depth = tt(c1), tt(d1[0])
    ident = (np.array([1, 0, 0, 0], np.float32), np.zeros(3, np.float32))
    pose = pkg.make_pose(*ident)
    inv_nd = 1.0 / max(1, int((d1 > 0).sum()))
    out = dict(shape=shape, gaussians=gmap.count)
    for name, fn in (("track", lambda: ss.track_step(ws, gmap, ps, rgb, depth)),
                     ("map", lambda: ss.map_step(ws, gmap, pose, rgb, depth, inv_nd))):
        ps = ss.PoseState(*ident, dev)
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for _ in range(K):
            fn()
        host = time.perf_counter() - t0
        e1.record()
        torch.cuda.synchronize()
        out[name] = dict(host_enqueue_us=1e6 * host / K, device_us=1e3 * e0.elapsed_time(e1) / K)
    return out


if __name__ == "__main__":
    for shape in sys.argv[1:] or ["tum", "replica"]:
        print(json.dumps(probe(shape)))
