"""TEST INFRASTRUCTURE: runs BASELINE.json configs[1] -- the TUM-shaped 640x480 synthetic sequence, 50 frames, 40
tracking iterations per frame + silhouette densification + mapping (30 iterations every 5th frame, prune at
iterations 0 / 20, overlap keyframe selection) -- ONCE on the CPU oracle and commits the result as a golden file:
the estimated poses, the ATE (Horn-aligned RMSE, utils/eval_helpers.py:24-78) and every random draw of the schedule
(selected keyframes, per-iteration keyframe picks), so that tests/test_gpu_ate.py only replays the schedule on the
device.  Takes ~40 minutes on 8 host cores.

  python oracle/pin_ate_tum.py [--shape tum] [--frames 50] [--out tests/golden/tum_50f_oracle_traj.npz]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import isogs_b200  # noqa: E402,F401
import slam_loop  # noqa: E402
from oracle import oracle as orc  # noqa: E402

syn = importlib.import_module("isogs_slam_cpp_b200.synthetic")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="tum")
    ap.add_argument("--frames", type=int, default=50)
    ap.add_argument("--track-iters", type=int, default=40)
    ap.add_argument("--map-iters", type=int, default=30)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    out = a.out or os.path.join(ROOT, "tests", "golden", f"{a.shape}_{a.frames}f_oracle_traj.npz")
    cfg = dict(slam_loop.CONFIG, frames=a.frames, track_iters=a.track_iters, map_iters=a.map_iters)
    intr = syn.SHAPES[a.shape][:6]
    W, H, fx, fy, cx, cy = intr
    frames, gt = slam_loop.frames_of(syn, a.shape, a.frames)
    c0, d0 = frames[0]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    pts, cols, msd = syn.backproject(c0, d0[None], K, np.eye(4))
    params = syn.init_params(pts, cols, msd, isotropic=True, seed=0)       # initialize_first_timestep
    ocam = orc.make_camera(W, H, fx, fy, cx, cy)
    be = slam_loop.OracleBackend(orc, ocam, params, intr)
    t0 = time.time()
    res = slam_loop.run(be, frames, intr, cfg, log=lambda s: print(f"[{time.time() - t0:6.0f} s] {s}", flush=True))
    ate, centres = slam_loop.ate_of(res["poses"], gt)
    print(f"oracle ATE RMSE {ate * 1e3:.3f} mm over {a.frames} frames, {be.count()} Gaussians, {time.time() - t0:.0f} s")
    sched = res["schedule"]
    np.savez_compressed(out, poses=res["poses"], ate=np.array([ate]), gt_centres=gt, centres=centres,
                        counts=np.array(res["counts"], np.int64), n_init=np.array([params["means3D"].shape[0]]),
                        config=np.array([json.dumps(cfg)]), shape=np.array([a.shape]),
                        schedule=np.array([json.dumps({k: {str(f): v for f, v in d.items()} for k, d in sched.items()})]))
    print("wrote", out)


if __name__ == "__main__":
    main()
