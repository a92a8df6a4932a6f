"""Time to the same loss for the keyframe-batched schedule (MEASUREMENT
INFRASTRUCTURE).  N ranks render N different keyframes per iteration and average the gradients before ONE Adam step,
so an iteration of the sharded run is a larger batch, not N sequential iterations.  The number of iterations the
batched schedule needs is a property of the arithmetic, not of the wiring: it is measured here on ONE GPU by running
the N renders of an iteration one after the other and averaging their gradient buffers (the sum order differs from
the peer kernel's at rounding level).  Wall time = iterations x the measured step time at N (bench records).

    python scripts/time_to_loss.py [--shape replica] [--iters 240] [--ms 1:0.813,2:0.864,8:0.918]"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="replica")
    ap.add_argument("--iters", type=int, default=240, help="sequential iterations that define the target loss")
    ap.add_argument("--ms", default="", help="measured ms per step at N, e.g. 1:0.813,2:0.864,4:0.89,8:0.918")
    a = ap.parse_args()
    import isogs_b200
    pkg = isogs_b200.pkg
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    n_gauss = {"tum": 307200, "scannetpp": 3000000}.get(a.shape, 1000000)
    syn, params, kfs = bench.build_workload(a.shape, n_gauss, 8)
    W, H, fx, fy, cx, cy = syn.SHAPES[a.shape][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    dev = torch.device("cuda")
    for k in kfs:
        k["pose"] = pkg.make_pose(k["q"], k["t"])
        k["rgb_d"], k["depth_d"] = torch.from_numpy(k["rgb"]).to(dev), torch.from_numpy(k["depth"]).to(dev)
    ms = {int(x.split(":")[0]): float(x.split(":")[1]) for x in a.ms.split(",") if x}

    def window_loss(ws, gmap):
        tot = 0.0
        for k in kfs:
            ss.render_forward(ws, gmap, pose=k["pose"])
            ws.losses.zero_()
            ss.loss_l1(ws, 1, k["rgb_d"], k["depth_d"], k["inv_nd"])
            tot += float(ws.losses[0])
        return tot / len(kfs)

    def run(world, max_iters, target=None):
        gmap = ss.GaussianMap(params, device=dev)
        ws = ss.StepWorkspace(cam, gmap.count, dev)
        acc = torch.zeros_like(gmap.grad_flat)
        curve = [(0, window_loss(ws, gmap))]
        for it in range(max_iters):
            acc.zero_()
            for r in range(world):
                k = kfs[(it * world + r) % len(kfs)]
                ss.render_forward(ws, gmap, pose=k["pose"])
                ws.losses.zero_()
                ss.loss_l1(ws, 1, k["rgb_d"], k["depth_d"], k["inv_nd"])
                ss.render_backward(ws, gmap, pose=k["pose"])
                acc += gmap.grad_flat
            gmap.grad_flat.copy_(acc / world)
            ss.adam_gaussians(gmap)
            if (it + 1) % 4 == 0:
                curve.append((it + 1, window_loss(ws, gmap)))
                if target is not None and curve[-1][1] <= target:
                    break
        assert not ws.check_overflow(wait=True)
        return curve

    base = run(1, a.iters)
    target = base[-1][1]
    out = {"shape": a.shape, "target_loss": target, "initial_loss": base[0][1],
           "defined_by": f"window-mean mapping loss of the sequential schedule (one keyframe per iteration) after {a.iters} iterations",
           "runs": {}}
    for world in (1, 2, 4, 8):
        c = base if world == 1 else run(world, a.iters, target)
        hit = next((i for i, l in c if l <= target), None)
        row = {"iterations_to_target": hit, "keyframe_renders_to_target": None if hit is None else hit * world,
               "final_loss": c[-1][1], "curve_every_40": [(i, round(l, 6)) for i, l in c if i % 40 == 0]}
        if hit is not None and world in ms:
            row["ms_per_step_measured"] = ms[world]
            row["time_to_target_ms"] = hit * ms[world]
        out["runs"][str(world)] = row
    if "1" in out["runs"] and "time_to_target_ms" in out["runs"]["1"]:
        t1 = out["runs"]["1"]["time_to_target_ms"]
        for w, row in out["runs"].items():
            if "time_to_target_ms" in row:
                row["speedup_to_same_loss"] = t1 / row["time_to_target_ms"]
    print("TIME_TO_LOSS " + json.dumps(out))


if __name__ == "__main__":
    main()
