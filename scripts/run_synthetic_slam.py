"""BASELINE.json configs[1]-shaped run: a synthetic RGB-D sequence through tracking (N iterations per
frame) + silhouette densification + mapping (N iterations per mapped frame, prune at iterations 0/20)
with the fused device steps (slam_step.track_step / map_step / add_new_gaussians / prune_gaussians),
i.e. the loop structure of rgbd_slam (python_ordinary/scripts/splatam.py:1200-1536) with the L1 terms.
Prints one JSON line: ATE RMSE, PSNR of the last frame, Gaussian count, tracking / mapping iters/s.

  python scripts/run_synthetic_slam.py --shape tum --frames 50 --track-iters 40 --map-iters 40
"""
import argparse
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


def centre(q, t):
    q = np.asarray(q, np.float64) / np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return -R.T @ np.asarray(t, np.float64)


def w2c_of(q, t):
    q = np.asarray(q, np.float64) / np.linalg.norm(q)
    w, x, y, z = q
    M = np.eye(4, dtype=np.float32)
    M[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                 [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                 [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    M[:3, 3] = np.asarray(t, np.float32)
    return M


def ate_rmse(est, gt):
    est, gt = np.asarray(est).T, np.asarray(gt).T       # Horn alignment, eval_helpers.py:24-78
    me, mg = est.mean(1, keepdims=True), gt.mean(1, keepdims=True)
    U, _, Vt = np.linalg.svd(((est - me) @ (gt - mg).T).T)
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    err = R @ est + (mg - R @ me) - gt
    return float(np.sqrt((err ** 2).sum(0).mean()))


def run(shape="tum", frames=50, track_iters=40, map_iters=40, map_every=5, kf_every=5, window=8, seed=0,
        isotropic=True, quiet=True, ssim=False, isoflat=False):
    dev = torch.device("cuda", 0)
    W, H, fx, fy, cx, cy = syn.SHAPES[shape][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(frames, step=1.0)
    rng = np.random.RandomState(seed)
    c0, d0, K = syn.render_frame(shape, traj[0], seed=0)
    pts, cols, msd = syn.backproject(c0, d0, K, traj[0])
    params = syn.init_params(pts, cols, msd, isotropic=isotropic, seed=seed)     # initialize_first_timestep
    gmap = ss.GaussianMap(params, device=dev)
    ws = ss.StepWorkspace(cam, gmap.count, dev)
    scene_radius = float(d0.max()) / 3.0                                          # splatam.py:448 (ratio 3)
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ident = (np.array([1, 0, 0, 0], np.float32), np.zeros(3, np.float32))
    poses, gt_c = [ident], [np.zeros(3)]
    keyframes = [dict(rgb=tt(c0), depth=tt(d0[0]), pose=ident, inv_nd=1.0 / max(1, int((d0 > 0).sum())))]
    t_track = t_map = 0.0
    n_track = n_map = 0
    added_total = pruned_total = 0
    for f in range(1, frames):
        c, d, _ = syn.render_frame(shape, traj[f], seed=f)
        rgb, depth = tt(c), tt(d[0])
        gt_c.append(-traj[f][:3, :3].T.astype(np.float64) @ traj[f][:3, 3].astype(np.float64))
        # ---- tracking (splatam.py:1231-1323)
        q, tr = ss.init_camera_pose(poses[-1], poses[-2] if f > 1 else None)
        ps = ss.PoseState(q, tr, dev)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(track_iters):
            ss.track_step(ws, gmap, ps, rgb, depth)
        torch.cuda.synchronize(); t_track += time.perf_counter() - t0; n_track += track_iters
        poses.append(ps.best_pose())
        # ---- densification + mapping (splatam.py:1376-1491)
        if f % map_every == 0:
            added_total += ss.add_new_gaussians(ws, gmap, poses[-1], rgb, depth, seed=seed + f)
            gmap.reset_optimizer()
            cur = dict(rgb=rgb, depth=depth, pose=poses[-1], inv_nd=1.0 / max(1, int((d > 0).sum())))
            # keyframe_selection_overlap on keyframe_list[:-1] + last keyframe + current frame (splatam.py:1391-1401)
            sel = []
            if len(keyframes) > 1:
                ids = ss.select_keyframes(cam, depth, w2c_of(*poses[-1]), [w2c_of(*k["pose"]) for k in keyframes[:-1]],
                                          k=window - 2, seed=seed + f, rng=rng)
                sel = [keyframes[int(i)] for i in ids]
            sel = sel + [keyframes[-1], cur]
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for it in range(map_iters):
                if it in (0, 20):                                                  # configs/replica/splatam.py:119-128
                    pruned_total += ss.prune_gaussians(ws, gmap, 0.005, 0.1 * scene_radius if it >= 0 else float("inf"))
                k = sel[rng.randint(len(sel))]
                ss.map_step(ws, gmap, pkg.make_pose(*k["pose"]), k["rgb"], k["depth"], k["inv_nd"], ssim=ssim, isoflat=isoflat,
                            iso_seed=seed + 1000 * f + it)
            torch.cuda.synchronize(); t_map += time.perf_counter() - t0; n_map += map_iters
            if f % kf_every == 0:
                keyframes.append(cur)
    ws.check_overflow(wait=True)
    # quality of the last frame
    ss.render_forward(ws, gmap, pose=pkg.make_pose(*poses[-1]))
    mse = float(((ws.render[:3] - rgb) ** 2).mean())
    est_c = [centre(*p) for p in poses]
    return dict(shape=shape, W=W, H=H, frames=frames, track_iters=track_iters, map_iters=map_iters,
                gaussians=int(gmap.count), added=int(added_total), pruned=int(pruned_total),
                ate_rmse_mm=1e3 * ate_rmse(est_c, gt_c), psnr_last=float(-10 * np.log10(max(mse, 1e-12))),
                tracking_iters_per_s=n_track / max(t_track, 1e-9), mapping_iters_per_s=n_map / max(t_map, 1e-9),
                note="wall clock with synchronize() around each tracking / mapping block; includes the densify / "
                     "prune host reads; mapping colour loss = " +
                     ("0.8*L1 + 0.2*(1-SSIM)" if ssim else "L1") + ", masked depth L1" +
                     (" + 50*flat + 2*iso (three-axis scales)" if isoflat else ""))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="tum")
    ap.add_argument("--frames", type=int, default=50)
    ap.add_argument("--track-iters", type=int, default=40)
    ap.add_argument("--map-iters", type=int, default=40)
    ap.add_argument("--map-every", type=int, default=5)
    ap.add_argument("--ssim", action="store_true", help="mapping colour loss 0.8*L1 + 0.2*(1-SSIM) (splatam.py:568)")
    ap.add_argument("--isoflat", action="store_true",
                    help="three-axis scales + the IsoGS flat / iso-surface terms (splatam.py:572-675): the complete loss")
    a = ap.parse_args()
    print(json.dumps(run(a.shape, a.frames, a.track_iters, a.map_iters, a.map_every, ssim=a.ssim,
                         isotropic=not a.isoflat, isoflat=a.isoflat)))
