"""TEST INFRASTRUCTURE: the frame loop of rgbd_slam (python_ordinary/scripts/splatam.py:1200-1536) -- constant-velocity
pose initialisation, N tracking iterations with the best-candidate rule, silhouette densification at the tracked pose,
overlap keyframe selection, M mapping iterations on randomly drawn keyframes with pruning at iterations 0 and 20,
keyframe insertion -- written ONCE against a small backend interface, so that the CPU oracle
(oracle/pin_ate_tum.py, which records the schedule's random draws) and the fused device steps
(tests/test_gpu_ate.py, which replays them) execute the identical schedule.  L1 terms, isotropic scales
(BASELINE.json configs[1])."""
import numpy as np

import oracle_slam as osl

CONFIG = dict(frames=50, track_iters=40, map_iters=30, map_every=5, kf_every=5, window=8, sil_add=0.5,
              prune_opac=0.005, prune_at=(0, 20), overlap_pixels=1600, seed=0)


def w2c_of(q, t):
    q = np.asarray(q, np.float64) / np.linalg.norm(q)
    w, x, y, z = q
    M = np.eye(4, dtype=np.float32)
    M[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                 [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                 [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    M[:3, 3] = np.asarray(t, np.float32)
    return M


def init_camera_pose(prev1, prev2=None):
    """initialize_camera_pose (splatam.py:844-863), float32 like slam_step.init_camera_pose."""
    q1, t1 = (np.asarray(a, dtype=np.float32) for a in prev1)
    if prev2 is None:
        return q1.copy(), t1.copy()
    q2, t2 = (np.asarray(a, dtype=np.float32) for a in prev2)
    n1 = q1 / max(float(np.linalg.norm(q1)), 1e-12)
    n2 = q2 / max(float(np.linalg.norm(q2)), 1e-12)
    q = n1 + (n1 - n2)
    q = q / max(float(np.linalg.norm(q)), 1e-12)
    return q.astype(np.float32), (t1 + (t1 - t2)).astype(np.float32)


def frames_of(syn, shape, n):
    """The synthetic sequence: (rgb [3,H,W], depth [H,W]) per frame, ground-truth camera centres."""
    traj = syn.trajectory(n, step=1.0)
    out, gt = [], []
    for f in range(n):
        c, d, _ = syn.render_frame(shape, traj[f], seed=f)
        out.append((np.ascontiguousarray(c), np.ascontiguousarray(d[0])))
        gt.append(-traj[f][:3, :3].T.astype(np.float64) @ traj[f][:3, 3].astype(np.float64))
    return out, np.array(gt)


def run(backend, frames, intr, cfg=CONFIG, schedule=None, log=None):
    """backend: init(params-from-frame-0 is done by the caller) and
         track(frame_idx, rgb, depth, q0, t0, iters) -> (q, t) best candidate
         add_new(frame_idx, rgb, depth, q, t) -> number added
         prune(max_scale) -> number removed
         map_iter(kf) with kf = dict(idx, rgb, depth, q, t, inv_nd)
         count() -> Gaussians
    schedule: None = draw (and record) the random choices; else the dict recorded by an earlier run.
    Returns dict(poses [n,7], schedule, counts)."""
    W, H, fx, fy, cx, cy = intr
    rng = np.random.RandomState(cfg["seed"])
    rec = dict(sel={}, picks={}, samples={}) if schedule is None else schedule
    ident = (np.array([1, 0, 0, 0], np.float32), np.zeros(3, np.float32))
    poses = [ident]
    d0 = frames[0][1]
    scene_radius = float(d0.max()) / 3.0                                   # splatam.py:448
    kfs = [dict(idx=0, rgb=frames[0][0], depth=d0, q=ident[0], t=ident[1], inv_nd=1.0 / max(1, int((d0 > 0).sum())))]
    counts = []
    for f in range(1, cfg["frames"]):
        rgb, depth = frames[f]
        q0, t0 = init_camera_pose(poses[-1], poses[-2] if f > 1 else None)
        q, t = backend.track(f, rgb, depth, q0, t0, cfg["track_iters"])
        poses.append((np.asarray(q, np.float32), np.asarray(t, np.float32)))
        if f % cfg["map_every"] == 0:
            added = backend.add_new(f, rgb, depth, poses[-1][0], poses[-1][1])
            cur = dict(idx=f, rgb=rgb, depth=depth, q=poses[-1][0], t=poses[-1][1],
                       inv_nd=1.0 / max(1, int((depth > 0).sum())))
            # keyframe_selection_overlap on keyframe_list[:-1], + the last keyframe + the current frame (:1391-1401)
            if schedule is None:
                sel = []
                if len(kfs) > 1:
                    valid = np.nonzero(depth.reshape(-1) > 0)[0]
                    samp = valid[rng.randint(0, len(valid), cfg["overlap_pixels"])]
                    pct = osl.keyframe_overlap(depth, (W, H, fx, fy, cx, cy), w2c_of(*poses[-1]),
                                               [w2c_of(k["q"], k["t"]) for k in kfs[:-1]], samp)
                    order = sorted(range(len(pct)), key=lambda i: pct[i], reverse=True)
                    keep = [i for i in order if pct[i] > 0.0]
                    sel = [int(i) for i in rng.permutation(np.array(keep, dtype=np.int64))[:cfg["window"] - 2]]
                picks = [int(rng.randint(len(sel) + 2)) for _ in range(cfg["map_iters"])]
                rec["sel"][f], rec["picks"][f] = sel, picks
            else:
                sel, picks = [int(i) for i in rec["sel"][f]], [int(i) for i in rec["picks"][f]]
            cand = [kfs[i] for i in sel] + [kfs[-1], cur]
            backend.begin_mapping()
            removed = 0
            for it in range(cfg["map_iters"]):
                if it in cfg["prune_at"]:
                    removed += backend.prune(cfg["prune_opac"], 0.1 * scene_radius)
                backend.map_iter(cand[picks[it]])
            if f % cfg["kf_every"] == 0:
                kfs.append(cur)
            counts.append((f, int(added), int(removed), int(backend.count())))
            if log:
                log(f"frame {f}: +{added} -{removed} -> {backend.count()} Gaussians, {len(kfs)} keyframes")
    return dict(poses=np.array([np.concatenate([q, t]) for q, t in poses], np.float32), schedule=rec, counts=counts)


def ate_of(poses, gt_centres):
    c = np.array([osl.pose_to_centre(p[:4], p[4:]) for p in poses])
    return osl.ate_rmse(c, gt_centres), c


class OracleBackend:
    """The loop on the CPU oracle (numpy state)."""

    def __init__(self, orc, ocam, params, intr):
        self.orc, self.ocam, self.intr = orc, ocam, intr
        self.state = osl.new_state(params)

    def count(self):
        return self.state["p"]["means3D"].shape[0]

    def track(self, f, rgb, depth, q0, t0, iters):
        ts = osl.TrackState(q0, t0)
        for _ in range(iters):
            osl.track_iteration(self.orc, self.ocam, self.state["p"], ts, rgb, depth)
        return ts.best[:4].copy(), ts.best[4:].copy()

    def add_new(self, f, rgb, depth, q, t):
        """add_new_gaussians (splatam.py:799-841) + get_pointcloud (:298-348) + initialize_new_params (:763-796)."""
        orc, p = self.orc, self.state["p"]
        W, H, fx, fy, cx, cy = self.intr
        flags = orc.F_LOG_SCALES | orc.F_LOGIT_OPAC
        o = orc.render_full(self.ocam, p["means3D"], p["log_scales"], None, p["logit_opacities"].reshape(-1),
                            p["rgb_colors"], ch=5, flags=flags, pose=orc.make_pose(q, t))
        rd, sil = o["render"][3], o["alpha"]
        thr = orc.depth_outlier_threshold(rd, depth, 50.0)
        err = np.abs(depth - rd) * (depth > 0)
        mask = ((sil < 0.5) | ((rd > depth) & (err > thr))) & (depth > 0)
        ys, xs = np.nonzero(mask)
        if len(ys) == 0:
            return 0
        z = depth[ys, xs].astype(np.float32)
        pc = np.stack([(xs.astype(np.float32) - np.float32(cx)) / np.float32(fx) * z,
                       (ys.astype(np.float32) - np.float32(cy)) / np.float32(fy) * z, z], -1)
        c2w = np.linalg.inv(w2c_of(q, t).astype(np.float64))
        pw = (pc.astype(np.float64) @ c2w[:3, :3].T + c2w[:3, 3]).astype(np.float32)
        n = len(z)
        new = dict(means3D=pw, rgb_colors=np.ascontiguousarray(rgb[:, ys, xs].T, dtype=np.float32),
                   unnorm_rotations=np.tile(np.array([[1, 0, 0, 0]], np.float32), (n, 1)),
                   logit_opacities=np.zeros((n, 1), np.float32),
                   log_scales=np.log(z / np.float32((fx + fy) / 2.0)).astype(np.float32)[:, None])
        for k in list(p.keys()):
            self.state["p"][k] = np.ascontiguousarray(np.concatenate([self.state["p"][k], new[k].reshape(n, -1)]))
            self.state["m"][k] = np.ascontiguousarray(np.concatenate([self.state["m"][k], np.zeros((n,) + self.state["m"][k].shape[1:], np.float32)]))
            self.state["v"][k] = np.ascontiguousarray(np.concatenate([self.state["v"][k], np.zeros((n,) + self.state["v"][k].shape[1:], np.float32)]))
        return n

    def begin_mapping(self):
        osl.reset_optimizer(self.state)                                    # splatam.py:1406

    def prune(self, opac_thres, max_scale):
        """prune_gaussians + remove_points (slam_external.py:139-188)."""
        p = self.state["p"]
        sig = 1.0 / (1.0 + np.exp(-p["logit_opacities"].reshape(-1).astype(np.float64)))
        big = np.exp(p["log_scales"].astype(np.float64)).max(axis=1) > max_scale
        keep = ~((sig < opac_thres) | big)
        removed = int((~keep).sum())
        if removed:
            for d in (self.state["p"], self.state["m"], self.state["v"]):
                for k in d:
                    d[k] = np.ascontiguousarray(d[k][keep])
        return removed

    def map_iter(self, kf):
        osl.map_iteration(self.orc, self.ocam, self.state, kf["q"], kf["t"], kf["rgb"], kf["depth"])
