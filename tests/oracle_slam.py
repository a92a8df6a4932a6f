"""TEST INFRASTRUCTURE: the tracking / mapping iteration of python_ordinary/scripts/splatam.py
(1247-1290, 1412-1476; L1 terms) executed with the CPU oracle.  Used by the GPU parity tests as the
checker for slam_step.map_step / track_step and for the 50-frame ATE comparison."""
import numpy as np

MAP_LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=0.05, log_scales=1e-3)


def _quats(p):
    """Rotations matter only for three-axis scales (transform_to_frame, slam_helpers.py:276-279)."""
    return p["unnorm_rotations"] if p["log_scales"].ndim == 2 and p["log_scales"].shape[1] == 3 else None


def new_state(params):
    p = {k: np.ascontiguousarray(v, dtype=np.float32).copy() for k, v in params.items()}
    return dict(p=p, m={k: np.zeros_like(v) for k, v in p.items()}, v={k: np.zeros_like(v) for k, v in p.items()},
                step=0)


def reset_optimizer(state):
    for k in state["m"]:
        state["m"][k][...] = 0
        state["v"][k][...] = 0
    state["step"] = 0


def _flags(orc):
    return orc.F_LOG_SCALES | orc.F_LOGIT_OPAC


def map_iteration(orc, ocam, state, q, t, gt_rgb, gt_depth, grads_out=None, extra_grads=None, w_im=0.5, w_depth=1.0,
                  ssim=False, isoflat=False, ignore_outlier_depth=False, ba=None, lr_rot=4e-4, lr_trans=2e-3):
    """One mapping iteration on one keyframe.  extra_grads: list of gradient dicts from OTHER ranks of the
    data-parallel schedule (averaged with this one before Adam).  ba: a TrackState -- do_ba=True
    (splatam.py:505-510), the pose is a parameter as well (q, t ignored)."""
    p = state["p"]
    pose = orc.make_pose(q, t) if ba is None else orc.make_pose(ba.pose[:4], ba.pose[4:])
    lo = p["logit_opacities"].reshape(-1)
    qs = _quats(p)
    o = orc.render_full(ocam, p["means3D"], p["log_scales"], qs, lo, p["rgb_colors"], ch=5, flags=_flags(orc),
                        pose=pose)
    v_render, losses = orc.loss_l1(ocam, 1, o["render"], o["alpha"], gt_rgb, gt_depth,
                                   w_im=(0.8 * w_im if ssim else w_im), w_depth=w_depth,
                                   ignore_outlier_depth=ignore_outlier_depth)
    if ssim:  # colour term 0.8 * L1 + 0.2 * (1 - SSIM), splatam.py:568
        v_c = np.ascontiguousarray(v_render[:3])
        s_val = orc.ssim(o["render"][:3], gt_rgb, -0.2 * w_im, v_c)
        v_render[:3] = v_c
        losses[0] += 0.2 * w_im * (1.0 - s_val)
    v_rec = orc.blend_bwd(ocam, 5, o["rec"], o["vals"], o["ranges"], o["alpha"], o["last"], v_render)
    g = orc.preprocess_bwd(ocam, p["means3D"], p["log_scales"], qs, lo, p["rgb_colors"], o["radii"], o["rec"], v_rec,
                           ch=5, flags=_flags(orc), pose=pose)
    grads = dict(means3D=g["means"], rgb_colors=g["colors"], logit_opacities=g["opacities"], log_scales=g["scales"])
    if qs is not None:
        grads["unnorm_rotations"] = g["quats"]
    if isoflat:  # 50 * flat + 2 * iso on every centre (N < 8192) -- splatam.py:572-675,733-739
        total, _, gg, _ = orc.iso_flat_loss(p["means3D"], p["log_scales"], p["unnorm_rotations"], lo)
        losses[0] += total
        for k in gg:
            grads[k] = (np.asarray(grads[k]).reshape(gg[k].shape) + gg[k]).astype(np.float32)
    if grads_out is not None:
        grads_out.update({k: np.array(v, copy=True) for k, v in grads.items()})
    if extra_grads:
        n = 1 + len(extra_grads)
        grads = {k: (sum([v.reshape(-1)] + [e[k].reshape(-1) for e in extra_grads]) / n).astype(np.float32)
                 for k, v in grads.items()}
    state["step"] += 1
    for k, lr in MAP_LRS.items():
        if k not in grads:
            continue
        orc.adam(p[k].reshape(-1), np.ascontiguousarray(grads[k]).reshape(-1), state["m"][k].reshape(-1),
                 state["v"][k].reshape(-1), lr, eps=1e-15, step=state["step"])
    if ba is not None:
        gp = np.ascontiguousarray(g["pose"], dtype=np.float32)
        if grads_out is not None:
            grads_out["pose"] = gp.copy()
        ba.step += 1
        orc.adam(ba.pose[:4], gp[:4], ba.m[:4], ba.v[:4], lr_rot, eps=1e-15, step=ba.step)
        orc.adam(ba.pose[4:], gp[4:], ba.m[4:], ba.v[4:], lr_trans, eps=1e-15, step=ba.step)
    return float(losses[0])


def gradients_only(orc, ocam, params, q, t, gt_rgb, gt_depth):
    g = {}
    st = new_state(params)
    map_iteration(orc, ocam, st, q, t, gt_rgb, gt_depth, grads_out=g)
    return g


class TrackState:
    def __init__(self, q, t):
        self.pose = np.array(list(q) + list(t), dtype=np.float32)
        self.m = np.zeros(7, np.float32)
        self.v = np.zeros(7, np.float32)
        self.step = 0
        self.best = self.pose.copy()
        self.best_loss = 1e20


def track_iteration(orc, ocam, params, ts, gt_rgb, gt_depth, lr_rot=4e-4, lr_trans=2e-3, use_sil=True,
                    ignore_outlier_depth=False):
    """One tracking iteration: sum-L1 with the silhouette mask, pose gradient, Adam on the 7 pose floats,
    best candidate kept as in splatam.py:1287-1290."""
    pose = orc.make_pose(ts.pose[:4], ts.pose[4:])
    lo = params["logit_opacities"].reshape(-1)
    qs = _quats(params)
    o = orc.render_full(ocam, params["means3D"], params["log_scales"], qs, lo, params["rgb_colors"], ch=5,
                        flags=_flags(orc), pose=pose)
    v_render, losses = orc.loss_l1(ocam, 0, o["render"], o["alpha"], gt_rgb, gt_depth, use_sil=use_sil,
                                   ignore_outlier_depth=ignore_outlier_depth)
    v_rec = orc.blend_bwd(ocam, 5, o["rec"], o["vals"], o["ranges"], o["alpha"], o["last"], v_render)
    g = orc.preprocess_bwd(ocam, params["means3D"], params["log_scales"], qs, lo, params["rgb_colors"], o["radii"],
                           o["rec"], v_rec, ch=5, flags=_flags(orc), pose=pose)
    gp = np.ascontiguousarray(g["pose"], dtype=np.float32)
    ts.step += 1
    orc.adam(ts.pose[:4], gp[:4], ts.m[:4], ts.v[:4], lr_rot, eps=1e-8, step=ts.step)
    orc.adam(ts.pose[4:], gp[4:], ts.m[4:], ts.v[4:], lr_trans, eps=1e-8, step=ts.step)
    loss = float(np.float32(losses[0]))
    if loss < ts.best_loss:
        ts.best_loss = loss
        ts.best = ts.pose.copy()
    return loss, gp


def ate_rmse(est_t, gt_t):
    """Horn-aligned translational RMSE (python_ordinary/utils/eval_helpers.py:24-78: align + evaluate_ate).
    est_t, gt_t: [n,3] camera centres."""
    est, gt = np.asarray(est_t, np.float64).T, np.asarray(gt_t, np.float64).T
    mu_e, mu_g = est.mean(1, keepdims=True), gt.mean(1, keepdims=True)
    Wm = (est - mu_e) @ (gt - mu_g).T
    U, _, Vt = np.linalg.svd(Wm.T)
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    tr = mu_g - R @ mu_e
    err = R @ est + tr - gt
    return float(np.sqrt((err ** 2).sum(0).mean()))


def pose_to_centre(q, t):
    """camera centre of the w2c pose (q, t): c = -R^T t."""
    q = np.asarray(q, np.float64)
    q = q / np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return -R.T @ np.asarray(t, np.float64)


def keyframe_overlap(depth, intr, cur_w2c, kf_w2cs, sample_idx, edge=20):
    """percent_inside of keyframe_selection_overlap (python_ordinary/utils/keyframe_selection.py:9-84) in numpy
    float32: back-projection of the sampled pixels, the unique/isin duplicate filter, projection into each
    keyframe.  depth [H,W]; intr = (W, H, fx, fy, cx, cy); sample_idx flat pixel indices."""
    W, H, fx, fy, cx, cy = intr
    f32 = np.float32
    ys, xs = (sample_idx // int(W)).astype(np.int64), (sample_idx % int(W)).astype(np.int64)
    z = depth[ys, xs].astype(f32)
    pc = np.stack([(xs.astype(f32) - f32(cx)) / f32(fx) * z, (ys.astype(f32) - f32(cy)) / f32(fy) * z, z], -1)
    c2w = np.linalg.inv(np.asarray(cur_w2c, np.float64)).astype(f32)
    pts = (pc @ c2w[:3, :3].T + c2w[:3, 3]).astype(f32)
    key = np.abs(np.rint(pts * f32(1e4)) / f32(1e4))
    _, inv, counts = np.unique(np.concatenate([key, np.zeros((1, 3), f32)]), axis=0, return_inverse=True,
                               return_counts=True)
    valid = counts[inv.reshape(-1)[:len(key)]] == 1
    pts = pts[valid]
    out = []
    for M in np.asarray(kf_w2cs, f32):
        p = pts @ M[:3, :3].T + M[:3, 3]
        pz = p[:, 2] + f32(1e-5)
        u, v = (f32(fx) * p[:, 0] + f32(cx) * p[:, 2]) / pz, (f32(fy) * p[:, 1] + f32(cy) * p[:, 2]) / pz
        m = (u < W - edge) & (u > edge) & (v < H - edge) & (v > edge) & (pz > 0)
        out.append(m.sum() / max(1, len(pts)))
    return np.array(out, f32)
