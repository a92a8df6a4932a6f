"""Deterministic synthetic RGB-D sequences shaped like the reference's datasets:
an axis-aligned box room with procedural texture and a few boxes, analytic ray-cast depth, a smooth
Lissajous trajectory with pose 0 = identity (relative_pose=True, splatam.py:959,1213-1214), and
Gaussians initialised exactly as the reference does (one per valid-depth pixel, s = z/((fx+fy)/2),
logit_o = 0, q = [1,0,0,0]; splatam.py:298-399).  numpy only; no dataset files."""
import numpy as np

SHAPES = {
    # name: (W, H, fx, fy, cx, cy, depth_quantum, hole_fraction)
    "tum": (640, 480, 525.0, 525.0, 319.5, 239.5, 1.0 / 5000.0, 0.05),
    "replica": (1200, 680, 600.0, 600.0, 599.5, 339.5, 1.0 / 6553.5, 0.0),
    "scannetpp": (1752, 1168, 1200.0, 1200.0, 875.5, 583.5, 1.0 / 1000.0, 0.0),
    "tiny": (160, 120, 130.0, 130.0, 79.5, 59.5, 0.0, 0.02),
}

ROOM_MIN = np.array([-3.0, -1.5, -2.0])
ROOM_MAX = np.array([3.0, 1.5, 4.0])


def _boxes(seed=1234):
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(8):
        c = np.array([rng.uniform(-2.2, 2.2), rng.uniform(0.2, 1.2), rng.uniform(1.5, 3.5)])
        h = np.array([rng.uniform(0.15, 0.5), rng.uniform(0.15, 0.4), rng.uniform(0.15, 0.5)])
        out.append((c - h, c + h))
    return out


def _texture(p, n_id):
    """Procedural albedo in [0,1]^3 from hit point and a surface id."""
    f = 3.0 + n_id
    r = 0.5 + 0.25 * np.sin(f * p[..., 0] + 1.3 * n_id) + 0.25 * np.sin(2.1 * f * p[..., 2] + 0.4)
    g = 0.5 + 0.25 * np.sin(f * p[..., 1] + 2.1 * n_id) + 0.25 * np.cos(1.7 * f * p[..., 0] + 0.9)
    b = 0.5 + 0.25 * np.cos(f * p[..., 2] + 0.7 * n_id) + 0.25 * np.sin(1.3 * f * p[..., 1] + 1.9)
    return np.clip(np.stack([r, g, b], -1), 0.0, 1.0)


def trajectory(n_frames, step=1.0):
    """w2c 4x4 per frame; <= ~2 cm and <= ~0.5 deg per frame; frame 0 = identity."""
    out = []
    for k in range(n_frames):
        a = k * step
        t = np.array([0.30 * np.sin(0.05 * a), 0.10 * np.sin(0.07 * a), 0.25 * (1 - np.cos(0.04 * a))])
        yaw, pitch = 0.12 * np.sin(0.045 * a), 0.05 * np.sin(0.06 * a)
        cy_, sy_ = np.cos(yaw), np.sin(yaw)
        cp_, sp_ = np.cos(pitch), np.sin(pitch)
        Ry = np.array([[cy_, 0, sy_], [0, 1, 0], [-sy_, 0, cy_]])
        Rx = np.array([[1, 0, 0], [0, cp_, -sp_], [0, sp_, cp_]])
        c2w = np.eye(4)
        c2w[:3, :3] = Ry @ Rx
        c2w[:3, 3] = t
        out.append(np.linalg.inv(c2w))
    return out


def render_frame(shape, w2c=None, seed=0):
    """Returns color [3,H,W] float32 in [0,1], depth [1,H,W] float32 metres (0 = hole), K [3,3]."""
    W, H, fx, fy, cx, cy, quantum, holes = SHAPES[shape] if isinstance(shape, str) else shape
    w2c = np.eye(4) if w2c is None else np.asarray(w2c, dtype=np.float64)
    c2w = np.linalg.inv(w2c)
    u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64), indexing="xy")
    d_cam = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], -1)
    d = d_cam @ c2w[:3, :3].T
    o = c2w[:3, 3]
    best_t = np.full((H, W), np.inf)
    best_id = np.zeros((H, W), dtype=np.int32)
    # room (inside-out): exit distance of the slab intersection
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (ROOM_MIN - o) / d
        t2 = (ROOM_MAX - o) / d
    tfar = np.maximum(t1, t2)
    t_exit = tfar.min(-1)
    axis = tfar.argmin(-1)
    best_t = t_exit
    best_id = axis.astype(np.int32) * 2 + (np.take_along_axis(d, axis[..., None], -1)[..., 0] > 0)
    for bi, (bmin, bmax) in enumerate(_boxes()):
        with np.errstate(divide="ignore", invalid="ignore"):
            a = (bmin - o) / d
            b = (bmax - o) / d
        tn = np.minimum(a, b).max(-1)
        tf = np.maximum(a, b).min(-1)
        hit = (tn < tf) & (tn > 1e-3) & (tn < best_t)
        best_t = np.where(hit, tn, best_t)
        best_id = np.where(hit, 6 + bi, best_id)
    p = o + d * best_t[..., None]
    color = _texture(p, best_id.astype(np.float64))
    depth = best_t  # d_cam.z == 1  =>  t is z-depth in the camera frame
    if quantum > 0:
        depth = np.round(depth / quantum) * quantum
    if holes > 0:
        rng = np.random.RandomState(seed + 17)
        # blocky holes, like real sensor dropouts
        hb = rng.rand((H + 7) // 8, (W + 7) // 8) < holes
        depth = np.where(np.kron(hb, np.ones((8, 8), bool))[:H, :W], 0.0, depth)
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float32)
    return (np.ascontiguousarray(color.transpose(2, 0, 1), dtype=np.float32),
            depth[None].astype(np.float32), K)


def backproject(color, depth, K, w2c=None, mask=None):
    """get_pointcloud (splatam.py:298-348): integer pixel coordinates, returns pts [n,3], cols [n,3],
    mean3_sq_dist [n]."""
    H, W = depth.shape[-2:]
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    u, v = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    z = depth.reshape(-1).astype(np.float32)
    pts = np.stack([((u.reshape(-1) - cx) / fx) * z, ((v.reshape(-1) - cy) / fy) * z, z], -1)
    if w2c is not None:
        c2w = np.linalg.inv(np.asarray(w2c, dtype=np.float64))
        pts = (pts.astype(np.float64) @ c2w[:3, :3].T + c2w[:3, 3]).astype(np.float32)
    cols = color.transpose(1, 2, 0).reshape(-1, 3)
    s = z / ((fx + fy) / 2.0)
    sel = (z > 0) if mask is None else (mask.reshape(-1) & (z > 0))
    return pts[sel], cols[sel].astype(np.float32), (s * s)[sel]


def init_params(pts, cols, mean3_sq_dist, isotropic=True, seed=0):
    """initialize_params (splatam.py:351-399).  isotropic=True -> log_scales [N,1] (the BASELINE
    benchmark configuration); False -> the IsoGS [N,3] + randn*0.01 variant (splatam.py:356-368)."""
    n = pts.shape[0]
    base = np.log(np.sqrt(mean3_sq_dist))[:, None].astype(np.float32)
    if isotropic:
        log_scales = base
    else:
        rng = np.random.RandomState(seed)
        log_scales = np.tile(base, (1, 3)) + (rng.randn(n, 3) * 0.01).astype(np.float32)
    rots = np.zeros((n, 4), np.float32)
    rots[:, 0] = 1.0
    return dict(means3D=np.ascontiguousarray(pts, dtype=np.float32), rgb_colors=np.ascontiguousarray(cols),
                unnorm_rotations=rots, logit_opacities=np.zeros((n, 1), np.float32),
                log_scales=np.ascontiguousarray(log_scales, dtype=np.float32))


def make_map(shape, n_target=None, isotropic=True, seed=0, max_frames=64):
    """Gaussians from frame 0, grown toward n_target by appending back-projected pixels of later
    frames (same rule as add_new_gaussians, a18) -- returns params dict, list of (color, depth, w2c), K."""
    traj = trajectory(max_frames, step=4.0)
    frames = []
    color, depth, K = render_frame(shape, traj[0], seed)
    frames.append((color, depth, traj[0]))
    pts, cols, msd = backproject(color, depth, K, traj[0])
    k = 1
    rng = np.random.RandomState(seed + 1)
    while n_target is not None and pts.shape[0] < n_target and k < max_frames:
        color, depth, _ = render_frame(shape, traj[k], seed + k)
        frames.append((color, depth, traj[k]))
        p2, c2, m2 = backproject(color, depth, K, traj[k])
        need = n_target - pts.shape[0]
        if p2.shape[0] > need:
            sel = np.sort(rng.choice(p2.shape[0], need, replace=False))
            p2, c2, m2 = p2[sel], c2[sel], m2[sel]
        pts, cols, msd = np.concatenate([pts, p2]), np.concatenate([cols, c2]), np.concatenate([msd, m2])
        k += 1
    if n_target is not None and pts.shape[0] > n_target:
        pts, cols, msd = pts[:n_target], cols[:n_target], msd[:n_target]
    return init_params(pts, cols, msd, isotropic, seed), frames, K


def w2c_to_pose(w2c):
    """4x4 w2c -> (unit quaternion wxyz, translation) as used by cam_unnorm_rots / cam_trans."""
    R = np.asarray(w2c, dtype=np.float64)[:3, :3]
    t = np.asarray(w2c, dtype=np.float64)[:3, 3]
    tr = np.trace(R)
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k]) * 2
        q = np.zeros(4)
        q[0] = (R[k, j] - R[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (R[j, i] + R[i, j]) / s
        q[1 + k] = (R[k, i] + R[i, k]) / s
    if q[0] < 0:
        q = -q
    return q.astype(np.float32), t.astype(np.float32)
