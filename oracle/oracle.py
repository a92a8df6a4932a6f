"""ctypes/numpy front-end of the CPU ORACLE (test infrastructure, not product code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  See isogs_oracle.h for the reference file:line map.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libisogs_oracle.so")
REC = 12

F_LOG_SCALES, F_LOGIT_OPAC, F_NO_OPAC_EXTENT = 1, 2, 4


class Camera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("R", C.c_float * 9), ("t", C.c_float * 3),
                ("near_plane", C.c_float), ("far_plane", C.c_float), ("eps2d", C.c_float),
                ("bg", C.c_float * 3)]


class Pose(C.Structure):
    _fields_ = [("q", C.c_float * 4), ("t", C.c_float * 3)]


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("isogs_oracle.cpp", "isogs_oracle.h")]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in src)):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_logf.restype = C.c_float
        _lib.orc_logf.argtypes = [C.c_float]
        _lib.orc_expf.restype = C.c_float
        _lib.orc_expf.argtypes = [C.c_float]
        _lib.orc_tile_bin.restype = C.c_int64
        _lib.orc_compact_rows.restype = C.c_int64
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def make_camera(width, height, fx, fy, cx, cy, viewmat=None, near=0.01, far=100.0, eps2d=0.3, bg=(0, 0, 0)):
    """viewmat: 4x4 (or 3x4) row-major world->camera."""
    cam = Camera()
    cam.width, cam.height = int(width), int(height)
    cam.fx, cam.fy, cam.cx, cam.cy = float(fx), float(fy), float(cx), float(cy)
    vm = np.eye(4, dtype=np.float32) if viewmat is None else np.asarray(viewmat, dtype=np.float32)
    cam.R[:] = [float(v) for v in vm[:3, :3].reshape(-1)]
    cam.t[:] = [float(v) for v in vm[:3, 3]]
    cam.near_plane, cam.far_plane, cam.eps2d = float(near), float(far), float(eps2d)
    cam.bg[:] = [float(b) for b in bg]
    return cam


def make_pose(q, t):
    p = Pose()
    p.q[:] = [float(v) for v in q]
    p.t[:] = [float(v) for v in t]
    return p


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))


def n_tiles(cam):
    return ((cam.width + 15) // 16) * ((cam.height + 15) // 16)


def logf(x):
    return float(lib().orc_logf(float(x)))


def expf(x):
    return float(lib().orc_expf(float(x)))


def preprocess_fwd(cam, means, scales, quats, opacities, colors, ch=4, flags=0, pose=None):
    means, scales, quats, opacities, colors = _f(means), _f(scales), _f(quats), _f(opacities), _f(colors)
    N = means.shape[0]
    scale_dim = 1 if scales.ndim == 1 or scales.shape[1] == 1 else 3
    rec = np.empty((N, REC), np.float32)
    radii = np.empty((N, 2), np.int32)
    tiles = np.empty((N,), np.int32)
    lib().orc_preprocess_fwd(C.byref(cam), C.byref(pose) if pose is not None else None, C.c_int64(N),
                             _p(means), _p(scales), C.c_int(scale_dim), _p(quats), _p(opacities), _p(colors),
                             C.c_int(ch), C.c_int(flags), _p(rec), _p(radii), _p(tiles))
    return rec, radii, tiles


def tile_bin(cam, rec, radii, tiles, want_unsorted=False):
    N = rec.shape[0]
    M = int(tiles.astype(np.int64).sum())
    ks = np.empty((M,), np.uint64)
    vs = np.empty((M,), np.int32)
    ku = np.empty((M,), np.uint64) if want_unsorted else None
    vu = np.empty((M,), np.int32) if want_unsorted else None
    ranges = np.empty((n_tiles(cam) + 1,), np.int32)
    m2 = lib().orc_tile_bin(C.byref(cam), C.c_int64(N), _p(rec), _p(radii), _p(tiles), C.c_int64(M),
                            _p(ku), _p(vu), _p(ks), _p(vs), _p(ranges))
    assert m2 == M
    if want_unsorted:
        return ks, vs, ranges, ku, vu
    return ks, vs, ranges


def blend_fwd(cam, ch, rec, vals, ranges):
    H, W = cam.height, cam.width
    render = np.empty((ch, H, W), np.float32)
    alpha = np.empty((H, W), np.float32)
    last = np.empty((H, W), np.int32)
    lib().orc_blend_fwd(C.byref(cam), C.c_int(ch), _p(rec), _p(vals), _p(ranges), _p(render), _p(alpha), _p(last))
    return render, alpha, last


def blend_bwd(cam, ch, rec, vals, ranges, alpha, last, v_render, v_alpha=None):
    N = rec.shape[0]
    v_render, v_alpha = _f(v_render), _f(v_alpha)
    v_rec = np.empty((N, REC), np.float32)
    lib().orc_blend_bwd(C.byref(cam), C.c_int(ch), C.c_int64(N), _p(rec), _p(vals), _p(ranges), _p(alpha),
                        _p(last), _p(v_render), _p(v_alpha), _p(v_rec))
    return v_rec


def preprocess_bwd(cam, means, scales, quats, opacities, colors, radii, rec, v_rec, ch=4, flags=0, pose=None):
    means, scales, quats, opacities, colors = _f(means), _f(scales), _f(quats), _f(opacities), _f(colors)
    N = means.shape[0]
    scale_dim = 1 if scales.ndim == 1 or scales.shape[1] == 1 else 3
    v_means = np.empty((N, 3), np.float32)
    v_scales = np.empty((N, scale_dim), np.float32)
    v_quats = np.empty((N, 4), np.float32) if quats is not None else None
    v_opac = np.empty((N,), np.float32)
    v_colors = np.empty((N, 3), np.float32)
    v_pose = np.zeros((7,), np.float64)
    lib().orc_preprocess_bwd(C.byref(cam), C.byref(pose) if pose is not None else None, C.c_int64(N),
                             _p(means), _p(scales), C.c_int(scale_dim), _p(quats), _p(opacities), _p(colors),
                             C.c_int(ch), C.c_int(flags), _p(radii), _p(rec), _p(_f(v_rec)),
                             _p(v_means), _p(v_scales), _p(v_quats), _p(v_opac), _p(v_colors), _p(v_pose))
    return dict(means=v_means, scales=v_scales, quats=v_quats, opacities=v_opac, colors=v_colors, pose=v_pose)


def depth_outlier_threshold(render_depth, gt_depth, factor=10.0):
    rd, gd = _f(render_depth), _f(gt_depth)
    fn = lib().orc_depth_outlier_threshold
    fn.restype = C.c_float
    return float(fn(C.c_int64(rd.size), _p(rd), _p(gd), C.c_float(factor)))


def loss_l1(cam, mode, render, alpha, gt_rgb, gt_depth, w_im=0.5, w_depth=1.0, sil_thres=0.99, use_sil=True,
            ignore_outlier_depth=False):
    render, alpha, gt_rgb, gt_depth = _f(render), _f(alpha), _f(gt_rgb), _f(gt_depth)
    v_render = np.empty_like(render)
    losses = np.zeros((3,), np.float64)
    thr = depth_outlier_threshold(render[3], gt_depth) if ignore_outlier_depth else -1.0
    lib().orc_loss_l1(C.byref(cam), C.c_int(mode), _p(render), _p(alpha), _p(gt_rgb), _p(gt_depth),
                      C.c_float(w_im), C.c_float(w_depth), C.c_float(sil_thres), C.c_int(int(use_sil)),
                      C.c_float(thr), _p(v_render), _p(losses))
    return v_render, losses


def ssim(x, y, coef=0.0, v_x=None):
    """Mean SSIM of [3,H,W] images; if v_x is given, v_x += coef * d(mean SSIM)/dx in place."""
    x, y = _f(x), _f(y)
    H, W = x.shape[1], x.shape[2]
    fn = lib().orc_ssim
    fn.restype = C.c_double
    return float(fn(C.c_int(W), C.c_int(H), _p(x), _p(y), C.c_float(coef), _p(v_x)))


def iso_flat_loss(means, log_scales, quats, logit_opac, query_idx=None, w_iso=2.0, w_flat=50.0, target=1.0,
                  with_grad=True):
    """Returns (weighted loss, {flat, iso, mean density}, gradient dict or None, neighbour table [Q,16])."""
    means, log_scales, quats, lo = _f(means), _f(log_scales), _f(quats), _f(logit_opac).reshape(-1)
    N = means.shape[0]
    qi = None if query_idx is None else np.ascontiguousarray(query_idx, dtype=np.int32)
    Q = N if qi is None else qi.shape[0]
    g = None
    if with_grad:
        g = dict(means3D=np.zeros((N, 3), np.float32), log_scales=np.zeros((N, 3), np.float32),
                 unnorm_rotations=np.zeros((N, 4), np.float32), logit_opacities=np.zeros((N,), np.float32))
    out = np.zeros((3,), np.float64)
    nbr = np.empty((Q, 16), np.int32)
    fn = lib().orc_iso_flat_loss
    fn.restype = C.c_double
    gp = (lambda k: _p(g[k])) if with_grad else (lambda k: None)
    total = fn(C.c_int64(N), _p(means), _p(log_scales), _p(quats), _p(lo), C.c_int(Q), _p(qi), C.c_float(w_iso),
               C.c_float(w_flat), C.c_float(target), gp("means3D"), gp("log_scales"), gp("unnorm_rotations"),
               gp("logit_opacities"), _p(out), _p(nbr))
    return float(total), out, g, nbr


def adam(p, g, m, v, lr, b1=0.9, b2=0.999, eps=1e-8, step=1):
    """In-place on p, m, v (float32 contiguous)."""
    lib().orc_adam(C.c_int64(p.size), _p(p), _p(_f(g)), _p(m), _p(v), C.c_float(lr), C.c_float(b1),
                   C.c_float(b2), C.c_float(eps), C.c_int(step))


def compact_rows(keep, rows):
    rows = _f(rows).reshape(rows.shape[0], -1)
    keep = np.ascontiguousarray(keep, dtype=np.uint8)
    out = np.empty_like(rows)
    n = lib().orc_compact_rows(C.c_int64(rows.shape[0]), C.c_int(rows.shape[1]), _p(keep), _p(rows), _p(out))
    return out[:n]


def render_full(cam, means, scales, quats, opacities, colors, ch=4, flags=0, pose=None):
    """Forward pipeline in one call; returns a dict of every intermediate."""
    rec, radii, tiles = preprocess_fwd(cam, means, scales, quats, opacities, colors, ch, flags, pose)
    keys, vals, ranges = tile_bin(cam, rec, radii, tiles)
    render, alpha, last = blend_fwd(cam, ch, rec, vals, ranges)
    return dict(rec=rec, radii=radii, tiles=tiles, keys=keys, vals=vals, ranges=ranges,
                render=render, alpha=alpha, last=last)
