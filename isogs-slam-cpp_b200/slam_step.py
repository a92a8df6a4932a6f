"""The tracking / mapping step around the rasterizer (north_star (4)-(5)), as fused device ops:

  map_step   = transform_to_frame + activations + render(rgb, depth, depth^2, silhouette) + masked-L1
               loss + backward + (optional gradient all-reduce) + fused Adam on all Gaussian params
  track_step = same forward with the frame pose on the device, sum-L1 loss with the silhouette
               mask, backward chained to the 7-dof pose, fused Adam on the pose only

Mirrors get_loss + loss.backward() + optimizer.step() of python_ordinary/scripts/splatam.py:494-760,
1247-1317 (tracking), 1412-1476 (mapping) with SSIM / flat / iso terms left out.
Parameters live in a capacity-doubling SoA pool (cpp/include/core/gaussian_model.hpp:18-124
without the unused 48-slot SH block).  torch is used for memory, streams and NCCL only.
"""
import ctypes as C

import torch

from . import _lib, raster
from ._lib import REC, check, lib
from .raster import _ptr, _stream

PARAM_ORDER = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
VAR_ORDER = ("means2D_gradient_accum", "denom", "max_2D_radius")

# reference learning rates (python_ordinary/configs/replica/splatam.py:88-96,109-117)
MAPPING_LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=0.05, log_scales=1e-3)
TRACKING_LRS = dict(cam_unnorm_rots=4e-4, cam_trans=2e-3)


class GaussianMap:
    """Replicated Gaussian map: parameters, Adam moments and one flat gradient buffer."""

    def __init__(self, params, capacity=None, device="cuda", grad_alloc=None):
        """`grad_alloc(numel) -> zeroed fp32 CUDA tensor` supplies the flat gradient buffer (dp.PeerAllReduce.alloc
        makes it NVLink-peer-mapped memory); default: a plain torch allocation."""
        self.grad_alloc = grad_alloc
        n = params["means3D"].shape[0]
        self.isotropic = params["log_scales"].shape[1] == 1
        self.sd = 1 if self.isotropic else 3
        self.widths = dict(means3D=3, rgb_colors=3, unnorm_rotations=4, logit_opacities=1, log_scales=self.sd)
        self.capacity = int(capacity or n)
        self.count = n
        self.device = torch.device(device)
        self.p, self.m, self.v = {}, {}, {}
        for k in PARAM_ORDER:
            w = self.widths[k]
            buf = torch.zeros((self.capacity, w), dtype=torch.float32, device=self.device)
            src = params[k]
            src = torch.from_numpy(src) if not torch.is_tensor(src) else src
            buf[:n] = src.reshape(n, w).to(self.device)
            self.p[k] = buf
            self.m[k] = torch.zeros_like(buf)
            self.v[k] = torch.zeros_like(buf)
        # per-Gaussian statistics of the gradient-driven densification (variables[...] of splatam.py:394-397)
        self.vars = {k: torch.zeros((self.capacity, 1), dtype=torch.float32, device=self.device) for k in VAR_ORDER}
        # one contiguous gradient buffer so a single all-reduce covers every parameter group
        self.grad_names = [k for k in PARAM_ORDER if not (self.isotropic and k == "unnorm_rotations")]
        total = sum(self.widths[k] for k in self.grad_names) * self.capacity + 4 * len(self.grad_names)
        self.grad_flat = self._new_grad_flat(total)
        self._carve()
        self.step_count = 0

    def _new_grad_flat(self, total):
        if self.grad_alloc is None:
            return torch.zeros((total,), dtype=torch.float32, device=self.device)
        self.grad_flat, self.g = None, {}                       # drop the views before the old buffer is freed
        buf = self.grad_alloc(total)
        if buf.numel() < total or buf.dtype != torch.float32 or buf.data_ptr() % 16:
            raise ValueError("grad_alloc returned an unusable buffer")
        return buf

    def _carve(self):
        # every group starts on a 16-byte boundary (the kernels store float4 rows); the padding floats stay 0
        self.g, off = {}, 0
        for k in self.grad_names:
            sz = self.widths[k] * self.count
            self.g[k] = self.grad_flat[off:off + sz].view(self.count, self.widths[k])
            off = (off + sz + 3) & ~3
        self.grad_used = off

    def view(self, k):
        return self.p[k][:self.count]

    def reset_optimizer(self):
        """The reference re-creates Adam for every tracking frame / mapping call (splatam.py:1237,1406)."""
        for k in PARAM_ORDER:
            self.m[k].zero_()
            self.v[k].zero_()
        self.step_count = 0

    def reset_stats(self):
        """add_new_gaussians and densify re-create the three statistics as zeros (splatam.py:835-837,
        slam_external.py:219-222)."""
        for k in VAR_ORDER:
            self.vars[k].zero_()

    def set_count(self, n):
        self.count = int(n)
        self._carve()

    def reserve(self, n):
        """Capacity-doubling pool (cpp/include/core/gaussian_model.hpp:18-124 pre-allocates 20 M slots; here
        the pool grows geometrically and keeps parameters + Adam moments; one copy per growth, not per frame
        as the reference's torch.cat does, splatam.py:832-833)."""
        if n <= self.capacity:
            return
        cap = max(int(n), int(self.capacity * 1.5) + 1024)
        for d in (self.p, self.m, self.v):
            for k in PARAM_ORDER:
                buf = torch.zeros((cap, self.widths[k]), dtype=torch.float32, device=self.device)
                buf[:self.count] = d[k][:self.count]
                d[k] = buf
        for k in VAR_ORDER:
            buf = torch.zeros((cap, 1), dtype=torch.float32, device=self.device)
            buf[:self.count] = self.vars[k][:self.count]
            self.vars[k] = buf
        self.capacity = cap
        total = sum(self.widths[k] for k in self.grad_names) * cap + 4 * len(self.grad_names)
        self.grad_flat = self._new_grad_flat(total)
        self._carve()


class StepWorkspace:
    """Per-(resolution, N) scratch that is reused across iterations (no per-step allocation)."""

    def __init__(self, cam, n_cap, device):
        self.cam = cam
        self.device = device
        H, W = cam.height, cam.width
        f32, i32 = torch.float32, torch.int32
        self.rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.v_rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.radii = torch.empty((n_cap, 2), dtype=i32, device=device)
        self.tiles = torch.empty((n_cap,), dtype=i32, device=device)
        self.cum = torch.empty((n_cap,), dtype=i32, device=device)
        self.n_cap = n_cap
        self.scan_b, _ = _lib.temp_bytes(n_cap, 0)
        self.scan_tmp = torch.empty((self.scan_b,), dtype=torch.uint8, device=device)
        self.ranges = torch.empty((cam.n_tiles + 1,), dtype=i32, device=device)
        self.render = torch.empty((5, H, W), dtype=f32, device=device)
        self.v_render = torch.empty((5, H, W), dtype=f32, device=device)
        self.alpha = torch.empty((H, W), dtype=f32, device=device)
        self.last = torch.empty((H, W), dtype=i32, device=device)
        self.small = torch.zeros((32,), dtype=f32, device=device)  # losses[4] | v_pose[7] | pose_acc[16]
        self.losses, self.v_pose, self.pose_acc = self.small[0:4], self.small[4:11], self.small[11:27]
        self.ssim = self.small[27:28]                               # mean SSIM of the last map_step(ssim=True)
        self.dmaps = None                                           # [9,H,W] SSIM derivative maps, on first use
        self.outlier_scratch = None                                 # [2,H,W] depth-error sort buffers, on first use
        self.isect_cap = 0
        self._grow_isects(max(1024, 3 * n_cap))
        # device-side intersection count {n_isects, overflow} + its lagged pinned mirror
        self.totals = torch.zeros((2,), dtype=i32, device=device)
        self.totals_host = torch.zeros((2,), dtype=i32).pin_memory()
        self._totals_ev = None

    @property
    def M(self):
        """Intersection count of the last forward (synchronises; for reports and tests only)."""
        return int(self.totals[0].item())

    def post_totals(self):
        """Queue an async copy of {n_isects, overflow} to pinned memory (no sync)."""
        self.totals_host.copy_(self.totals, non_blocking=True)
        self._totals_ev = torch.cuda.Event()
        self._totals_ev.record()

    def check_overflow(self, wait=False):
        """Look at the PREVIOUS forward's count if its copy has landed.  On overflow the tile lists of
        that forward were truncated (never out of bounds): grow the buffers and report it."""
        ev = self._totals_ev
        if ev is None or not (wait or ev.query()):
            return False
        if wait:
            ev.synchronize()
        self._totals_ev = None
        n, over = int(self.totals_host[0]), int(self.totals_host[1])
        if over:
            self._grow_isects(int(n * 1.25) + 1024)
            import warnings
            warnings.warn(f"tile-intersection buffers overflowed (needed {n}); grown to {self.isect_cap}. "
                          "The affected step rendered truncated tile lists.")
            return True
        return False

    def ensure_gaussians(self, n):
        """Grow the per-Gaussian scratch when the map has grown (densification)."""
        if n <= self.n_cap:
            return
        n_cap = int(n * 1.25) + 1024
        f32, i32, device = torch.float32, torch.int32, self.device
        self.rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.v_rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.radii = torch.empty((n_cap, 2), dtype=i32, device=device)
        self.tiles = torch.empty((n_cap,), dtype=i32, device=device)
        self.cum = torch.empty((n_cap,), dtype=i32, device=device)
        self.n_cap = n_cap
        self.scan_b, _ = _lib.temp_bytes(n_cap, 0)
        self.scan_tmp = torch.empty((self.scan_b,), dtype=torch.uint8, device=device)
        if self.isect_cap < 3 * n_cap:
            self._grow_isects(3 * n_cap)

    def _grow_isects(self, cap):
        dev = self.device
        self.isect_cap = int(cap)
        self.ku = torch.empty((cap,), dtype=torch.int64, device=dev)   # (depth | id) tile buckets (scratch)
        self.vs = torch.empty((cap,), dtype=torch.int32, device=dev)   # depth-sorted Gaussian ids per tile
        _, self.sort_b = _lib.temp_bytes(0, cap)
        self.sort_tmp = torch.empty((self.sort_b,), dtype=torch.uint8, device=dev)


def render_forward(ws, gmap, pose=None, pose_dev=None, ch=5, marks=None):
    """preprocess (raw params, fused activations + frame transform) -> bin -> blend, with NO host sync:
    the intersection count stays on the device (ws.totals) and is checked one step late
    (StepWorkspace.check_overflow), replacing the blocking `.item()` of gsplat Intersect.cpp:80."""
    cam, N = ws.cam, gmap.count
    flags = _lib.F_LOG_SCALES | _lib.F_LOGIT_OPAC
    L, st = lib(), _stream()
    ws.check_overflow()
    quats = None if gmap.isotropic else gmap.p["unnorm_rotations"]
    check(L.isogs_preprocess_fwd(C.byref(cam), C.byref(pose) if pose is not None else None, _ptr(pose_dev), N,
                                 _ptr(gmap.p["means3D"]), _ptr(gmap.p["log_scales"]), gmap.sd, _ptr(quats),
                                 _ptr(gmap.p["logit_opacities"]), _ptr(gmap.p["rgb_colors"]), ch, flags,
                                 _ptr(ws.rec), _ptr(ws.radii), None, None, None, 0, None, st))
    _mark(marks, "preprocess_fwd")
    check(L.isogs_tile_bin_async(C.byref(cam), N, ws.isect_cap, _ptr(ws.rec), _ptr(ws.radii), _ptr(ws.ku), None,
                                 _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.totals), _ptr(ws.sort_tmp), ws.sort_b, st))
    _mark(marks, "tile_bin")
    check(L.isogs_blend_fwd(C.byref(cam), ch, _ptr(ws.rec), _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.render),
                            _ptr(ws.alpha), _ptr(ws.last), st))
    ws.post_totals()


def render_backward(ws, gmap, pose=None, pose_dev=None, ch=5, want_gaussians=True, want_pose=False, marks=None):
    cam, N = ws.cam, gmap.count
    flags = _lib.F_LOG_SCALES | _lib.F_LOGIT_OPAC
    L, st = lib(), _stream()
    ws.v_rec[:N].zero_()
    _mark(marks, "zero_v_rec")
    if N > 0:
        # tracking needs only the geometric + depth-channel gradients (the pose's inputs)
        bwd = L.isogs_blend_bwd if want_gaussians else L.isogs_blend_bwd_geom
        check(bwd(C.byref(cam), ch, N, _ptr(ws.rec), _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.alpha),
                                _ptr(ws.last), _ptr(ws.v_render), None, _ptr(ws.v_rec), st))
    _mark(marks, "blend_bwd")
    quats = None if gmap.isotropic else gmap.p["unnorm_rotations"]
    g = gmap.g if want_gaussians else {}
    check(L.isogs_preprocess_bwd(C.byref(cam), C.byref(pose) if pose is not None else None, _ptr(pose_dev), N,
                                 _ptr(gmap.p["means3D"]), _ptr(gmap.p["log_scales"]), gmap.sd, _ptr(quats),
                                 _ptr(gmap.p["logit_opacities"]), ch, flags, _ptr(ws.radii), _ptr(ws.v_rec),
                                 _ptr(g.get("means3D")), _ptr(g.get("log_scales")), _ptr(g.get("unnorm_rotations")),
                                 _ptr(g.get("logit_opacities")), _ptr(g.get("rgb_colors")), None,
                                 _ptr(ws.v_pose) if want_pose else None, _ptr(ws.pose_acc) if want_pose else None, st))
    _mark(marks, "preprocess_bwd")


def loss_l1(ws, mode, gt_rgb, gt_depth, inv_n_depth=1.0, w_im=0.5, w_depth=1.0, sil_thres=0.99, use_sil=True,
            ignore_outlier_depth=False):
    thr = None
    if ignore_outlier_depth:    # splatam.py:551-554: 10 x median depth error, computed on the device (no sync)
        P = ws.cam.width * ws.cam.height
        if ws.outlier_scratch is None:
            ws.outlier_scratch = torch.empty((2 * P,), dtype=torch.float32, device=ws.device)
        if ws.sort_b < _lib.temp_bytes(0, P)[1]:
            ws._grow_isects(max(ws.isect_cap, P))
        thr = ws.small[28:29]
        check(lib().isogs_depth_outlier_threshold(C.byref(ws.cam), _ptr(ws.render[3]), _ptr(gt_depth), 10.0,
                                                  _ptr(ws.outlier_scratch), _ptr(ws.sort_tmp), ws.sort_b, _ptr(thr),
                                                  _stream()))
    check(lib().isogs_loss_l1_fused(C.byref(ws.cam), mode, _ptr(ws.render), _ptr(ws.alpha), _ptr(gt_rgb),
                                    _ptr(gt_depth), w_im, w_depth, sil_thres, int(use_sil), inv_n_depth, _ptr(thr),
                                    _ptr(ws.v_render), _ptr(ws.losses), _stream()))


def loss_ssim(ws, gt_rgb, weight):
    """losses[0] += weight * (1 - SSIM(render, gt)); v_render[0:3] += its gradient (calc_ssim,
    slam_external.py:54-97).  Must follow loss_l1, which overwrites v_render.  ws.ssim receives mean SSIM."""
    if ws.dmaps is None:
        ws.dmaps = torch.empty((9, ws.cam.height, ws.cam.width), dtype=torch.float32, device=ws.device)
    check(lib().isogs_ssim_loss_fused(C.byref(ws.cam), _ptr(ws.render), _ptr(gt_rgb), weight, _ptr(ws.dmaps),
                                      _ptr(ws.v_render), _ptr(ws.losses), _ptr(ws.ssim), _stream()))


def iso_flat_loss(ws, gmap, w_flat=50.0, w_iso=2.0, seed=0, query_idx=None, sample_size=8192, res=None,
                  with_grad=True):
    """The IsoGS geometric terms of get_loss (splatam.py:572-675, weights :733-739): losses[0] += 50 * flat +
    2 * iso and their gradients accumulated into gmap.g (call after render_backward).  Exact 16-NN through a
    device-built uniform grid instead of the reference's 8192 x N cdist.  ws.geo receives {flat, iso, mean
    density}; ws.nbr the neighbour table.  Three-axis scales only (the reference skips both terms otherwise)."""
    if gmap.isotropic:
        raise ValueError("iso / flat losses need three-axis scales (splatam.py:586,615)")
    N, dev = gmap.count, gmap.device
    if query_idx is not None:
        Q = int(query_idx.numel())
    else:
        Q = N if N < sample_size else sample_size          # :648-656
    if res is None:
        res = max(8, min(160, int(round(1.28 * N ** (1.0 / 3.0)))))
    key = (N, Q, res)
    if getattr(ws, "_geo_key", None) != key:
        gb, sb = C.c_size_t(0), C.c_size_t(0)
        check(lib().isogs_knn_workspace(N, res, C.byref(gb), C.byref(sb)))
        ws.knn_grid = torch.empty((gb.value // 4,), dtype=torch.int32, device=dev)
        ws.knn_sorted = torch.empty((sb.value // 4,), dtype=torch.float32, device=dev)
        ws.knn_scan_b, _ = _lib.temp_bytes((res + 1) ** 3 + 1, 0)
        ws.knn_scan = torch.empty((ws.knn_scan_b,), dtype=torch.uint8, device=dev)
        ws.nbr = torch.empty((Q, 16), dtype=torch.int32, device=dev)
        ws.query_idx = torch.empty((Q,), dtype=torch.int32, device=dev)
        ws._geo_key = key
    L, st = lib(), _stream()
    sampled = None
    if query_idx is None and Q != N:
        sampled, qptr = ws.query_idx, None
    else:
        qptr = query_idx
    check(L.isogs_knn16(N, _ptr(gmap.p["means3D"]), res, Q, _ptr(qptr), seed, _ptr(sampled), _ptr(ws.knn_grid),
                        _ptr(ws.knn_sorted), _ptr(ws.knn_scan), ws.knn_scan_b, _ptr(ws.nbr), st))
    q_used = qptr if qptr is not None else sampled
    ws.geo = ws.small[29:32]
    ws.geo.zero_()
    g = gmap.g if with_grad else {}
    check(L.isogs_iso_flat_loss(N, _ptr(gmap.p["means3D"]), _ptr(gmap.p["log_scales"]), _ptr(gmap.p["unnorm_rotations"]),
                                _ptr(gmap.p["logit_opacities"]), Q, _ptr(q_used), _ptr(ws.nbr), w_iso, w_flat, 1.0,
                                _ptr(g.get("means3D")), _ptr(g.get("log_scales")), _ptr(g.get("unnorm_rotations")),
                                _ptr(g.get("logit_opacities")), _ptr(ws.losses), _ptr(ws.geo), st))


def adam_gaussians(gmap, lrs=MAPPING_LRS, eps=1e-15, beta1=0.9, beta2=0.999):
    gmap.step_count += 1
    names = gmap.grad_names
    arr = (_lib.AdamGroup * len(names))()
    for i, k in enumerate(names):
        arr[i].param = gmap.p[k].data_ptr()
        arr[i].grad = gmap.g[k].data_ptr()
        arr[i].exp_avg = gmap.m[k].data_ptr()
        arr[i].exp_avg_sq = gmap.v[k].data_ptr()
        arr[i].numel = gmap.count * gmap.widths[k]
        arr[i].lr = lrs[k]
    check(lib().isogs_adam_fused(arr, len(names), beta1, beta2, eps, gmap.step_count, _stream()))


def _mark(marks, name):
    if marks is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))


def map_step(ws, gmap, pose, gt_rgb, gt_depth, inv_n_depth, lrs=MAPPING_LRS, allreduce=None, marks=None, w_im=0.5,
             w_depth=1.0, ssim=False, isoflat=False, iso_seed=0):
    """One mapping iteration on one keyframe (splatam.py:1412-1476, L1 terms).  `pose` is the keyframe's
    HOST pose (camera_grad=False).  `allreduce(flat_grad_tensor)` is called between backward and Adam.
    `marks` (a list) receives (stage, cuda event) pairs for per-stage timing.  `w_im` weighs the mean colour
    term; `ssim=True` makes that term the reference's 0.8 * L1 + 0.2 * (1 - SSIM) (splatam.py:568), else pure L1;
    `isoflat=True` adds 50 * flat + 2 * iso (:572-675,733-739; three-axis scales) -- with both flags the loss is the
    reference's complete mapping loss."""
    _mark(marks, "begin")
    ws.small.zero_()
    render_forward(ws, gmap, pose=pose, marks=marks)
    _mark(marks, "blend_fwd")
    loss_l1(ws, 1, gt_rgb, gt_depth, inv_n_depth, w_im=(0.8 * w_im if ssim else w_im), w_depth=w_depth)
    _mark(marks, "loss")
    if ssim:
        loss_ssim(ws, gt_rgb, 0.2 * w_im)
        _mark(marks, "ssim")
    render_backward(ws, gmap, pose=pose, want_gaussians=True, marks=marks)
    if isoflat:
        iso_flat_loss(ws, gmap, seed=iso_seed)
        _mark(marks, "isoflat")
    if allreduce is not None:
        allreduce(gmap.grad_flat[:gmap.grad_used])
        _mark(marks, "allreduce")
    adam_gaussians(gmap, lrs)
    _mark(marks, "adam")
    return ws.losses


def _pose_of(pose):
    return pose if isinstance(pose, _lib.Pose) else _lib.make_pose(pose[0], pose[1])


def add_new_gaussians(ws, gmap, pose, gt_rgb, gt_depth, sil_thres=0.5, noise_std=0.01, seed=0):
    """Silhouette-driven densification of one frame (add_new_gaussians, splatam.py:799-841): silhouette +
    depth render at `pose` (no gradient), non-presence mask `sil < sil_thres or (render_d > gt_d and
    |err| > 50 median(err))`, back-projection of the masked valid-depth pixels, append to the pool.
    Appended Gaussians start with zero Adam moments.  Two host reads per mapped frame (the median
    threshold and the new count).  Returns the number of Gaussians added."""
    pose = _pose_of(pose)
    cam = ws.cam
    P = cam.width * cam.height
    ws.ensure_gaussians(gmap.count)
    render_forward(ws, gmap, pose=pose)
    # 50 x median(|gt_d - d| * (gt_d > 0)) (splatam.py:811-812) by the device-side median kernel; one host read
    if ws.outlier_scratch is None:
        ws.outlier_scratch = torch.empty((2 * P,), dtype=torch.float32, device=ws.device)
    if ws.sort_b < _lib.temp_bytes(0, P)[1]:
        ws._grow_isects(max(ws.isect_cap, P))
    check(lib().isogs_depth_outlier_threshold(C.byref(cam), _ptr(ws.render[3]), _ptr(gt_depth), 50.0,
                                              _ptr(ws.outlier_scratch), _ptr(ws.sort_tmp), ws.sort_b,
                                              _ptr(ws.small[28:29]), _stream()))
    thr = float(ws.small[28])
    gmap.reserve(gmap.count + P)
    count = torch.tensor([gmap.count], dtype=torch.int32, device=gmap.device)
    scan = torch.empty((2 * P,), dtype=torch.int32, device=gmap.device)
    sb, _ = _lib.temp_bytes(P, 0)
    tmp = torch.empty((sb,), dtype=torch.uint8, device=gmap.device)
    quats = gmap.p["unnorm_rotations"]
    check(lib().isogs_backproject_append(C.byref(cam), C.byref(pose), _ptr(gt_rgb), _ptr(gt_depth), _ptr(ws.render[3]),
                                         _ptr(ws.alpha), sil_thres, thr, _ptr(gmap.p["means3D"]),
                                         _ptr(gmap.p["rgb_colors"]), _ptr(quats), _ptr(gmap.p["logit_opacities"]),
                                         _ptr(gmap.p["log_scales"]), gmap.sd, noise_std, seed, gmap.capacity, _ptr(scan),
                                         _ptr(tmp), sb, _ptr(count), _stream()))
    new_count = int(count.item())
    added = new_count - gmap.count
    if added > 0:
        for k in PARAM_ORDER:
            gmap.m[k][gmap.count:new_count].zero_()
            gmap.v[k][gmap.count:new_count].zero_()
        gmap.set_count(new_count)
        gmap.reset_stats()
        ws.ensure_gaussians(new_count)
    return added


def _compact(gmap, keep, N):
    """remove_points (slam_external.py:139-160): ONE order-preserving gather of the 5 parameter tensors, their
    10 Adam moments and the 3 densification statistics.  Returns the new count (one host read)."""
    dev = gmap.device
    rows_in, rows_out = [], []
    for d, names in ((gmap.p, PARAM_ORDER), (gmap.m, PARAM_ORDER), (gmap.v, PARAM_ORDER), (gmap.vars, VAR_ORDER)):
        for k in names:
            rows_in.append((d, k))
            rows_out.append(torch.zeros_like(d[k]))
    arr = (_lib.Rows * len(rows_in))()
    for i, ((d, k), out) in enumerate(zip(rows_in, rows_out)):
        arr[i].inp, arr[i].out, arr[i].width = d[k].data_ptr(), out.data_ptr(), d[k].shape[1]
    scan = torch.empty((N,), dtype=torch.int32, device=dev)
    sb, _ = _lib.temp_bytes(N, 0)
    tmp = torch.empty((sb,), dtype=torch.uint8, device=dev)
    cnt = torch.zeros((1,), dtype=torch.int32, device=dev)
    check(lib().isogs_compact(N, _ptr(keep), arr, len(rows_in), _ptr(scan), _ptr(tmp), sb, _ptr(cnt), _stream()))
    new_count = int(cnt.item())
    if new_count != N:
        for (d, k), out in zip(rows_in, rows_out):
            d[k] = out
        gmap.set_count(new_count)
    return new_count


def prune_gaussians(ws, gmap, opac_thres=0.005, max_scale=float("inf")):
    """prune_gaussians + remove_points (slam_external.py:139-188): drop `sigmoid(o) < opac_thres or
    max exp(s) > max_scale`; one compaction.  Returns the number removed (one host read)."""
    N = gmap.count
    if N == 0:
        return 0
    keep = torch.empty((N,), dtype=torch.uint8, device=gmap.device)
    check(lib().isogs_prune_mask(N, _ptr(gmap.p["logit_opacities"]), _ptr(gmap.p["log_scales"]), gmap.sd, opac_thres,
                                 max_scale, _ptr(keep), _stream()))
    return N - _compact(gmap, keep, N)


def accumulate_densify_stats(ws, gmap, v_xy=None, scale=(1.0, 1.0)):
    """accumulate_mean2d_gradient + the max_2D_radius update of get_loss (slam_external.py:100-104,
    splatam.py:751-753) for the Gaussians seen by the last render.  v_xy defaults to the screen-space gradient
    of the last backward (ws.v_rec, pixel units).  NOTE: the reference accumulates the gradient of the COLOUR
    render only (splatam.py:528-530); the fused 5-channel step's v_rec also carries the depth-loss part -- use
    `colour_only_screen_gradient` for the reference's exact statistic."""
    v = ws.v_rec if v_xy is None else v_xy
    check(lib().isogs_densify_stats(gmap.count, _ptr(ws.radii), _ptr(v), v.shape[1], scale[0], scale[1],
                                    _ptr(gmap.vars["means2D_gradient_accum"]), _ptr(gmap.vars["denom"]),
                                    _ptr(gmap.vars["max_2D_radius"]), _stream()))


def colour_only_screen_gradient(ws, gmap):
    """Second blend backward with the depth planes of v_render zeroed: dL_colour/d(mean2D) exactly as the
    reference's separate colour render sees it.  Costs one extra blend_bwd; only for densification statistics."""
    N = gmap.count
    if getattr(ws, "v_rec2", None) is None or ws.v_rec2.shape[0] < ws.v_rec.shape[0]:
        ws.v_rec2 = torch.empty_like(ws.v_rec)
        ws.v_render2 = torch.zeros_like(ws.v_render)
    ws.v_render2[:3].copy_(ws.v_render[:3])
    ws.v_rec2[:N].zero_()
    check(lib().isogs_blend_bwd(C.byref(ws.cam), 5, N, _ptr(ws.rec), _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.alpha),
                                _ptr(ws.last), _ptr(ws.v_render2), None, _ptr(ws.v_rec2), _stream()))
    return ws.v_rec2


DENSIFY_DEFAULTS = dict(start_after=500, remove_big_after=3000, stop_after=5000, densify_every=100, grad_thresh=0.0002,
                        num_to_split_into=2, removal_opacity_threshold=0.005, final_removal_opacity_threshold=0.005,
                        reset_opacities=False, reset_opacities_every=3000)   # configs/replica/splatam.py:130-141


def densify(ws, gmap, it, scene_radius, dd=DENSIFY_DEFAULTS, seed=0, noise=None, accumulate=True, v_xy=None):
    """densify (slam_external.py:191-243): accumulate the statistics, and every `densify_every` iterations clone
    the small / split the large high-gradient Gaussians, drop the split parents, prune by opacity / size -- two
    device passes (select, emit) + ONE compaction instead of the reference's ~40 torch ops and 4 torch.cat
    re-allocations.  Two host reads (the new count to size the pool, the count after compaction).
    `noise` ([n*n_split,3] standard normals, reference row order) overrides the built-in generator (tests).
    Returns (n_clone, n_split, n_removed) or None when nothing was due."""
    if it > dd["stop_after"]:
        return None
    if accumulate:
        accumulate_densify_stats(ws, gmap, v_xy)
    out = None
    if it >= dd["start_after"] and it % dd["densify_every"] == 0:
        N, n_into, dev = gmap.count, int(dd["num_to_split_into"]), gmap.device
        L, st = lib(), _stream()
        scratch = torch.empty((4 * N,), dtype=torch.int32, device=dev)
        sb, _ = _lib.temp_bytes(N, 0)
        tmp = torch.empty((sb,), dtype=torch.uint8, device=dev)
        counts = torch.zeros((3,), dtype=torch.int32, device=dev)
        check(L.isogs_densify_select(N, _ptr(gmap.vars["means2D_gradient_accum"]), _ptr(gmap.vars["denom"]),
                                     _ptr(gmap.p["log_scales"]), gmap.sd, dd["grad_thresh"], 0.01 * scene_radius, n_into,
                                     _ptr(scratch), _ptr(tmp), sb, _ptr(counts), st))
        total, n_clone, n_split = (int(x) for x in counts.cpu())
        gmap.reserve(total)
        keep = torch.empty((total,), dtype=torch.uint8, device=dev)
        check(L.isogs_densify_emit(N, gmap.capacity, _ptr(scratch), _ptr(gmap.p["means3D"]), _ptr(gmap.p["rgb_colors"]),
                                   _ptr(gmap.p["unnorm_rotations"]), _ptr(gmap.p["logit_opacities"]),
                                   _ptr(gmap.p["log_scales"]), gmap.sd, n_into, _ptr(noise), seed, _ptr(keep), st))
        for k in PARAM_ORDER:                                   # cat_params_to_optimizer: new rows start at zero
            gmap.m[k][N:total].zero_()
            gmap.v[k][N:total].zero_()
        gmap.set_count(total)
        gmap.reset_stats()                                      # :219-222
        thr = dd["final_removal_opacity_threshold"] if it == dd["stop_after"] else dd["removal_opacity_threshold"]
        big = 0.1 * scene_radius if it >= dd["remove_big_after"] else float("inf")
        keep2 = torch.empty((total,), dtype=torch.uint8, device=dev)
        check(L.isogs_prune_mask(total, _ptr(gmap.p["logit_opacities"]), _ptr(gmap.p["log_scales"]), gmap.sd, thr, big,
                                 _ptr(keep2), st))
        check(L.isogs_mask_and(total, _ptr(keep), _ptr(keep2), st))
        left = _compact(gmap, keep, total)
        ws.ensure_gaussians(left)
        out = (n_clone, n_split, total - left)
    if it > 0 and it % dd["reset_opacities_every"] == 0 and dd["reset_opacities"]:
        # inverse_sigmoid(0.01), fresh Adam state for the group (update_params_and_optimizer, :106-119)
        gmap.p["logit_opacities"][:gmap.count].fill_(-4.59511985013459)
        gmap.m["logit_opacities"].zero_()
        gmap.v["logit_opacities"].zero_()
    return out


def keyframe_overlap(cam, gt_depth, cur_w2c, kf_w2cs, pixels=1600, seed=0, sample_idx=None, edge=20):
    """percent_inside of keyframe_selection_overlap (keyframe_selection.py:40-84) for every keyframe, on the
    device: DEVICE float[K].  cur_w2c: 4x4 (or 3x4) array-like; kf_w2cs: [K,4,4] / [K,3,4] array-like or tensor."""
    import numpy as np
    dev = gt_depth.device
    P = cam.width * cam.height
    kf = torch.as_tensor(np.asarray(kf_w2cs, dtype=np.float32)[:, :3, :4].copy()).contiguous().to(dev)
    K = kf.shape[0]
    w2c = (C.c_float * 12)(*np.asarray(cur_w2c, dtype=np.float32)[:3, :4].reshape(-1).tolist())
    si = torch.empty((P + pixels,), dtype=torch.int32, device=dev)
    sf = torch.empty((6 * pixels,), dtype=torch.float32, device=dev)
    sb, _ = _lib.temp_bytes(P, 0)
    tmp = torch.empty((sb,), dtype=torch.uint8, device=dev)
    out = torch.zeros((max(K, 1),), dtype=torch.float32, device=dev)
    check(lib().isogs_keyframe_overlap(C.byref(cam), _ptr(gt_depth), w2c, _ptr(sample_idx), pixels, seed, _ptr(kf), K,
                                       edge, _ptr(si), _ptr(sf), _ptr(tmp), sb, _ptr(out), _stream()))
    return out[:K]


def select_keyframes(cam, gt_depth, cur_w2c, kf_w2cs, k, pixels=1600, seed=0, rng=None, sample_idx=None):
    """keyframe_selection_overlap (keyframe_selection.py:40-95): rank by overlap (stable, descending), keep those
    with overlap > 0, return a random choice of k of them (one host read of K floats)."""
    import numpy as np
    pct = keyframe_overlap(cam, gt_depth, cur_w2c, kf_w2cs, pixels, seed, sample_idx).cpu().numpy()
    order = sorted(range(len(pct)), key=lambda i: pct[i], reverse=True)
    sel = [i for i in order if pct[i] > 0.0]
    rng = rng or np.random
    return list(rng.permutation(np.array(sel, dtype=np.int64))[:k])


class PoseState:
    """cam_unnorm_rots / cam_trans of the frame being tracked, on the device, with Adam moments."""

    def __init__(self, q, t, device="cuda"):
        self.pose = torch.tensor(list(q) + list(t), dtype=torch.float32, device=device)
        self.m = torch.zeros_like(self.pose)
        self.v = torch.zeros_like(self.pose)
        # best candidate {q, t, min loss} (splatam.py:1240-1242,1287-1290), kept on the device
        self.best = torch.tensor(list(q) + list(t) + [1e20], dtype=torch.float32, device=device)
        self.step_count = 0

    def best_pose(self):
        """(q, t) of the best candidate -- what the reference writes back (splatam.py:1321-1323)."""
        b = self.best.cpu().numpy()
        return b[:4].copy(), b[4:7].copy()


def init_camera_pose(prev1, prev2=None, forward_prop=True):
    """initialize_camera_pose (splatam.py:844-863): constant-velocity initialisation of the next frame's
    pose from the two previous (q, t) pairs; HOST arithmetic on 7 floats."""
    import numpy as np
    q1, t1 = (np.asarray(a, dtype=np.float32) for a in prev1)
    if prev2 is None or not forward_prop:
        return q1.copy(), t1.copy()
    q2, t2 = (np.asarray(a, dtype=np.float32) for a in prev2)
    n1 = q1 / max(float(np.linalg.norm(q1)), 1e-12)
    n2 = q2 / max(float(np.linalg.norm(q2)), 1e-12)
    q = n1 + (n1 - n2)
    q = q / max(float(np.linalg.norm(q)), 1e-12)
    return q.astype(np.float32), (t1 + (t1 - t2)).astype(np.float32)


def track_step(ws, gmap, ps, gt_rgb, gt_depth, lrs=TRACKING_LRS, eps=1e-8, use_sil=True, sil_thres=0.99,
               ignore_outlier_depth=False):
    """One tracking iteration (splatam.py:1247-1290): only the pose gets a gradient and a step."""
    ws.small.zero_()
    render_forward(ws, gmap, pose_dev=ps.pose)
    loss_l1(ws, 0, gt_rgb, gt_depth, 1.0, sil_thres=sil_thres, use_sil=use_sil,
            ignore_outlier_depth=ignore_outlier_depth)
    render_backward(ws, gmap, pose_dev=ps.pose, want_gaussians=False, want_pose=True)
    ps.step_count += 1
    arr = (_lib.AdamGroup * 2)()
    for i, (off, n, lr) in enumerate(((0, 4, lrs["cam_unnorm_rots"]), (4, 3, lrs["cam_trans"]))):
        arr[i].param = ps.pose.data_ptr() + 4 * off
        arr[i].grad = ws.v_pose.data_ptr() + 4 * off
        arr[i].exp_avg = ps.m.data_ptr() + 4 * off
        arr[i].exp_avg_sq = ps.v.data_ptr() + 4 * off
        arr[i].numel = n
        arr[i].lr = lr
    check(lib().isogs_adam_fused(arr, 2, 0.9, 0.999, eps, ps.step_count, _stream()))
    check(lib().isogs_pose_keep_best(_ptr(ws.losses), _ptr(ps.pose), _ptr(ps.best), _stream()))
    return ws.losses
