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
        # isogs_adam_state on the DEVICE (step, bias corrections): advanced by the step's first kernel
        self.adam_state = torch.zeros((4,), dtype=torch.float64, device=self.device)

    def c_struct(self, with_grads=True):
        """isogs_gaussians over the current buffers (pointers only)."""
        key = (self.count, with_grads, tuple(self.p[k].data_ptr() for k in PARAM_ORDER), self.grad_flat.data_ptr())
        hit = getattr(self, "_cs", None)
        if hit is not None and hit[0] == key:
            return hit[1]
        g = _lib.Gaussians()
        g.N, g.scale_dim = self.count, self.sd
        for i, k in enumerate(PARAM_ORDER):
            g.p[i] = self.p[k].data_ptr()
            has = with_grads and k in self.g
            g.g[i] = self.g[k].data_ptr() if has else None
            g.m[i] = self.m[k].data_ptr() if has else None
            g.v[i] = self.v[k].data_ptr() if has else None
        self._cs = (key, g)
        return g

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
        self.adam_state.zero_()

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
        self.tile_rects = torch.empty((n_cap, 4), dtype=i32, device=device)   # tile rect + depth bits per Gaussian
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
        self.median_tmp, self.median_b = None, 0                    # CUB temp of the median sort (its OWN buffer)
        self.isect_cap = 0
        self._grow_isects(max(1024, 3 * n_cap))
        # device-side intersection count {n_isects, overflow} + its lagged pinned mirror
        self.totals = torch.zeros((2,), dtype=i32, device=device)
        self.totals_host = torch.zeros((2,), dtype=i32).pin_memory()
        self._totals_ev = None
        # the DEVICE-side current-frame slot (isogs_frame) a captured graph reads, per-keyframe frame structs and
        # the captured graphs themselves
        self.cur_frame = torch.zeros((C.sizeof(_lib.Frame),), dtype=torch.uint8, device=device)
        self._frames = {}
        self._graphs = {}
        self._cap_stream = None
        self._cws = None

    def c_struct(self):
        """isogs_workspace over the current buffers (rebuilt when a buffer was re-allocated)."""
        bufs = (self.rec, self.v_rec, self.radii, self.ku, self.vs, self.ranges, self.totals, self.sort_tmp, self.render,
                self.v_render, self.alpha, self.last, self.small, self.dmaps, self.outlier_scratch, self.median_tmp,
                self.tile_rects)
        key = tuple(0 if b is None else b.data_ptr() for b in bufs) + (self.isect_cap,)
        if self._cws is not None and self._cws[0] == key:
            return self._cws[1]
        w = _lib.Workspace()
        w.rec, w.v_rec, w.radii = self.rec.data_ptr(), self.v_rec.data_ptr(), self.radii.data_ptr()
        w.bucket, w.vals, w.ranges = self.ku.data_ptr(), self.vs.data_ptr(), self.ranges.data_ptr()
        w.totals, w.sort_temp, w.sort_temp_bytes = self.totals.data_ptr(), self.sort_tmp.data_ptr(), self.sort_b
        w.isect_cap = self.isect_cap
        w.render, w.v_render = self.render.data_ptr(), self.v_render.data_ptr()
        w.alpha, w.last, w.small = self.alpha.data_ptr(), self.last.data_ptr(), self.small.data_ptr()
        w.adam_state = None                                   # filled per call (the map's or the pose's state)
        w.dmaps = None if self.dmaps is None else self.dmaps.data_ptr()
        w.median_scratch = None if self.outlier_scratch is None else self.outlier_scratch.data_ptr()
        w.median_temp = None if self.median_tmp is None else self.median_tmp.data_ptr()
        w.median_temp_bytes = self.median_b
        w.tile_rects = self.tile_rects.data_ptr()
        self._cws = (key, w)
        self.drop_graphs()                                    # captured launches hold the old pointers
        return w

    def frame(self, gt_rgb, gt_depth, pose=None, inv_n_depth=1.0):
        """DEVICE isogs_frame for (ground truth, keyframe pose, 1/n_valid): built once per distinct value (one small
        H2D copy), then reused -- the per-iteration path has no host->device traffic for it."""
        q = (1.0, 0.0, 0.0, 0.0) if pose is None else tuple(float(v) for v in pose.q)
        t = (0.0, 0.0, 0.0) if pose is None else tuple(float(v) for v in pose.t)
        key = (gt_rgb.data_ptr(), gt_depth.data_ptr(), q, t, float(inv_n_depth))
        hit = self._frames.get(key)
        if hit is not None:
            return hit
        f = _lib.Frame()
        f.gt_rgb, f.gt_depth = gt_rgb.data_ptr(), gt_depth.data_ptr()
        f.pose[:] = list(q) + list(t)
        f.inv_n_depth = float(inv_n_depth)
        host = torch.frombuffer(bytearray(bytes(f)), dtype=torch.uint8)
        dev = host.to(self.device)
        if len(self._frames) >= 256:
            self._frames.pop(next(iter(self._frames)))
        self._frames[key] = dev
        return dev

    def drop_graphs(self):
        for ge in self._graphs.values():
            lib().isogs_graph_destroy(ge)
        self._graphs = {}

    def __del__(self):
        try:
            self.drop_graphs()
        except Exception:
            pass

    @property
    def M(self):
        """Intersection count of the last forward (synchronises; for reports and tests only)."""
        return int(self.totals[0].item())

    def post_totals(self):
        """Queue an async copy of {n_isects, overflow} to pinned memory (no sync)."""
        self.totals_host.copy_(self.totals, non_blocking=True)
        if self._totals_ev is None:
            self._totals_ev = torch.cuda.Event()
        self._totals_ev.record()

    def check_overflow(self, wait=False):
        """The overflow word is STICKY on the device (tile_scan_kernel ORs into totals[1]; the largest count seen
        while it is set is kept in totals[0]) and mirrored to pinned memory after every forward, so the host reads
        the mirror WITHOUT gating on the newest copy: whichever copy has landed carries the flag.  While the flag is
        set the Adam / best-pose kernels of the affected steps are device-side no-ops (they are handed totals + 1),
        so no parameter was updated from a truncated render.  On overflow: grow the buffers, clear the flag."""
        if wait and self._totals_ev is not None:
            self._totals_ev.synchronize()
        if int(self.totals_host[1]) == 0:
            return False
        torch.cuda.current_stream().synchronize()              # rare path: settle, then read the final count
        n = int(self.totals[0].item())
        self._grow_isects(max(int(n * 1.25) + 1024, self.isect_cap))
        self.totals.zero_()
        self.totals_host.zero_()
        import warnings
        warnings.warn(f"tile-intersection buffers overflowed (needed {n}); grown to {self.isect_cap}. "
                      "The affected steps were skipped on the device (no Adam / best-pose update).")
        return True

    def median_temp(self):
        """Temp storage of the device-side median (CUB key sort over P depth errors).  Its own buffer: the
        intersection buffers (vs / ku / sort_tmp) must not be re-allocated between a forward and its backward."""
        P = self.cam.width * self.cam.height
        if self.outlier_scratch is None:
            self.outlier_scratch = torch.empty((2 * P,), dtype=torch.float32, device=self.device)
        if self.median_tmp is None:
            _, self.median_b = _lib.temp_bytes(0, P)
            self.median_tmp = torch.empty((self.median_b,), dtype=torch.uint8, device=self.device)
        return self.outlier_scratch, self.median_tmp, self.median_b

    def ensure_gaussians(self, n):
        """Grow the per-Gaussian scratch when the map has grown (densification)."""
        if n <= self.n_cap:
            return
        n_cap = int(n * 1.25) + 1024
        f32, i32, device = torch.float32, torch.int32, self.device
        self.rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.v_rec = torch.empty((n_cap, REC), dtype=f32, device=device)
        self.radii = torch.empty((n_cap, 2), dtype=i32, device=device)
        self.tile_rects = torch.empty((n_cap, 4), dtype=i32, device=device)
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


def render_forward(ws, gmap, pose=None, pose_dev=None, ch=5, marks=None, zero_v_rec=False):
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
                                 _ptr(ws.rec), _ptr(ws.radii), None, None, None, 0, None,
                                 _ptr(ws.v_rec) if zero_v_rec else None, _ptr(ws.tile_rects), None, st))
    _mark(marks, "preprocess_fwd")
    check(L.isogs_tile_bin_async(C.byref(cam), N, ws.isect_cap, _ptr(ws.rec), _ptr(ws.radii), _ptr(ws.ku), None,
                                 _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.totals), _ptr(ws.sort_tmp), ws.sort_b,
                                 _ptr(ws.tile_rects), 0, st))
    _mark(marks, "tile_bin")
    check(L.isogs_blend_fwd(C.byref(cam), ch, _ptr(ws.rec), _ptr(ws.vs), _ptr(ws.ranges), _ptr(ws.render),
                            _ptr(ws.alpha), _ptr(ws.last), st))
    ws.post_totals()


def render_backward(ws, gmap, pose=None, pose_dev=None, ch=5, want_gaussians=True, want_pose=False, marks=None,
                    zero_v_rec=True):
    """Per-op backward (the fused steps do not use it).  zero_v_rec=False when the forward already cleared v_rec
    (render_forward(zero_v_rec=True))."""
    cam, N = ws.cam, gmap.count
    flags = _lib.F_LOG_SCALES | _lib.F_LOGIT_OPAC
    L, st = lib(), _stream()
    if zero_v_rec:
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
        scratch, tmp, tmp_b = ws.median_temp()
        thr = ws.small[28:29]
        check(lib().isogs_depth_outlier_threshold(C.byref(ws.cam), _ptr(ws.render[3]), _ptr(gt_depth), 10.0,
                                                  _ptr(scratch), _ptr(tmp), tmp_b, _ptr(thr), _stream()))
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


def adam_gaussians(gmap, lrs=MAPPING_LRS, eps=1e-15, beta1=0.9, beta2=0.999, skip_flag=None):
    """Stand-alone fused Adam with the HOST step count (torch.optim.Adam); the fused map_step keeps its own count on
    the device (gmap.adam_state) -- do not mix the two on one map without reset_optimizer()."""
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
    check(lib().isogs_adam_fused(arr, len(names), beta1, beta2, eps, gmap.step_count, None, _ptr(skip_flag), _stream()))


def _mark(marks, name):
    if marks is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))


def _params(flags, phases, w_im, w_depth, sil_thres, lr5, eps, beta1=0.9, beta2=0.999, lr_pose=(0.0, 0.0)):
    p = _lib.StepParams()
    p.lr_pose[:] = [float(v) for v in lr_pose]
    p.flags, p.phases = int(flags), int(phases)
    p.w_im, p.w_depth, p.sil_thres = float(w_im), float(w_depth), float(sil_thres)
    p.lr[:] = [float(v) for v in lr5]
    p.beta1, p.beta2, p.eps = float(beta1), float(beta2), float(eps)
    return p


def _event_array(marks):
    """ISOGS_STEP_EVENTS timing events for one fused step; appended to `marks` as (stage, event) pairs."""
    if marks is None:
        return None
    arr = (C.c_void_p * _lib.STEP_EVENTS)()
    for i, name in enumerate(_lib.STEP_EVENT_NAMES):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()                               # creates the underlying cudaEvent_t; re-recorded by the step
        arr[i] = ev.cuda_event
        marks.append((name, ev))
    return arr


def _graph_launch(ws, key, enqueue):
    """Replay the captured launch sequence for `key`, capturing it on first use (capture does not execute).
    `enqueue(stream)` issues the work.  Capture happens on a private side stream (the legacy default stream cannot
    capture); the replay is launched on the caller's current stream."""
    L = lib()
    ge = ws._graphs.get(key)
    if ge is None:
        if ws._cap_stream is None:
            ws._cap_stream = torch.cuda.Stream(device=ws.device)
        cs = C.c_void_p(ws._cap_stream.cuda_stream)
        check(L.isogs_graph_begin(cs))
        try:
            enqueue(cs)
        finally:
            h = C.c_void_p()
            rc = L.isogs_graph_end(cs, C.byref(h))
        check(rc)
        ge = ws._graphs[key] = h
    check(L.isogs_graph_launch(ge, _stream()))


def graph_node_counts(ws):
    """{key: (kernel nodes, other nodes)} of every graph this workspace has captured (counted from the graphs)."""
    out = {}
    for key, h in ws._graphs.items():
        k, o = C.c_int32(0), C.c_int32(0)
        check(lib().isogs_graph_info(h, C.byref(k), C.byref(o)))
        out[key] = (int(k.value), int(o.value))
    return out


def map_step(ws, gmap, pose, gt_rgb, gt_depth, inv_n_depth, lrs=MAPPING_LRS, allreduce=None, marks=None, w_im=0.5,
             w_depth=1.0, ssim=False, isoflat=False, iso_seed=0, graph=False, eps=1e-15, ignore_outlier_depth=False,
             ba=None):
    """One mapping iteration on one keyframe (splatam.py:1412-1476): ONE host call (isogs_map_step) enqueues
    transform_to_frame + activations + projection, binning, the 5-channel blend, the masked-L1 (+ SSIM) loss, both
    backwards and the fused Adam; no host sync.  `pose` is the keyframe's HOST pose (camera_grad=False).
    `allreduce(flat_grad_tensor)` runs between backward and Adam, as do the iso / flat terms (`isoflat=True`, :572-675,
    three-axis scales) -- with ssim and isoflat the loss is the reference's complete mapping loss.  `marks` (a list)
    receives (stage, cuda event) pairs.  graph=True replays a captured CUDA graph of the step: one graph serves
    every keyframe, because the frame is read through the device-side slot ws.cur_frame.
    ignore_outlier_depth: the depth term over |err| < 10 median(err) only, mean over the masked pixels (:545-549).
    ba: a PoseState -- get_loss(do_ba=True) (:505-510): the keyframe's pose is optimised together with the map
    (learning rates lrs['cam_unnorm_rots'], lrs['cam_trans']); `pose` is then ignored."""
    if ba is not None:
        if allreduce is not None or isoflat or graph:
            raise ValueError("map_step(ba=...) runs the eager single-call path")
        ws.check_overflow()
        if ssim and ws.dmaps is None:
            ws.dmaps = torch.empty((9, ws.cam.height, ws.cam.width), dtype=torch.float32, device=ws.device)
        if ignore_outlier_depth:
            ws.median_temp()
        frame = ws.frame(gt_rgb, gt_depth, pose, inv_n_depth)
        cws, cgs = ws.c_struct(), gmap.c_struct()
        cws.adam_state = gmap.adam_state.data_ptr()
        flags = (_lib.STEP_SSIM if ssim else 0) | (_lib.STEP_OUTLIER if ignore_outlier_depth else 0)
        prm = _params(flags, _lib.PHASE_RENDER | _lib.PHASE_ADAM, w_im, w_depth, 0.99, [lrs.get(k, 0.0) for k in PARAM_ORDER],
                      eps, lr_pose=(lrs.get("cam_unnorm_rots", 0.0), lrs.get("cam_trans", 0.0)))
        gmap.step_count += 1
        ba.step_count += 1
        check(lib().isogs_map_step_ba(C.byref(ws.cam), frame.data_ptr(), C.byref(cws), C.byref(cgs), ba.pose.data_ptr(),
                                      ba.m.data_ptr(), ba.v.data_ptr(), ba.adam_state.data_ptr(), C.byref(prm),
                                      _event_array(marks), _stream()))
        ws.post_totals()
        return ws.losses
    if ssim and ws.dmaps is None:
        ws.dmaps = torch.empty((9, ws.cam.height, ws.cam.width), dtype=torch.float32, device=ws.device)
    if ignore_outlier_depth:
        ws.median_temp()
    ws.check_overflow()
    L, st = lib(), _stream()
    frame = ws.frame(gt_rgb, gt_depth, pose, inv_n_depth)
    cws, cgs = ws.c_struct(), gmap.c_struct()
    cws.adam_state = gmap.adam_state.data_ptr()
    flags = (_lib.STEP_SSIM if ssim else 0) | (_lib.STEP_OUTLIER if ignore_outlier_depth else 0)
    lr5 = [lrs.get(k, 0.0) for k in PARAM_ORDER]
    split = allreduce is not None or isoflat
    ev = _event_array(marks)
    gmap.step_count += 1
    if graph and marks is None:
        ws.cur_frame.copy_(frame, non_blocking=True)
        fr = ws.cur_frame.data_ptr()
        key = ("map", flags, split, id(gmap), gmap.count, gmap.grad_flat.data_ptr(), tuple(lr5), w_im, w_depth, eps)
        p1 = _params(flags, _lib.PHASE_RENDER if split else _lib.PHASE_RENDER | _lib.PHASE_ADAM, w_im, w_depth, 0.99, lr5, eps)
        _graph_launch(ws, key, lambda cs: check(L.isogs_map_step(C.byref(ws.cam), fr, C.byref(cws), C.byref(cgs),
                                                                 C.byref(p1), None, cs)))
    else:
        p1 = _params(flags, _lib.PHASE_RENDER if split else _lib.PHASE_RENDER | _lib.PHASE_ADAM, w_im, w_depth, 0.99, lr5, eps)
        check(L.isogs_map_step(C.byref(ws.cam), frame.data_ptr(), C.byref(cws), C.byref(cgs), C.byref(p1), ev, st))
    if split:
        if isoflat:
            iso_flat_loss(ws, gmap, seed=iso_seed)
            _mark(marks, "isoflat")
        if allreduce is not None and hasattr(allreduce, "fused_adam"):
            # exchange fused with the optimizer step (dp.PeerReduceAdam): no separate Adam launch
            allreduce.fused_adam(gmap, lrs, eps)
            if marks is not None:                  # this path has no separate Adam launch: drop its (unrecorded) mark
                marks[:] = [m for m in marks if m[0] != "adam"]
            _mark(marks, "allreduce")
            ws.post_totals()
            return ws.losses
        if allreduce is not None:
            allreduce(gmap.grad_flat[:gmap.grad_used])
            _mark(marks, "allreduce")
        p2 = _params(flags, _lib.PHASE_ADAM, w_im, w_depth, 0.99, lr5, eps)
        if graph and marks is None:
            key2 = ("adam", id(gmap), gmap.count, gmap.grad_flat.data_ptr(), tuple(lr5), eps)
            _graph_launch(ws, key2, lambda cs: check(L.isogs_map_step(C.byref(ws.cam), ws.cur_frame.data_ptr(),
                                                                      C.byref(cws), C.byref(cgs), C.byref(p2), None, cs)))
        else:
            check(L.isogs_map_step(C.byref(ws.cam), frame.data_ptr(), C.byref(cws), C.byref(cgs), C.byref(p2), ev, st))
            if marks is not None:                  # keep the stage order: the Adam mark goes last
                k = next(i for i, m in enumerate(marks) if m[0] == "adam")
                marks.append(marks.pop(k))
    ws.post_totals()
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
    scratch, tmp, tmp_b = ws.median_temp()
    check(lib().isogs_depth_outlier_threshold(C.byref(cam), _ptr(ws.render[3]), _ptr(gt_depth), 50.0,
                                              _ptr(scratch), _ptr(tmp), tmp_b, _ptr(ws.small[28:29]), _stream()))
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


def accumulate_densify_stats(ws, gmap, v_xy=None, scale=None):
    """accumulate_mean2d_gradient + the max_2D_radius update of get_loss (slam_external.py:100-104,
    splatam.py:751-753) for the Gaussians seen by the last render.  ONE convention everywhere: the
    accumulated number is the norm of `means2D.grad` as the Inria rasterizer writes it, i.e. the screen-space
    gradient in NDC units v_xy * (W/2, H/2) -- the unit the reference's grad_thresh = 2e-4 was tuned on
    (configs/replica/splatam.py:135).  v_xy=None takes the pixel-unit gradient of the last backward (ws.v_rec) and
    applies that scaling; an explicit v_xy is a `means2D.grad` tensor (already scaled, e.g. from the drop-in
    GaussianRasterizer) and is accumulated as-is.  NOTE: the reference accumulates the gradient of the COLOUR render
    only (splatam.py:528-530); the fused 5-channel step's v_rec also carries the depth-loss part -- use
    `colour_only_screen_gradient` for the reference's exact statistic."""
    if scale is None:
        scale = (0.5 * ws.cam.width, 0.5 * ws.cam.height) if v_xy is None else (1.0, 1.0)
    v = ws.v_rec if v_xy is None else v_xy
    check(lib().isogs_densify_stats(gmap.count, _ptr(ws.radii), _ptr(v), v.shape[1], scale[0], scale[1],
                                    _ptr(gmap.vars["means2D_gradient_accum"]), _ptr(gmap.vars["denom"]),
                                    _ptr(gmap.vars["max_2D_radius"]), _stream()))


def colour_only_screen_gradient(ws, gmap):
    """Second blend backward with the depth planes of v_render zeroed: dL_colour/d(mean2D) exactly as the
    reference's separate colour render sees it, in PIXEL units (v_rec rows): pass it as
    accumulate_densify_stats(ws, gmap, v_xy=..., scale=(W/2, H/2)).  Costs one extra blend_bwd."""
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
        self.adam_state = torch.zeros((4,), dtype=torch.float64, device=device)   # isogs_adam_state of the pose optimizer
        self._stage = torch.zeros((15,), dtype=torch.float32).pin_memory()

    def reset(self, q, t):
        """Start the next frame IN PLACE (same device buffers, so a captured tracking graph stays valid): the
        reference re-creates the pose optimizer per frame (splatam.py:1237)."""
        self._stage[:7] = torch.tensor(list(q) + list(t), dtype=torch.float32)
        self._stage[7:14] = self._stage[:7]
        self._stage[14] = 1e20
        self.pose.copy_(self._stage[:7], non_blocking=True)
        self.best.copy_(self._stage[7:15], non_blocking=True)
        self.m.zero_()
        self.v.zero_()
        self.adam_state.zero_()
        self.step_count = 0
        return self

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
               ignore_outlier_depth=False, graph=False, marks=None):
    """One tracking iteration (splatam.py:1247-1290) as ONE host call (isogs_track_step): only the pose gets a
    gradient and an Adam step; the best candidate is kept on the device.  graph=True replays a captured graph (the
    same graph serves every frame as long as `ps` and the ground-truth buffers are reused in place)."""
    if ignore_outlier_depth:
        ws.median_temp()
    ws.check_overflow()
    L, st = lib(), _stream()
    frame = ws.frame(gt_rgb, gt_depth)
    cws, cgs = ws.c_struct(), gmap.c_struct(with_grads=False)
    cws.adam_state = None
    flags = (_lib.STEP_OUTLIER if ignore_outlier_depth else 0) | (0 if use_sil else _lib.STEP_NO_SIL)
    prm = _params(flags, _lib.PHASE_RENDER | _lib.PHASE_ADAM, 0.5, 1.0, sil_thres,
                  [lrs["cam_unnorm_rots"], lrs["cam_trans"], 0.0, 0.0, 0.0], eps)
    ps.step_count += 1
    args = (_ptr(ps.pose), _ptr(ps.m), _ptr(ps.v), _ptr(ps.best), _ptr(ps.adam_state))
    if graph and marks is None:
        ws.cur_frame.copy_(frame, non_blocking=True)
        fr = ws.cur_frame.data_ptr()
        key = ("track", flags, id(gmap), gmap.count, ps.pose.data_ptr(), float(sil_thres), tuple(prm.lr), eps)
        _graph_launch(ws, key, lambda cs: check(L.isogs_track_step(C.byref(ws.cam), fr, C.byref(cws), C.byref(cgs), *args,
                                                                   C.byref(prm), None, cs)))
    else:
        check(L.isogs_track_step(C.byref(ws.cam), frame.data_ptr(), C.byref(cws), C.byref(cgs), *args, C.byref(prm),
                                 _event_array(marks), st))
    ws.post_totals()
    return ws.losses
