"""Drop-in for `diff_gaussian_rasterization` (the module python_ordinary/scripts/splatam.py:39 and
python_ordinary/utils/recon_helpers.py:2 import): same two names, same keyword arguments, same
3-tuple return (color [3,H,W], radii [N] int32, depth [1,H,W]).

Semantics are gsplat's (the only rasterizer source in the reference tree):
alpha clamp 0.999, 1/255 threshold, opacity-aware 3.33-sigma extent, eps2d 0.3, near 0.01.
The third output is the un-normalised alpha-weighted depth (gsplat mode "D").

Extension: `scales` of shape [N,1] selects the isotropic kernel specialisation (rotations may
then be None); [N,3] is the general path the reference Python loop uses.
"""
import collections
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _lib, raster


class GaussianRasterizationSettings(NamedTuple):
    # field order of python_ordinary/utils/recon_helpers.py:14-26
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool = False


_CAM_CACHE = collections.OrderedDict()
_CAM_CACHE_MAX = 64


def camera_from_settings(s: GaussianRasterizationSettings) -> _lib.Camera:
    """Recover pinhole intrinsics + view matrix from the Inria-style settings
    (setup_camera, recon_helpers.py:4-27).  One device->host read per distinct settings object;
    cached (the cache keeps the tensors alive so data_ptr keys cannot be recycled)."""
    vm, pm, bg = s.viewmatrix, s.projmatrix, s.bg
    key = (s.image_height, s.image_width, float(s.tanfovx), float(s.tanfovy), vm.data_ptr(), vm._version,
           pm.data_ptr(), pm._version, bg.data_ptr(), bg._version)
    hit = _CAM_CACHE.get(key)
    if hit is not None:
        _CAM_CACHE.move_to_end(key)
        return hit[0]
    W, H = int(s.image_width), int(s.image_height)
    w2c = vm.detach().reshape(4, 4).t().double().cpu()      # viewmatrix is w2c^T
    full = pm.detach().reshape(4, 4).t().double().cpu()     # projmatrix^T = P @ w2c
    P = full @ torch.linalg.inv(w2c)
    fx = W / (2.0 * float(s.tanfovx))
    fy = H / (2.0 * float(s.tanfovy))
    cx = W * (1.0 + float(P[0, 2])) / 2.0                   # P[0][2] = -(w - 2cx)/w
    cy = H * (1.0 + float(P[1, 2])) / 2.0
    # near/far from the z rows of P: P[2][2] = f/(f-n), P[2][3] = -fn/(f-n)
    a, b = float(P[2, 2]), float(P[2, 3])
    near = -b / a if abs(a) > 1e-12 else 0.01
    far = -b / (a - 1.0) if abs(a - 1.0) > 1e-12 else 100.0
    if not (0.0 < near < far):
        near, far = 0.01, 100.0
    cam = _lib.make_camera(W, H, fx, fy, cx, cy, w2c.float().tolist(), near=near, far=far,
                           bg=[float(v) for v in bg.detach().float().cpu().reshape(-1)[:3]])
    _CAM_CACHE[key] = (cam, vm, pm, bg)
    if len(_CAM_CACHE) > _CAM_CACHE_MAX:
        _CAM_CACHE.popitem(last=False)
    return cam


class _Rasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, colors, opacities, scales, rotations, cam):
        means3D = means3D.contiguous()
        colors = colors.contiguous()
        opac = opacities.reshape(-1).contiguous()
        scales = scales.contiguous()
        rotations = None if rotations is None else rotations.contiguous()
        render, alpha, rctx = raster.forward(cam, means3D, scales, rotations, opac, colors, ch=4, flags=0)
        ctx.rctx = rctx
        ctx.opac_shape = opacities.shape
        ctx.save_for_backward(means3D, scales, rotations if rotations is not None else torch.empty(0), opac)
        color = render[:3]
        depth = render[3:4]
        radii = rctx.radii.max(dim=1).values
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, v_color, v_radii, v_depth, v_alpha):
        means3D, scales, rotations, opac = ctx.saved_tensors
        if rotations.numel() == 0:
            rotations = None
        rctx = ctx.rctx
        H, W = rctx.cam.height, rctx.cam.width
        v_render = torch.empty((4, H, W), dtype=torch.float32, device=means3D.device)
        v_render[:3] = v_color if v_color is not None else 0
        v_render[3:4] = v_depth if v_depth is not None else 0
        v_a = None if v_alpha is None else v_alpha.contiguous()
        v_rec = raster.blend_backward(rctx, v_render, v_a)
        need = ctx.needs_input_grad  # means3D, means2D, colors, opacities, scales, rotations, cam
        g = raster.preprocess_backward(rctx, v_rec, means3D, scales, rotations, opac,
                                       want_means=need[0], want_means2d=need[1], want_colors=need[2],
                                       want_opac=need[3], want_scales=need[4], want_quats=need[5])
        g_opac = g["opacities"].reshape(ctx.opac_shape) if g["opacities"] is not None else None
        return g["means"], g["means2d"], g["colors"], g_opac, g["scales"], g["quats"], None


class GaussianRasterizer(nn.Module):
    """`GaussianRasterizer(raster_settings=cam)(means3D=..., means2D=..., colors_precomp=..., rotations=...,
    opacities=..., scales=...)` -- keyword names fixed by python_ordinary/utils/slam_helpers.py:131-138."""

    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, return_alpha=False):
        if shs is not None or cov3D_precomp is not None:
            raise NotImplementedError("shs / cov3D_precomp are not on the IsoGS-SLAM hot path (sh_degree=0)")
        if colors_precomp is None or scales is None:
            raise ValueError("colors_precomp and scales are required")
        s = self.raster_settings
        cam = camera_from_settings(s)
        if s.scale_modifier != 1.0:
            scales = scales * s.scale_modifier
        if means2D is None:
            means2D = torch.zeros_like(means3D)
        color, radii, depth, alpha = _Rasterize.apply(means3D, means2D, colors_precomp, opacities, scales, rotations,
                                                      cam)
        if return_alpha:
            return color, radii, depth, alpha
        return color, radii, depth
