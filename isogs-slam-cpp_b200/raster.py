"""Functional front-end of the C ABI on torch CUDA tensors (pointer plumbing only).

forward()  = preprocess -> duplicateWithKeys/sort/identifyTileRanges -> renderCUDA forward
backward() = renderCUDA backward -> preprocess backward (+ pose chain)
Mirrors the call order of cpp/src/rendering/renderer.cpp:84-232 and gsplat/rendering.py.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import REC, check, lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk(name, t, shape_last=None, dtype=torch.float32):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU fallback)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if shape_last is not None and (t.dim() < 1 or t.shape[-1] != shape_last):
        raise RuntimeError(f"{name} must have last dim {shape_last}, got {tuple(t.shape)}")


@dataclass
class RasterCtx:
    """Everything the backward needs (saved by the autograd Function)."""
    cam: _lib.Camera
    ch: int
    flags: int
    N: int
    M: int
    scale_dim: int
    rec: torch.Tensor
    radii: torch.Tensor
    vals: torch.Tensor
    ranges: torch.Tensor
    alpha: torch.Tensor
    last: torch.Tensor
    pose: Optional[_lib.Pose] = None
    pose_dev: Optional[torch.Tensor] = None
    keys: Optional[torch.Tensor] = None
    tiles: Optional[torch.Tensor] = None
    keys_unsorted: Optional[torch.Tensor] = None
    vals_unsorted: Optional[torch.Tensor] = None


def preprocess(cam, means, scales, quats, opacities, colors, ch=4, flags=0, pose=None, pose_dev=None):
    """north_star (1).  Returns rec [N,12], radii [N,2] i32, tiles [N] i32, cum [N] i32, M (host int)."""
    N = means.shape[0]
    scale_dim = 1 if (scales.dim() == 1 or scales.shape[-1] == 1) else 3
    _chk("means", means, 3)
    _chk("scales", scales)
    _chk("quats", quats, 4)
    _chk("opacities", opacities)
    _chk("colors", colors, 3)
    if scale_dim == 3 and quats is None:
        raise RuntimeError("anisotropic scales [N,3] need rotations [N,4]")
    dev = means.device
    rec = torch.empty((N, REC), dtype=torch.float32, device=dev)
    radii = torch.empty((N, 2), dtype=torch.int32, device=dev)
    tiles = torch.empty((N,), dtype=torch.int32, device=dev)
    cum = torch.empty((N,), dtype=torch.int32, device=dev)
    scan_b, _ = _lib.temp_bytes(N, 0)
    scan_tmp = torch.empty((scan_b,), dtype=torch.uint8, device=dev)
    m = C.c_int64(0)
    check(lib().isogs_preprocess_fwd(C.byref(cam), C.byref(pose) if pose is not None else None, _ptr(pose_dev), N,
                                     _ptr(means), _ptr(scales), scale_dim, _ptr(quats), _ptr(opacities), _ptr(colors),
                                     ch, flags, _ptr(rec), _ptr(radii), _ptr(tiles), _ptr(cum), _ptr(scan_tmp),
                                     scan_b, C.byref(m), None, None, None, _stream()))
    return rec, radii, tiles, cum, int(m.value)


def tile_bin(cam, rec, radii, cum, M, literal=False):
    """north_star (2).  Returns keys_sorted u64-as-int64 [M], vals_sorted i32 [M], ranges i32 [n_tiles+1],
    plus the unsorted pair (for the bit-exact tests).  literal=True runs the emit-keys + CUB radix sort + ranges
    pipeline (isogs_tile_bin_cub) instead of the counting sort by tile."""
    N = rec.shape[0]
    dev = rec.device
    ku = torch.empty((M,), dtype=torch.int64, device=dev)
    vu = torch.empty((M,), dtype=torch.int32, device=dev)
    ks = torch.empty((M,), dtype=torch.int64, device=dev)
    vs = torch.empty((M,), dtype=torch.int32, device=dev)
    ranges = torch.empty((cam.n_tiles + 1,), dtype=torch.int32, device=dev)
    _, sort_b = _lib.temp_bytes(0, M)
    tmp = torch.empty((sort_b,), dtype=torch.uint8, device=dev)
    fn = lib().isogs_tile_bin_cub if literal else lib().isogs_tile_bin
    check(fn(C.byref(cam), N, M, _ptr(rec), _ptr(radii), _ptr(cum), _ptr(ku), _ptr(vu), _ptr(ks),
                               _ptr(vs), _ptr(ranges), _ptr(tmp), sort_b, _stream()))
    return ks, vs, ranges, ku, vu


def blend_forward(cam, ch, rec, vals, ranges):
    """north_star (3).  render [ch,H,W], alpha [H,W], last [H,W] i32."""
    dev = rec.device
    H, W = cam.height, cam.width
    render = torch.empty((ch, H, W), dtype=torch.float32, device=dev)
    alpha = torch.empty((H, W), dtype=torch.float32, device=dev)
    last = torch.empty((H, W), dtype=torch.int32, device=dev)
    check(lib().isogs_blend_fwd(C.byref(cam), ch, _ptr(rec), _ptr(vals), _ptr(ranges), _ptr(render), _ptr(alpha),
                                _ptr(last), _stream()))
    return render, alpha, last


def forward(cam, means, scales, quats, opacities, colors, ch=4, flags=0, pose=None, pose_dev=None, keep_keys=False):
    rec, radii, tiles, cum, M = preprocess(cam, means, scales, quats, opacities, colors, ch, flags, pose, pose_dev)
    ks, vs, ranges, ku, vu = tile_bin(cam, rec, radii, cum, M)
    render, alpha, last = blend_forward(cam, ch, rec, vs, ranges)
    scale_dim = 1 if (scales.dim() == 1 or scales.shape[-1] == 1) else 3
    ctx = RasterCtx(cam=cam, ch=ch, flags=flags, N=means.shape[0], M=M, scale_dim=scale_dim, rec=rec, radii=radii,
                    vals=vs, ranges=ranges, alpha=alpha, last=last, pose=pose, pose_dev=pose_dev)
    if keep_keys:
        ctx.keys, ctx.tiles, ctx.keys_unsorted, ctx.vals_unsorted = ks, tiles, ku, vu
    return render, alpha, ctx


def blend_backward(ctx, v_render, v_alpha=None):
    """north_star (4), pixel part.  Returns v_rec [N,12]."""
    _chk("v_render", v_render)
    _chk("v_alpha", v_alpha)
    v_rec = torch.zeros((ctx.N, REC), dtype=torch.float32, device=ctx.rec.device)
    if ctx.M > 0:
        check(lib().isogs_blend_bwd(C.byref(ctx.cam), ctx.ch, ctx.N, _ptr(ctx.rec), _ptr(ctx.vals), _ptr(ctx.ranges),
                                    _ptr(ctx.alpha), _ptr(ctx.last), _ptr(v_render), _ptr(v_alpha), _ptr(v_rec),
                                    _stream()))
    return v_rec


def preprocess_backward(ctx, v_rec, means, scales, quats, opacities, want_means=True, want_scales=True,
                        want_quats=True, want_opac=True, want_colors=True, want_means2d=False, want_pose=False):
    """north_star (4), Gaussian part + pose chain.  Returns a dict of gradients (None when not wanted)."""
    dev = means.device
    N = ctx.N
    sd = ctx.scale_dim

    def mk(want, shape):
        return torch.empty(shape, dtype=torch.float32, device=dev) if want else None
    g = dict(means=mk(want_means, (N, 3)), scales=mk(want_scales, (N, sd)),
             quats=mk(want_quats and sd == 3, (N, 4)), opacities=mk(want_opac, (N,)),
             colors=mk(want_colors, (N, 3)), means2d=mk(want_means2d, (N, 3)), pose=None)
    pose_acc = None
    if want_pose:
        if ctx.pose is None and ctx.pose_dev is None:
            raise RuntimeError("pose gradient requested but the forward had no pose")
        g["pose"] = torch.zeros((7,), dtype=torch.float32, device=dev)
        pose_acc = torch.zeros((16,), dtype=torch.float32, device=dev)
    check(lib().isogs_preprocess_bwd(C.byref(ctx.cam), C.byref(ctx.pose) if ctx.pose is not None else None,
                                     _ptr(ctx.pose_dev), N, _ptr(means), _ptr(scales), sd, _ptr(quats),
                                     _ptr(opacities), ctx.ch, ctx.flags, _ptr(ctx.radii), _ptr(v_rec),
                                     _ptr(g["means"]), _ptr(g["scales"]), _ptr(g["quats"]), _ptr(g["opacities"]),
                                     _ptr(g["colors"]), _ptr(g["means2d"]), _ptr(g["pose"]), _ptr(pose_acc),
                                     _stream()))
    return g
