"""GPU parity against the REFERENCE's own CUDA kernels: vendored gsplat 1.5.3
(cpp/third_party/gsplat/gsplat/cuda/csrc/RasterizeToPixels3DGS{Fwd,Bwd}.cu) compiled in place for
sm_100a by oracle/build_ref.sh into oracle/_ref/gsplat_ref.so.  This pins the blend forward /
backward -- the one stage with no CPU-runnable reference -- to the reference itself, on the same
projected records, tile offsets and sorted ids.  Skipped when the library was not built
(/root/reference absent at build time)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "gsplat_ref.so")


@pytest.fixture(scope="module")
def gsref():
    if not os.path.exists(SO):
        pytest.skip("oracle/_ref/gsplat_ref.so not built")
    spec = importlib.util.spec_from_file_location("gsplat_ref", SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def t(a, dtype=torch.float32):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype)


def gsplat_inputs(ctx, ch):
    rec = ctx.rec
    means2d = rec[:, 0:2].contiguous()[None]
    conics = torch.stack([2.0 * rec[:, 2], rec[:, 3], 2.0 * rec[:, 4]], -1).contiguous()[None]
    opac = rec[:, 6].contiguous()[None]
    colors = rec[:, 7:7 + ch].contiguous()[None]
    th, tw = (ctx.cam.height + 15) // 16, (ctx.cam.width + 15) // 16
    offsets = ctx.ranges[:-1].reshape(1, th, tw).contiguous()
    return means2d, conics, colors, opac, offsets, ctx.vals.contiguous()


def scene(pkg, syn, kind, iso):
    if kind == "splatam":
        params, frames, K = syn.make_map("tiny")
        W, H, fx, fy, cx, cy = syn.SHAPES["tiny"][:6]
        flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
        q, tr = syn.w2c_to_pose(syn.trajectory(4, step=3.0)[3])
        return (pkg.make_camera(W, H, fx, fy, cx, cy), t(params["means3D"]), t(params["log_scales"]), None,
                t(params["logit_opacities"].reshape(-1)), t(params["rgb_colors"]), flags, pkg.make_pose(q, tr))
    rng = np.random.RandomState(4)
    n, W, H, f = 6000, 208, 144, 170.0
    z = 0.8 + 3.5 * rng.rand(n)
    means = np.stack([(rng.rand(n) * W - W / 2) / f * z, (rng.rand(n) * H - H / 2) / f * z, z], -1).astype(np.float32)
    s = (z / f * (0.5 + 4.0 * rng.rand(n))).astype(np.float32)
    scales = s[:, None] if iso else (s[:, None] * (0.3 + rng.rand(n, 3))).astype(np.float32)
    quats = None if iso else rng.randn(n, 4).astype(np.float32)
    return (pkg.make_camera(W, H, f, f, W / 2 - 0.5, H / 2 - 0.5), t(means), t(scales), t(quats),
            t(rng.rand(n).astype(np.float32)), t(rng.rand(n, 3).astype(np.float32)), 0, None)


@pytest.mark.parametrize("kind,iso,ch", [("splatam", True, 5), ("random", True, 4), ("random", False, 5)])
def test_blend_matches_reference_kernels(pkg, syn, gsref, kind, iso, ch):
    cam, means, scales, quats, opac, cols, flags, pose = scene(pkg, syn, kind, iso)
    render, alpha, ctx = pkg.raster.forward(cam, means, scales, quats, opac, cols, ch=ch, flags=flags, pose=pose)
    m2d, conics, colors, op, offsets, ids = gsplat_inputs(ctx, ch)
    W, H = cam.width, cam.height
    r_ref, a_ref, l_ref = gsref.rast_fwd(m2d, conics, colors, op, W, H, offsets, ids)
    torch.cuda.synchronize()
    r_ref = r_ref[0].permute(2, 0, 1)
    scale = r_ref.abs().amax(dim=(1, 2)).clamp(min=1.0)[:, None, None]
    err = ((render - r_ref).abs() / scale).amax(0)
    # the reference tests alpha < 1/255 per pixel (fast-math), we test sigma > ln(255 o) (the same
    # inequality): a pair sitting exactly on the threshold may flip -> allow a handful of pixels
    bad = int((err > 1e-4).sum())
    assert bad <= max(2, int(2e-4 * W * H)), f"{bad} pixels differ from the reference kernel by more than 1e-4"
    assert float(err.max()) <= 6e-3
    assert float((alpha - a_ref[0, :, :, 0]).abs().max()) <= 6e-3
    hit = alpha > 0
    same_last = (ctx.last == l_ref[0])[hit].float().mean().item()
    assert same_last >= 0.999, same_last
    # backward on the reference's own forward state vs ours on ours
    g = torch.Generator(device="cuda").manual_seed(1)
    v_render = torch.randn((ch, H, W), device="cuda", generator=g)
    v_alpha = torch.randn((H, W), device="cuda", generator=g)
    v_rec = pkg.raster.blend_backward(ctx, v_render, v_alpha)
    vm, vc, vcol, vo = gsref.rast_bwd(m2d, conics, colors, op, W, H, offsets, ids, a_ref, l_ref,
                                      v_render.permute(1, 2, 0).contiguous()[None], v_alpha[None, :, :, None].contiguous())
    torch.cuda.synchronize()

    def rel(a, b):
        return float((a - b).abs().max() / (b.abs().max() + 1e-30))
    assert rel(v_rec[:, 0:2], vm[0]) <= 2e-3, "v_means2d"
    assert rel(v_rec[:, 2:5], vc[0]) <= 2e-3, "v_conics"
    assert rel(v_rec[:, 5], vo[0]) <= 2e-3, "v_opacities"
    assert rel(v_rec[:, 6:6 + ch], vcol[0]) <= 2e-3, "v_colors"


# ---------------------------------------------------------------------------------------------
# projection (WITH the opacity-aware extent) and tile intersection against the reference's kernels
# ---------------------------------------------------------------------------------------------
def _quat_R(q):
    q = q / np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]], np.float32)


def _proj_scene(iso, n=20000, seed=12):
    rng = np.random.RandomState(seed)
    W, H, f = 320, 208, 260.0
    cx, cy = W / 2 - 0.5, H / 2 - 0.5
    z = 0.6 + 4.0 * rng.rand(n)
    u, v = rng.rand(n) * (W + 60) - 30, rng.rand(n) * (H + 60) - 30     # some off-image, some straddling the border
    means = np.stack([(u - cx) / f * z, (v - cy) / f * z, z], -1).astype(np.float32)
    means[: n // 60, 2] = rng.rand(n // 60) * 0.02 - 0.005               # behind / at the near plane
    s = (np.abs(z) / f * (0.5 + 5.0 * rng.rand(n))).astype(np.float32)
    scales = np.repeat(s[:, None], 3, 1) if iso else (s[:, None] * (0.3 + rng.rand(n, 3))).astype(np.float32)
    quats = rng.randn(n, 4).astype(np.float32)
    quats /= np.linalg.norm(quats, axis=1, keepdims=True)
    opac = rng.rand(n).astype(np.float32)
    opac[: n // 30] = 0.002 + 0.004 * rng.rand(n // 30)                   # around the 1/255 cull threshold
    vm = np.eye(4, dtype=np.float32)
    vm[:3, :3] = _quat_R(np.array([0.995, 0.03, -0.05, 0.02]))
    vm[:3, 3] = [0.04, -0.03, 0.06]
    return dict(W=W, H=H, f=f, cx=cx, cy=cy, means=means, scales=scales, quats=quats, opac=opac, vm=vm)


@pytest.mark.parametrize("iso", [True, False])
def test_projection_matches_reference_kernel_with_opacity_extent(pkg, gsref, iso):
    """isogs_preprocess_fwd with its DEFAULT flags -- the opacity-aware extent min(3.33, sqrt(2 ln(255 o))) of
    ProjectionEWA3DGSFused.cu:164-184 included -- against projection_ewa_3dgs_fused_fwd_kernel itself
    (launcher of ProjectionEWA3DGSFused.cu:214-291, opacities passed).  Tolerances of the vendored gsplat's own
    test_projection (tests/test_basic.py:167-268): means2d / depth / conic 1e-4 relative-ish; the int radii
    equal except where the reference's fast-math __logf / rsqrt moves a ceil() across an integer."""
    sc = _proj_scene(iso)
    W, H = sc["W"], sc["H"]
    cam = pkg.make_camera(W, H, sc["f"], sc["f"], sc["cx"], sc["cy"], sc["vm"])
    K = torch.tensor([[sc["f"], 0, sc["cx"]], [0, sc["f"], sc["cy"]], [0, 0, 1]], dtype=torch.float32, device="cuda")
    means, quats, opac = t(sc["means"]), t(sc["quats"]), t(sc["opac"])
    scales3 = t(sc["scales"])
    r_ref, m_ref, d_ref, c_ref = gsref.proj_fwd(means, quats, scales3, opac, t(sc["vm"]), K, W, H, 0.3, 0.01, 100.0)
    ours_scales = scales3[:, :1].contiguous() if iso else scales3
    rec, radii, tiles, cum, M = pkg.raster.preprocess(cam, means, ours_scales, None if iso else quats, opac,
                                                      torch.zeros_like(means), ch=4, flags=0)
    torch.cuda.synchronize()
    r_ref, radii = r_ref.cpu().numpy(), radii.cpu().numpy()
    vis_ref, vis = (r_ref > 0).all(1), (radii > 0).all(1)
    n = len(vis)
    assert (vis_ref != vis).sum() <= max(2, int(1e-3 * n)), "visibility differs beyond threshold-level flips"
    both = vis_ref & vis
    assert both.sum() > 0.5 * n
    mism = (r_ref[both] != radii[both]).any(1)
    assert mism.mean() <= 1e-3, f"radii differ on {mism.mean():.2%} of the visible Gaussians"
    assert np.abs(r_ref[both] - radii[both]).max() <= 1
    rec = rec.cpu().numpy()[both]
    m_ref, d_ref, c_ref = m_ref.cpu().numpy()[both], d_ref.cpu().numpy()[both], c_ref.cpu().numpy()[both]
    assert np.abs(rec[:, 0:2] - m_ref).max() <= 2e-3, "means2d (pixels)"
    assert np.abs(rec[:, 10] - d_ref).max() <= 1e-5 * np.abs(d_ref).max()
    conic = np.stack([2 * rec[:, 2], rec[:, 3], 2 * rec[:, 4]], -1)
    assert np.abs(conic - c_ref).max() <= 1e-4 * max(1.0, np.abs(c_ref).max())
    # the extent really is the opacity-aware one: with the flag that disables it the radii of faint splats grow
    rec2, radii2, _, _, _ = pkg.raster.preprocess(cam, means, ours_scales, None if iso else quats, opac,
                                                  torch.zeros_like(means), ch=4, flags=pkg._lib.F_NO_OPAC_EXTENT)
    assert (radii2.cpu().numpy()[both] > radii[both]).any()


def test_tile_keys_match_reference_intersect_exactly(pkg, gsref):
    """isogs_tile_bin against the reference's own intersect_tile (two kernel passes + cumsum + CUB radix sort) and
    intersect_offset (Intersect.cpp:15-186, IntersectTile.cu:24-114,209-339) on IDENTICAL (means2d, radii, depths):
    tiles_per_gauss, sorted isect_ids, flatten_ids and offsets must be EQUAL -- the bar of gsplat's test_isect
    (tests/test_basic.py:451-489)."""
    sc = _proj_scene(True, n=30000, seed=5)
    W, H = sc["W"], sc["H"]
    cam = pkg.make_camera(W, H, sc["f"], sc["f"], sc["cx"], sc["cy"], sc["vm"])
    means, opac = t(sc["means"]), t(sc["opac"])
    rec, radii, tiles, cum, M = pkg.raster.preprocess(cam, means, t(sc["scales"][:, :1]), None, opac,
                                                      torch.zeros_like(means), ch=4, flags=0)
    ks, vs, ranges, ku, vu = pkg.raster.tile_bin(cam, rec, radii, cum, M)
    tw, th = (W + 15) // 16, (H + 15) // 16
    tpg, ids_ref, flat_ref, off_ref = gsref.isect(rec[:, 0:2].contiguous(), radii.contiguous(), rec[:, 10].contiguous(), tw, th)
    torch.cuda.synchronize()
    assert M == ids_ref.numel() > 3 * 10000
    assert torch.equal(tpg.to(torch.int32), tiles), "tiles_per_gauss"
    assert torch.equal(ids_ref, ks), "sorted (tile | depth) keys"
    assert torch.equal(flat_ref.to(torch.int32), vs), "sorted Gaussian ids"
    assert torch.equal(off_ref.reshape(-1).to(torch.int32), ranges[:-1]), "tile offsets"
    assert int(ranges[-1]) == M
