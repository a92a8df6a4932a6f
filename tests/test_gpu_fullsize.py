"""Full-size checks at BASELINE.json's shapes (configs 3 and 4: Replica-shaped 1200x680 with 1 M
Gaussians, ScanNet++-shaped 1752x1168 with 3 M), where the CPU oracle would take minutes: the
size-independent properties the domain offers -- sortedness of the (tile | depth) keys, consistency
of ranges / counts, sync-free == synchronised binning, image bounds, the colour-gradient checksum
(sum_g c_g . dL/dc_g == sum_p v_render . render, exact by linearity) and linearity of the backward."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [("replica", 1_000_000), ("scannetpp", 3_000_000)]


@pytest.mark.parametrize("shape,n_target", CASES, ids=[c[0] for c in CASES])
def test_fullsize_properties(pkg, syn, shape, n_target):
    params, frames, K = syn.make_map(shape, n_target, isotropic=True, seed=0, max_frames=8)
    W, H, fx, fy, cx, cy = syn.SHAPES[shape][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    means, ls = tt(params["means3D"]), tt(params["log_scales"])
    lo, rgb = tt(params["logit_opacities"].reshape(-1)), tt(params["rgb_colors"])
    N = means.shape[0]
    assert N >= 0.95 * n_target
    q, tr = syn.w2c_to_pose(syn.trajectory(8, step=4.0)[2])
    pose = pkg.make_pose(q, tr)
    render, alpha, ctx = pkg.raster.forward(cam, means, ls, None, lo, rgb, ch=5, flags=flags, pose=pose, keep_keys=True)
    torch.cuda.synchronize()
    M, n_tiles = ctx.M, cam.n_tiles
    assert M > N  # every visible splat touches >= 1 tile, most touch several
    # --- integer work: bit-exact structural properties -------------------------------------
    keys = ctx.keys
    assert bool((keys[1:] >= keys[:-1]).all()), "(tile | depth) keys are not sorted"
    tile_of = (keys >> 32).to(torch.int32)
    assert int(tile_of.min()) >= 0 and int(tile_of.max()) < n_tiles
    counts = torch.bincount(tile_of.long(), minlength=n_tiles)
    ranges = ctx.ranges.long()
    assert int(ranges[0]) == 0 and int(ranges[-1]) == M
    assert torch.equal(ranges[1:] - ranges[:-1], counts), "tile ranges disagree with the sorted keys"
    assert int(ctx.tiles.sum()) == M, "per-Gaussian tile counts do not add up to n_isects"
    assert int(ctx.vals.min()) >= 0 and int(ctx.vals.max()) < N
    # every (tile, gaussian) pair appears exactly once: the multiset of ids matches the tile counts
    assert torch.equal(torch.bincount(ctx.vals.long(), minlength=N), ctx.tiles.long())
    # depth bits of the key are the record's depth
    assert torch.equal((keys & 0xFFFFFFFF).to(torch.int32), ctx.rec[ctx.vals.long(), 10].view(torch.int32))
    # sync-free binning produces the identical lists
    import ctypes as C
    bucket = torch.empty((M + 7,), dtype=torch.int64, device="cuda")
    vs2 = torch.empty((M + 7,), dtype=torch.int32, device="cuda")
    rg2 = torch.empty((n_tiles + 1,), dtype=torch.int32, device="cuda")
    tot = torch.zeros((2,), dtype=torch.int32, device="cuda")
    _, sort_b = pkg._lib.temp_bytes(0, M + 7)
    tmp = torch.empty((sort_b,), dtype=torch.uint8, device="cuda")
    p = lambda x: C.c_void_p(x.data_ptr())
    pkg._lib.check(pkg.lib().isogs_tile_bin_async(C.byref(cam), N, M + 7, p(ctx.rec), p(ctx.radii), p(bucket), None,
                                                  p(vs2), p(rg2), p(tot), p(tmp), sort_b, None, 0,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert tot.tolist() == [M, 0] and torch.equal(rg2, ctx.ranges) and torch.equal(vs2[:M], ctx.vals)
    # --- images -------------------------------------------------------------------------------
    assert bool(torch.isfinite(render).all()) and bool(torch.isfinite(alpha).all())
    assert float(alpha.min()) >= 0.0 and float(alpha.max()) <= 1.0
    assert float(alpha.mean()) > 0.9, "the map should cover the frame"
    assert float(render[:3].min()) >= 0.0 and float(render[:3].max()) <= 1.0 + 1e-4
    tile_id = (torch.arange(H, device="cuda")[:, None] // 16) * ((W + 15) // 16) + torch.arange(W, device="cuda")[None] // 16
    start, end = ctx.ranges[tile_id.long()], ctx.ranges[tile_id.long() + 1]
    assert bool(((ctx.last >= start - 1) & (ctx.last < end)).all()), "last-contributor index outside its tile's list"
    # depth^2 >= depth^2/alpha-ish (Cauchy-Schwarz on the blending weights): sum w z^2 * sum w >= (sum w z)^2
    assert bool((render[4] * alpha + 1e-4 >= render[3] ** 2 * (1 - 1e-4)).all())
    # --- backward: checksum + linearity ---------------------------------------------------------
    g = torch.Generator(device="cuda").manual_seed(3)
    v_render = torch.randn((5, H, W), device="cuda", generator=g)
    v_rec = pkg.raster.blend_backward(ctx, v_render)
    torch.cuda.synchronize()
    cols = ctx.rec[:, 7:12].double()
    lhs = (cols * v_rec[:, 6:11].double()).sum(0)                       # sum_g c_g . dL/dc_g   per channel
    rhs = (v_render.double() * render.double()).sum(dim=(1, 2))         # sum_p v . render      (bg = 0)
    scale = (v_render.double().abs() * render.double().abs()).sum(dim=(1, 2))
    assert bool(((lhs - rhs).abs() <= 1e-5 * scale).all()), (lhs, rhs)
    v_rec2 = pkg.raster.blend_backward(ctx, 2.0 * v_render)
    num = (v_rec2 - 2.0 * v_rec).abs().amax(0)
    den = v_rec2.abs().amax(0) + 1e-30
    assert float((num / den)[:11].max()) <= 1e-4, "backward is not linear in v_render"
    gg = pkg.raster.preprocess_backward(ctx, v_rec, means, ls, None, lo)
    for k in ("means", "scales", "opacities", "colors"):
        assert bool(torch.isfinite(gg[k]).all()), k
    culled = ctx.radii[:, 0] <= 0
    assert float(gg["means"][culled].abs().max() if culled.any() else 0.0) == 0.0, "culled splats must get no gradient"
