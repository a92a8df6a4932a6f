"""GPU parity: the sm_100a kernels, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Bars (BASELINE.json): records / tile keys / ranges bit-exact; images <= 1e-4
max-abs; Gaussian and pose gradients <= 1e-3 relative (||a-b||_inf / ||b||_inf per tensor)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4
GRAD_TOL = 1e-3


def relerr(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


def t(a, dtype=torch.float32):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype)


def random_scene(n, W, H, f, iso, seed, big=False):
    rng = np.random.RandomState(seed)
    cx, cy = (W - 1) / 2.0, (H - 1) / 2.0
    u, v = rng.rand(n) * (W + 40) - 20, rng.rand(n) * (H + 40) - 20
    z = 0.5 + rng.rand(n) * 4
    z[: n // 50] = rng.rand(n // 50) * 0.02 - 0.005  # behind / at the near plane
    means = np.stack([(u - cx) / f * z, (v - cy) / f * z, z], -1).astype(np.float32)
    s = (np.abs(z) / f * (0.5 + (8.0 if big else 2.5) * rng.rand(n))).astype(np.float32)
    scales = s[:, None] if iso else (s[:, None] * (0.4 + rng.rand(n, 3))).astype(np.float32)
    quats = None if iso else rng.randn(n, 4).astype(np.float32)
    opac = rng.rand(n).astype(np.float32)
    opac[: n // 40] = 0.003  # below 1/255
    cols = rng.rand(n, 3).astype(np.float32)
    return dict(W=W, H=H, f=f, cx=cx, cy=cy, means=means, scales=scales, quats=quats, opac=opac, cols=cols)


def run_both(pkg, orc, sc, ch, flags=0, pose=None, viewmat=None, bg=(0, 0, 0)):
    ocam = orc.make_camera(sc["W"], sc["H"], sc["f"], sc["f"], sc["cx"], sc["cy"], viewmat, bg=bg)
    gcam = pkg.make_camera(sc["W"], sc["H"], sc["f"], sc["f"], sc["cx"], sc["cy"], viewmat, bg=bg)
    opose = None if pose is None else orc.make_pose(*pose)
    gpose = None if pose is None else pkg.make_pose(*pose)
    o = orc.render_full(ocam, sc["means"], sc["scales"], sc["quats"], sc["opac"], sc["cols"], ch=ch, flags=flags,
                        pose=opose)
    render, alpha, ctx = pkg.raster.forward(gcam, t(sc["means"]), t(sc["scales"]), t(sc["quats"]), t(sc["opac"]),
                                            t(sc["cols"]), ch=ch, flags=flags, pose=gpose, keep_keys=True)
    torch.cuda.synchronize()
    return ocam, opose, o, render, alpha, ctx


def assert_forward_parity(o, render, alpha, ctx):
    # bit-exact integer / record work
    assert np.array_equal(ctx.radii.cpu().numpy(), o["radii"]), "radii differ"
    assert np.array_equal(ctx.tiles.cpu().numpy(), o["tiles"]), "tile counts differ"
    rec = ctx.rec.cpu().numpy()
    assert np.array_equal(rec.view(np.uint32), o["rec"].view(np.uint32)), "projected records not bit-exact"
    assert ctx.M == o["keys"].shape[0]
    assert np.array_equal(ctx.keys.cpu().numpy().view(np.uint64), o["keys"]), "sorted keys differ"
    assert np.array_equal(ctx.vals.cpu().numpy(), o["vals"]), "sorted values differ"
    assert np.array_equal(ctx.ranges.cpu().numpy(), o["ranges"]), "tile ranges differ"
    # images
    err = np.abs(render.cpu().numpy() - o["render"])
    scale = np.maximum(1.0, np.abs(o["render"]).reshape(o["render"].shape[0], -1).max(1))[:, None, None]
    assert (err / scale).max() <= IMG_TOL, f"image error {(err / scale).max()}"
    assert np.abs(alpha.cpu().numpy() - o["alpha"]).max() <= IMG_TOL
    assert np.array_equal(ctx.last.cpu().numpy(), o["last"]), "last-contributor ids differ"


@pytest.mark.parametrize("iso", [True, False])
@pytest.mark.parametrize("ch", [4, 5])
def test_forward_parity_random(pkg, orc, iso, ch):
    sc = random_scene(6000, 200, 136, 180.0, iso, seed=11 + ch)
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, ch, bg=(0.2, 0.1, 0.4))
    assert_forward_parity(o, render, alpha, ctx)


def test_forward_parity_view_and_pose(pkg, orc):
    sc = random_scene(5000, 168, 120, 150.0, False, seed=5)
    vm = np.eye(4, dtype=np.float32)
    c, s = np.cos(0.05), np.sin(0.05)
    vm[:3, :3] = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float32)
    vm[:3, 3] = [0.03, -0.02, 0.05]
    pose = (np.array([0.99, 0.01, -0.02, 0.03], np.float32), np.array([0.01, 0.02, -0.03], np.float32))
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, 5, pose=pose, viewmat=vm)
    assert_forward_parity(o, render, alpha, ctx)


def test_forward_parity_raw_params(pkg, orc):
    """log-scales / logit-opacities activated inside the kernel with the deterministic exp."""
    sc = random_scene(5000, 160, 128, 150.0, True, seed=9)
    sc["scales"] = np.log(sc["scales"]).astype(np.float32)
    o_ = np.clip(sc["opac"], 1e-3, 1 - 1e-3)
    sc["opac"] = np.log(o_ / (1 - o_)).astype(np.float32)
    flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, 5, flags=flags)
    assert_forward_parity(o, render, alpha, ctx)


def test_forward_edge_cases(pkg, orc):
    # empty input, all culled, a single huge splat covering every tile, ragged image size
    sc = random_scene(64, 50, 37, 40.0, True, seed=1)
    for sel in (slice(0, 0), slice(0, 1)):
        sub = dict(sc, means=sc["means"][sel], scales=sc["scales"][sel], opac=sc["opac"][sel], cols=sc["cols"][sel])
        _, _, o, render, alpha, ctx = run_both(pkg, orc, sub, 4)
        assert_forward_parity(o, render, alpha, ctx)
    big = dict(sc)
    big["means"] = np.array([[0.0, 0.0, 1.0]], np.float32)
    big["scales"] = np.array([[0.6]], np.float32)
    big["opac"] = np.array([0.9], np.float32)
    big["cols"] = np.array([[0.3, 0.6, 0.9]], np.float32)
    _, _, o, render, alpha, ctx = run_both(pkg, orc, big, 5)
    assert ctx.M == o["keys"].shape[0] == 4 * 3
    assert_forward_parity(o, render, alpha, ctx)


def test_forward_many_per_tile(pkg, orc):
    """> 256 Gaussians per tile (several staging batches) and saturated pixels (early exit)."""
    sc = random_scene(20000, 64, 48, 60.0, True, seed=21, big=True)
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, 5)
    assert (np.diff(o["ranges"]) > 256).any()
    assert_forward_parity(o, render, alpha, ctx)


@pytest.mark.parametrize("n,lo", [(9000, 2048), (60000, 8192)])
def test_binning_long_tiles(pkg, orc, n, lo):
    """Tiles holding > 2048 (64 KB shared-memory sort) and > 8192 (in-place global sort) intersections."""
    sc = random_scene(n, 48, 32, 50.0, True, seed=77, big=True)
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, 4)
    assert np.diff(o["ranges"]).max() > lo
    assert_forward_parity(o, render, alpha, ctx)


@pytest.mark.parametrize("iso", [True, False])
def test_backward_parity(pkg, orc, iso):
    sc = random_scene(5000, 176, 120, 160.0, iso, seed=31)
    pose = (np.array([0.99, 0.02, -0.01, 0.02], np.float32), np.array([0.02, -0.01, 0.02], np.float32))
    ocam, opose, o, render, alpha, ctx = run_both(pkg, orc, sc, 5, pose=pose, bg=(0.1, 0.2, 0.3))
    rng = np.random.RandomState(2)
    v_render = rng.randn(5, sc["H"], sc["W"]).astype(np.float32)
    v_alpha = rng.randn(sc["H"], sc["W"]).astype(np.float32)
    ov = orc.blend_bwd(ocam, 5, o["rec"], o["vals"], o["ranges"], o["alpha"], o["last"], v_render, v_alpha)
    gv = pkg.raster.blend_backward(ctx, t(v_render), t(v_alpha))
    torch.cuda.synchronize()
    for c in range(11):
        assert relerr(gv[:, c].cpu().numpy(), ov[:, c]) <= GRAD_TOL, f"v_rec column {c}"
    # feed the ORACLE's v_rec to both projection backwards so only that stage differs
    og = orc.preprocess_bwd(ocam, sc["means"], sc["scales"], sc["quats"], sc["opac"], sc["cols"], o["radii"],
                            o["rec"], ov, ch=5, pose=opose)
    gg = pkg.raster.preprocess_backward(ctx, t(ov), t(sc["means"]), t(sc["scales"]), t(sc["quats"]), t(sc["opac"]),
                                        want_pose=True, want_means2d=True)
    torch.cuda.synchronize()
    for k in ("means", "scales", "opacities", "colors") + (() if iso else ("quats",)):
        assert relerr(gg[k].cpu().numpy().reshape(og[k].shape), og[k]) <= GRAD_TOL, k
    assert relerr(gg["pose"].cpu().numpy(), og["pose"]) <= GRAD_TOL, "pose gradient"
    # end to end (GPU v_rec -> GPU projection backward)
    ge = pkg.raster.preprocess_backward(ctx, gv, t(sc["means"]), t(sc["scales"]), t(sc["quats"]), t(sc["opac"]),
                                        want_pose=True)
    for k in ("means", "scales", "opacities", "colors"):
        assert relerr(ge[k].cpu().numpy().reshape(og[k].shape), og[k]) <= GRAD_TOL, "e2e " + k
    assert relerr(ge["pose"].cpu().numpy(), og["pose"]) <= GRAD_TOL
    # means2D side channel = v_xy * (W/2, H/2)
    m2 = gg["means2d"].cpu().numpy()
    assert relerr(m2[:, 0], ov[:, 0] * 0.5 * sc["W"]) <= 1e-6 and relerr(m2[:, 1], ov[:, 1] * 0.5 * sc["H"]) <= 1e-6


def test_splatam_shaped_frame(pkg, orc, syn):
    """BASELINE config 1 shape at reduced resolution: one Gaussian per valid-depth pixel."""
    params, frames, K = syn.make_map("tiny")
    W, H, fx, fy, cx, cy = syn.SHAPES["tiny"][:6]
    sc = dict(W=W, H=H, f=fx, cx=cx, cy=cy, means=params["means3D"], scales=params["log_scales"], quats=None,
              opac=params["logit_opacities"].reshape(-1), cols=params["rgb_colors"])
    flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
    q, tr = syn.w2c_to_pose(syn.trajectory(3)[2])
    _, _, o, render, alpha, ctx = run_both(pkg, orc, sc, 5, flags=flags, pose=(q, tr))
    assert_forward_parity(o, render, alpha, ctx)
    assert alpha.mean().item() > 0.8


@pytest.mark.parametrize("n,W,H", [(6000, 200, 136), (60000, 48, 32)])
def test_literal_cub_binning_equals_counting_sort(pkg, orc, n, W, H):
    """isogs_tile_bin_cub -- the literal duplicateWithKeys + cub::DeviceRadixSort + identifyTileRanges pipeline that
    north_star (2) names -- against the default counting-sort path AND the oracle: unsorted keys, sorted keys, values
    and ranges all bit-identical (incl. tiles with > 8192 intersections)."""
    sc = random_scene(n, W, H, 60.0 if W < 100 else 180.0, True, seed=3, big=W < 100)
    cam = pkg.make_camera(sc["W"], sc["H"], sc["f"], sc["f"], sc["cx"], sc["cy"])
    ocam = orc.make_camera(sc["W"], sc["H"], sc["f"], sc["f"], sc["cx"], sc["cy"])
    rec, radii, tiles, cum, M = pkg.raster.preprocess(cam, t(sc["means"]), t(sc["scales"]), None, t(sc["opac"]),
                                                      t(sc["cols"]), ch=4)
    a = pkg.raster.tile_bin(cam, rec, radii, cum, M)
    b = pkg.raster.tile_bin(cam, rec, radii, cum, M, literal=True)
    torch.cuda.synchronize()
    for x, y, name in zip(a[:3], b[:3], ("sorted keys", "sorted values", "ranges")):
        assert torch.equal(x, y), name
    orec, oradii, otiles = orc.preprocess_fwd(ocam, sc["means"], sc["scales"], None, sc["opac"], sc["cols"], ch=4)
    ks, vs, rg, ku, vu = orc.tile_bin(ocam, orec, oradii, otiles, want_unsorted=True)
    assert np.array_equal(b[3].cpu().numpy().view(np.uint64), ku), "unsorted keys (duplicateWithKeys emission order)"
    assert np.array_equal(b[4].cpu().numpy(), vu)
    assert np.array_equal(b[0].cpu().numpy().view(np.uint64), ks) and np.array_equal(b[2].cpu().numpy(), rg)
