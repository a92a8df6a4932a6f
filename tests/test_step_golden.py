"""The tracking / mapping step against the reference's OWN Python code.

tests/golden/splatam_step_64x48.npz was written by oracle/pin_against_splatam.py, which imports
python_ordinary/scripts/splatam.py (get_pointcloud, initialize_params, get_loss, add_new_gaussians,
initialize_camera_pose, initialize_optimizer) and utils/slam_external.prune_gaussians and runs them
on the CPU with the oracle standing in for the CUDA rasterizer.  Here:
  * CPU (`-m "not gpu"`): the oracle's step restatement (tests/oracle_slam.py) against the fixture;
  * GPU (`-m gpu`): slam_step.track_step / map_step / add_new_gaussians / prune_gaussians /
    init_camera_pose through the C ABI against the fixture.
Tolerances: losses 1e-5 relative, gradients 1e-3 of the largest component (north_star), Adam-updated
values 1e-6 absolute where the gradient is not numerical noise (Adam's first step is lr*sign(g))."""
import os

import numpy as np
import pytest

import oracle_slam as osl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "splatam_step_64x48.npz")
GAUSS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")


@pytest.fixture(scope="module")
def fx():
    return dict(np.load(FIX))


def _cam_args(fx):
    W, H, fx_, fy, cx, cy = [float(v) for v in fx["intr"]]
    return int(W), int(H), fx_, fy, cx, cy


def _relmax(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _row_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a.reshape(b.shape) - b).reshape(b.shape[0], -1).max(1) / max(np.abs(b).max(), 1e-30)


def _check_map_gradients(get_grad, fx):
    """Gradients within 1e-3 of the largest component (north_star) for every Gaussian.  (An earlier fixture
    had two appended Gaussians at EXACTLY equal depth by mirror symmetry of the analytic frame; torch's
    `rel_w2c @ pts4.T` and the fused transform differ by 1 ulp there and swapped their compositing order.
    The frame now has no such symmetry.)  The iso term tolerates a handful of rows: the centres come from a pixel
    grid, so the 16th neighbour has near-ties, and torch.cdist evaluates |a|^2 + |b|^2 - 2ab (1e-3 relative on
    these distances) while the oracle / kernel compute |a - b|^2 -- 3 of 3199 rows pick another 16th neighbour.
    Returns the rows to compare."""
    bad = np.zeros(fx["map_in_means3D"].shape[0], bool)
    for k in GAUSS:
        bad |= _row_err(get_grad(k), fx[f"map_grad_{k}"]) > 1e-3
    assert bad.sum() <= 8, np.nonzero(bad)[0]
    return ~bad


def _check_adam_update(p1, ref1, grad, lr, steps, rows, name):
    """Adam's update is lr * f(g_1/|g_2|...): an element's update error is lr x the RELATIVE error of that
    gradient element.  Elements above 1% of the largest gradient must agree to 5% of lr; the others (down to
    numerical noise, where the sign-like first step may flip) are bounded by 2*lr*steps."""
    g = np.abs(grad.reshape(ref1.shape))
    solid = (g > 1e-2 * g.max()) & rows.reshape((-1,) + (1,) * (g.ndim - 1))
    assert solid.mean() > 0.02, (name, solid.mean())
    d = np.abs(np.asarray(p1).reshape(ref1.shape) - ref1)
    # >= 99.5 %: a near-tie in the 16-NN of the iso term (see _check_map_gradients) can also appear in a later
    # iteration, for rows that were fine in the first one
    frac = float((d[solid] < 1e-6 + 0.05 * lr).mean())
    assert frac >= 0.995, (name, frac, d[solid].max())
    assert d.max() <= 2.0 * lr * steps + 1e-6, name


# ------------------------------------------------------------------ CPU: oracle vs fixture
def test_fixture_notes(fx):
    assert "get_loss(tracking=True" in str(fx["notes"])


def test_oracle_init_matches_reference(orc, fx):
    """get_pointcloud + initialize_params (splatam.py:296-399): integer-pixel back-projection, s = z / f."""
    W, H, f_x, f_y, cx, cy = _cam_args(fx)
    d = fx["d0"][0]
    ys, xs = np.nonzero(d > 0)
    z = d[ys, xs]
    means = np.stack([(xs - cx) / f_x * z, (ys - cy) / f_y * z, z], -1)
    assert np.abs(means - fx["init_means3D"]).max() < 1e-5
    assert np.abs(fx["rgb0"][:, ys, xs].T - fx["init_rgb_colors"]).max() == 0
    base = np.log(z / ((f_x + f_y) / 2))[:, None]
    assert np.abs(fx["init_log_scales"] - base).max() < 0.06          # N(0, 0.01) noise, splatam.py:366
    assert np.abs(fx["init_logit_opacities"]).max() == 0


def test_oracle_ssim_matches_reference(orc, fx):
    """calc_ssim + autograd (slam_external.py:54-97) on x = rgb0, y = rgb1."""
    v = np.zeros_like(fx["rgb0"])
    val = orc.ssim(fx["rgb0"], fx["rgb1"], 1.0, v)
    # smooth images: sigma^2 = E[x^2] - mu^2 cancels to ~c2 in the reference's float32 (the oracle accumulates
    # in double), hence 2e-5 rather than 1e-7
    assert abs(val - fx["ssim_value"][0]) < 2e-5
    assert _relmax(v, fx["ssim_grad"]) < 1e-3


def test_oracle_tracking_matches_reference(orc, fx):
    ocam = orc.make_camera(*_cam_args(fx))
    params = {k: fx[f"init_{k}"] for k in GAUSS}
    ts = osl.TrackState(fx["track_pose0"][:4], fx["track_pose0"][4:])
    for it in range(3):
        loss, gp = osl.track_iteration(orc, ocam, params, ts, fx["rgb1"], fx["d1"][0])
        assert abs(loss - fx["track_losses"][it, 0]) < 2e-5 * fx["track_losses"][it, 0], it
        assert _relmax(gp, fx["track_grads"][it]) < 1e-3, (it, gp, fx["track_grads"][it])
        assert np.abs(ts.pose - fx["track_poses"][it]).max() < 2e-6, it


def test_oracle_outlier_mask_matches_reference(orc, fx):
    """get_loss(tracking=True, ignore_outlier_depth_loss=True): mask = depth error < 10 x median (splatam.py:551-554)."""
    ocam = orc.make_camera(*_cam_args(fx))
    params = {k: fx[f"init_{k}"] for k in GAUSS}
    ts = osl.TrackState(fx["track_pose0"][:4], fx["track_pose0"][4:])
    loss, gp = osl.track_iteration(orc, ocam, params, ts, fx["rgb1"], fx["d1"][0], ignore_outlier_depth=True)
    assert abs(loss - fx["track_outlier_losses"][0]) < 2e-5 * fx["track_outlier_losses"][0]
    assert fx["track_outlier_losses"][0] < 0.98 * fx["track_losses"][0, 0], "the fixture must contain outliers"
    assert _relmax(gp, fx["track_outlier_grad"]) < 1e-3


def test_oracle_iso_flat_matches_reference(orc, fx):
    """compute_flat_loss + compute_iso_surface_loss_sampled (splatam.py:42-54,162-262) behind the preamble of
    get_loss (:578-640), on 600 three-axis Gaussians with 150 sampled queries."""
    total, out, g, nbr = orc.iso_flat_loss(fx["geo_means3D"], fx["geo_log_scales"], fx["geo_unnorm_rotations"],
                                           fx["geo_logit_opacities"], fx["geo_query_idx"])
    assert np.abs(out - fx["geo_losses"]).max() < 2e-5 * np.abs(fx["geo_losses"]).max(), (out, fx["geo_losses"])
    assert (nbr[:, 0] == fx["geo_query_idx"]).all(), "the nearest centre of a query is the query itself"
    for k in ("means3D", "log_scales", "unnorm_rotations", "logit_opacities"):
        assert _relmax(g[k].reshape(-1), fx[f"geo_grad_{k}"].reshape(-1)) < 1e-3, k


def test_oracle_mapping_matches_reference(orc, fx):
    ocam = orc.make_camera(*_cam_args(fx))
    st = osl.new_state({k: fx[f"map_in_{k}"] for k in GAUSS})
    q, t = fx["track_poses"][2][:4], fx["track_poses"][2][4:]
    g0 = {}
    for it in range(2):
        loss = osl.map_iteration(orc, ocam, st, q, t, fx["rgb1"], fx["d1"][0], grads_out=g0 if it == 0 else None,
                                 ssim=True, isoflat=True)
        assert abs(loss - fx["map_losses"][it, 0]) < 2e-5 * fx["map_losses"][it, 0], it
    rows = _check_map_gradients(lambda k: g0[k], fx)
    for k in GAUSS:
        _check_adam_update(st["p"][k], fx[f"map_out_{k}"], fx[f"map_grad_{k}"], osl.MAP_LRS[k], 2, rows, k)


MODES = os.path.join(ROOT, "tests", "golden", "splatam_modes_64x48.npz")


@pytest.mark.parametrize("mode", ["outlier", "ba"])
def test_oracle_mapping_side_modes_match_reference(orc, fx, mode):
    """The two get_loss modes the shipped configs leave off, against the reference's OWN get_loss
    (tests/golden/splatam_modes_64x48.npz from oracle/pin_modes_splatam.py, same inputs as the mapping fixture):
    mapping with ignore_outlier_depth_loss (mean over the masked pixels, splatam.py:545-559) and do_ba=True (pose AND
    Gaussians get gradients, :505-510), each with the complete mapping loss."""
    fm = dict(np.load(MODES))
    ocam = orc.make_camera(*_cam_args(fx))
    st = osl.new_state({k: fx[f"map_in_{k}"] for k in GAUSS})
    q, t = fm["pose"][:4], fm["pose"][4:]
    g = {}
    ts = osl.TrackState(q, t) if mode == "ba" else None
    loss = osl.map_iteration(orc, ocam, st, q, t, fx["rgb1"], fx["d1"][0], grads_out=g, ssim=True, isoflat=True,
                             ignore_outlier_depth=(mode == "outlier"), ba=ts)
    ref = fm[f"{mode}_losses"]
    assert abs(loss - ref[0]) < 2e-5 * ref[0], (loss, ref)
    if mode == "outlier":   # the mask bites: the depth term differs from the unmasked evaluation of the same inputs
        assert fm["outlier_losses"][2] < 0.95 * fm["ba_losses"][2]
    rows = _check_map_gradients(lambda k: g[k], dict(fx, **{f"map_grad_{k}": fm[f"{mode}_grad_{k}"] for k in GAUSS}))
    assert rows.mean() > 0.99
    if mode == "ba":
        assert _relmax(g["pose"], fm["ba_grad_pose"]) < 1e-3, (g["pose"], fm["ba_grad_pose"])


def test_init_camera_pose_matches_reference(fx):
    """initialize_camera_pose (splatam.py:844-863) on the fixture's poses of frames 0 and 1 (host arithmetic)."""
    import importlib

    import isogs_b200  # noqa: F401
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    p1 = fx["track_poses"][2]
    q, t = ss.init_camera_pose((p1[:4], p1[4:]), (np.array([1, 0, 0, 0], np.float32), np.zeros(3, np.float32)))
    assert np.abs(np.concatenate([q, t]) - fx["init_pose2"]).max() < 1e-6


def test_oracle_keyframe_overlap_matches_reference(fx):
    """keyframe_selection_overlap (utils/keyframe_selection.py:40-95) with the reference's own pixel draw."""
    pct = osl.keyframe_overlap(fx["d1"][0], fx["intr"], fx["kf_cur_w2c"], fx["kf_w2cs"], fx["kf_sample_idx"])
    assert np.abs(pct - fx["kf_percent"]).max() < 1e-6
    assert sorted(np.nonzero(pct > 0)[0].tolist()) == fx["kf_selected"].tolist()


# ------------------------------------------------------------------ GPU: product path vs fixture
def _gpu(fx):
    import importlib

    import isogs_b200
    import torch
    pkg = isogs_b200.pkg
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()  # noqa: E731
    return pkg, ss, torch, dev, pkg.make_camera(*_cam_args(fx))


@pytest.mark.gpu
def test_gpu_track_step_matches_reference(fx):
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"init_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    ps = ss.PoseState(fx["track_pose0"][:4], fx["track_pose0"][4:], "cuda")
    rgb, depth = dev(fx["rgb1"]), dev(fx["d1"][0])
    for it in range(3):
        losses = ss.track_step(ws, gmap, ps, rgb, depth)
        torch.cuda.synchronize()
        ref = fx["track_losses"][it]
        assert abs(float(losses[0]) - ref[0]) < 1e-4 * ref[0], (it, float(losses[0]), ref[0])
        assert _relmax(ws.v_pose.cpu().numpy(), fx["track_grads"][it]) < 1e-3, it
        assert np.abs(ps.pose.cpu().numpy() - fx["track_poses"][it]).max() < 5e-6, it
    assert not ws.check_overflow(wait=True)


@pytest.mark.gpu
def test_gpu_track_step_outlier_mask_matches_reference(fx):
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"init_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    ps = ss.PoseState(fx["track_pose0"][:4], fx["track_pose0"][4:], "cuda")
    losses = ss.track_step(ws, gmap, ps, dev(fx["rgb1"]), dev(fx["d1"][0]), ignore_outlier_depth=True)
    torch.cuda.synchronize()
    ref = fx["track_outlier_losses"]
    assert abs(float(losses[0]) - ref[0]) < 1e-4 * ref[0], (float(losses[0]), ref[0])
    assert _relmax(ws.v_pose.cpu().numpy(), fx["track_outlier_grad"]) < 1e-3


@pytest.mark.gpu
def test_gpu_map_step_matches_reference(fx):
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"map_in_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    pose = pkg.make_pose(fx["track_poses"][2][:4], fx["track_poses"][2][4:])
    rgb, depth = dev(fx["rgb1"]), dev(fx["d1"][0])
    inv_nd = 1.0 / int((fx["d1"] > 0).sum())
    rows = None
    for it in range(2):
        losses = ss.map_step(ws, gmap, pose, rgb, depth, inv_nd, ssim=True, isoflat=True)
        torch.cuda.synchronize()
        ref = fx["map_losses"][it]
        assert abs(float(losses[0]) - ref[0]) < 2e-4 * ref[0], (it, float(losses[0]), ref[0])
        if it == 0:
            g = {k: gmap.g[k].cpu().numpy() for k in GAUSS}
            rows = _check_map_gradients(lambda k: g[k], fx)
            # a12: means2D.grad of the reference comes from the colour render alone (splatam.py:528-530) and carries
            # the Inria NDC scaling v_xy * (W/2, H/2) -- the repo's one convention for the screen-space gradient
            v = ss.colour_only_screen_gradient(ws, gmap)[:gmap.count, :2].cpu().numpy()
            W, H = gcam.width, gcam.height
            ref2d = fx["map_means2D_grad"]
            assert np.all(ref2d[:, 2] == 0)
            got2d = v * np.array([0.5 * W, 0.5 * H], np.float32)
            assert np.abs(got2d - ref2d[:, :2]).max() <= 1e-3 * np.abs(ref2d).max()
    for k in GAUSS:
        _check_adam_update(gmap.view(k).cpu().numpy(), fx[f"map_out_{k}"], fx[f"map_grad_{k}"], ss.MAPPING_LRS[k], 2,
                           rows, k)


@pytest.mark.gpu
def test_gpu_map_step_outlier_mode_matches_reference(fx):
    """map_step(ignore_outlier_depth=True) with the complete mapping loss against the reference's own
    get_loss(mapping=True, ignore_outlier_depth_loss=True) (tests/golden/splatam_modes_64x48.npz): the depth term is
    the mean over the pixels with |err| < 10 * median(err), normaliser counted on the device."""
    fm = dict(np.load(MODES))
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"map_in_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    pose = pkg.make_pose(fm["pose"][:4], fm["pose"][4:])
    inv_nd = 1.0 / int((fx["d1"] > 0).sum())            # ignored in this mode: the normaliser is 1 / n(mask)
    losses = ss.map_step(ws, gmap, pose, dev(fx["rgb1"]), dev(fx["d1"][0]), inv_nd, ssim=True, isoflat=True,
                         ignore_outlier_depth=True)
    torch.cuda.synchronize()
    ref = fm["outlier_losses"]
    assert abs(float(losses[0]) - ref[0]) < 2e-4 * ref[0], (float(losses[0]), ref[0])
    g = {k: gmap.g[k].cpu().numpy() for k in GAUSS}
    rows = _check_map_gradients(lambda k: g[k], dict(fx, **{f"map_grad_{k}": fm[f"outlier_grad_{k}"] for k in GAUSS}))
    assert rows.mean() > 0.99


@pytest.mark.gpu
def test_gpu_map_step_ba_pose_gradient_matches_reference(fx):
    """map_step(ba=PoseState) (isogs_map_step_ba) against the reference's own get_loss(mapping=True, do_ba=True): the
    flat / iso terms do not depend on the pose, so the keyframe-pose gradient of the reference's complete loss is the
    pose gradient of colour (0.8 L1 + 0.2 (1 - SSIM)) + depth -- what map_step(ba=..., ssim=True) evaluates."""
    fm = dict(np.load(MODES))
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"map_in_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    ps = ss.PoseState(fm["pose"][:4], fm["pose"][4:], "cuda")
    inv_nd = 1.0 / int((fx["d1"] > 0).sum())
    lrs = dict(ss.MAPPING_LRS, cam_unnorm_rots=4e-4, cam_trans=2e-3)
    losses = ss.map_step(ws, gmap, pkg.make_pose(fm["pose"][:4], fm["pose"][4:]), dev(fx["rgb1"]), dev(fx["d1"][0]), inv_nd,
                         lrs=lrs, ssim=True, ba=ps)
    torch.cuda.synchronize()
    ref = fm["ba_losses"]
    want = ref[1] + ref[2]                      # the recorded terms are the weighted ones: colour + depth, no flat / iso
    assert abs(float(losses[0]) - want) < 5e-4 * want, (float(losses[0]), want, ref)
    v = ws.v_pose.cpu().numpy()
    assert _relmax(v, fm["ba_grad_pose"]) < 1e-3, (v, fm["ba_grad_pose"])
    assert np.abs(ps.pose.cpu().numpy() - fm["pose"]).max() > 1e-4, "the pose must have taken its Adam step"


@pytest.mark.gpu
def test_gpu_add_new_gaussians_matches_reference(fx):
    """add_new_gaussians (splatam.py:799-841) at the tracked pose of frame 1: same pixels selected, in the
    same (row-major) order, same back-projected parameters ("anisotropic": no noise)."""
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"init_{k}"] for k in GAUSS}, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    n0 = gmap.count
    pose = (fx["track_poses"][2][:4], fx["track_poses"][2][4:])
    added = ss.add_new_gaussians(ws, gmap, pose, dev(fx["rgb1"]), dev(fx["d1"][0]), sil_thres=0.5, noise_std=0.0)
    assert added == int(fx["added"][0])
    assert np.abs(gmap.view("means3D")[n0:].cpu().numpy() - fx["new_means3D"]).max() < 2e-6
    assert np.array_equal(gmap.view("rgb_colors")[n0:].cpu().numpy(), fx["new_rgb_colors"])
    assert np.abs(gmap.view("log_scales")[n0:].cpu().numpy() - fx["new_log_scales"]).max() < 1e-6
    assert np.array_equal(gmap.view("logit_opacities")[n0:].cpu().numpy(), fx["new_logit_opacities"])
    assert np.array_equal(gmap.view("unnorm_rotations")[n0:].cpu().numpy(), fx["new_unnorm_rotations"])
    assert np.array_equal(gmap.view("means3D")[:n0].cpu().numpy(), fx["init_means3D"])


@pytest.mark.gpu
def test_gpu_prune_matches_reference(fx):
    """prune_gaussians + remove_points (slam_external.py:139-188): opacity < 0.005 or max scale >
    0.1 * scene_radius; survivors keep their order."""
    pkg, ss, torch, dev, gcam = _gpu(fx)
    params = {k: fx[f"map_out_{k}"] for k in GAUSS}
    params["logit_opacities"], params["log_scales"] = fx["prune_in_logit_opacities"], fx["prune_in_log_scales"]
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    removed = ss.prune_gaussians(ws, gmap, 0.005, 0.1 * float(fx["scene_radius"][0]))
    assert gmap.count == fx["prune_out_means3D"].shape[0] and removed == params["means3D"].shape[0] - gmap.count
    assert np.array_equal(gmap.view("means3D").cpu().numpy(), fx["prune_out_means3D"])


@pytest.mark.gpu
@pytest.mark.parametrize("W,H", [(64, 48), (70, 50), (33, 7), (1200, 680)])
def test_gpu_ssim_kernel_matches_oracle(orc, fx, W, H):
    """isogs_ssim_loss_fused against the oracle (double accumulation) on partial tiles / images smaller than
    the window / the full Replica size, and against the reference's own value + autograd gradient at 64x48."""
    import ctypes as C
    pkg, ss, torch, dev, _ = _gpu(fx)
    rng = np.random.RandomState(W * 1000 + H)
    if (W, H) == (64, 48):
        x, y = fx["rgb0"], fx["rgb1"]
    else:
        x = rng.rand(3, H, W).astype(np.float32)
        y = np.clip(x + 0.3 * (rng.rand(3, H, W).astype(np.float32) - 0.5), 0, 1).astype(np.float32)
    gcam = pkg.make_camera(W, H, 60.0, 60.0, W / 2, H / 2)
    weight = 0.1
    base = np.float32(0.4 / (3 * W * H))                               # what the L1 kernel leaves there: w_im / (3P)
    v_ref = np.full((3, H, W), base, np.float32)                       # accumulates on top of what is there
    s_ref = orc.ssim(x, y, -weight, v_ref)
    dx, dy = dev(x), dev(y)
    v = torch.full((5, H, W), float(base), dtype=torch.float32, device="cuda")
    dmaps = torch.empty((9, H, W), dtype=torch.float32, device="cuda")
    small = torch.zeros((8,), dtype=torch.float32, device="cuda")
    small[0] = 2.0                                                     # losses[0] already holds the L1 part
    p = lambda t_: C.c_void_p(t_.data_ptr())                           # noqa: E731
    pkg._lib.check(pkg.lib().isogs_ssim_loss_fused(C.byref(gcam), p(dx), p(dy), weight, p(dmaps), p(v), p(small),
                                                   p(small[4:]), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    got = small.cpu().numpy()
    assert abs(got[4] - s_ref) < 2e-5, (got[4], s_ref)
    assert abs(got[0] - (2.0 + weight * (1.0 - s_ref))) < 5e-5   # ~2500 float atomics onto 2.0
    gv = v.cpu().numpy()
    assert np.all(gv[3:] == base), "depth planes of v_render must be untouched"
    scale = np.abs(v_ref - base).max()
    assert np.abs(gv[:3] - v_ref).max() < 1e-3 * scale
    if (W, H) == (64, 48):
        assert abs(got[4] - fx["ssim_value"][0]) < 2e-5
        assert _relmax((gv[:3].astype(np.float64) - base) / -weight, fx["ssim_grad"]) < 1e-3


@pytest.mark.gpu
def test_gpu_densify_matches_reference(fx):
    """accumulate_mean2d_gradient + densify (slam_external.py:100-104,191-243) run by the reference itself on 300
    isotropic Gaussians: statistics, clone / split selection, appended rows (fed the reference's own normal draw),
    split-parent removal + opacity / size prune, Adam-moment bookkeeping."""
    pkg, ss, torch, dev, gcam = _gpu(fx)
    gmap = ss.GaussianMap({k: fx[f"dens_in_{k}"] for k in GAUSS}, device="cuda")
    N = gmap.count
    for k in GAUSS:
        gmap.m[k][:N] = dev(fx[f"dens_in_m_{k}"]).reshape(N, -1)
        gmap.v[k][:N] = dev(fx[f"dens_in_v_{k}"]).reshape(N, -1)
    gmap.vars["means2D_gradient_accum"][:N, 0] = dev(fx["dens_accum0"])
    gmap.vars["denom"][:N, 0] = dev(fx["dens_denom0"])
    ws = ss.StepWorkspace(gcam, N, torch.device("cuda"))
    ws.radii[:N] = torch.from_numpy(np.stack([3 * fx["dens_seen"], 2 * fx["dens_seen"]], 1).astype(np.int32)).cuda()
    v_xy = dev(fx["dens_v_xy"])
    ss.accumulate_densify_stats(ws, gmap, v_xy=v_xy)
    assert np.allclose(gmap.vars["means2D_gradient_accum"][:N, 0].cpu().numpy(), fx["dens_accum1"], rtol=1e-6, atol=1e-10)
    assert np.array_equal(gmap.vars["denom"][:N, 0].cpu().numpy(), fx["dens_denom1"])
    assert np.array_equal(gmap.vars["max_2D_radius"][:N, 0].cpu().numpy(), 3.0 * fx["dens_seen"])
    dd = dict(ss.DENSIFY_DEFAULTS, start_after=0, remove_big_after=0, stop_after=1000, densify_every=1)
    n_clone, n_split, n_removed = ss.densify(ws, gmap, 1, 2.0, dd, noise=dev(fx["dens_noise"]), v_xy=v_xy)
    assert 2 * n_split == fx["dens_noise"].shape[0] and n_clone > 0
    assert gmap.count == fx["dens_out_means3D"].shape[0] == N + n_clone + 2 * n_split - n_removed
    for k in GAUSS:
        got, ref = gmap.view(k).cpu().numpy(), fx[f"dens_out_{k}"].reshape(gmap.count, -1)
        assert np.abs(got - ref).max() < 2e-6, k
    assert np.abs(gmap.m["means3D"][:gmap.count].cpu().numpy() - fx["dens_out_m_means3D"]).max() < 1e-9
    assert float(gmap.vars["means2D_gradient_accum"].abs().max()) == 0.0 and fx["dens_out_accum"].max() == 0.0
    # built-in generator: same counts, children spread around the parents with the parents' scale
    gmap2 = ss.GaussianMap({k: fx[f"dens_in_{k}"] for k in GAUSS}, device="cuda")
    gmap2.vars["means2D_gradient_accum"][:N, 0] = dev(fx["dens_accum1"])
    gmap2.vars["denom"][:N, 0] = dev(fx["dens_denom1"])
    ws.radii[:N] = torch.from_numpy(np.stack([3 * fx["dens_seen"], 2 * fx["dens_seen"]], 1).astype(np.int32)).cuda()
    c2 = ss.densify(ws, gmap2, 1, 2.0, dd, seed=3, v_xy=v_xy)
    assert c2 == (n_clone, n_split, n_removed)
    a, b = gmap2.view("means3D").cpu().numpy(), fx["dens_out_means3D"]
    same = np.abs(a - b).max(1) < 1e-6
    assert same.sum() >= gmap2.count - 2 * n_split and np.isfinite(a).all()
    spread = np.abs(a - b)[~same]
    assert 0.0 < spread.mean() < 0.2                               # scales are 0.005 .. 0.08


@pytest.mark.gpu
def test_gpu_keyframe_overlap_matches_reference(fx):
    pkg, ss, torch, dev, gcam = _gpu(fx)
    depth = dev(fx["d1"][0])
    idx = torch.from_numpy(fx["kf_sample_idx"]).cuda()
    pct = ss.keyframe_overlap(gcam, depth, fx["kf_cur_w2c"], fx["kf_w2cs"], pixels=len(fx["kf_sample_idx"]),
                              sample_idx=idx).cpu().numpy()
    assert np.abs(pct - fx["kf_percent"]).max() < 1e-6, (pct, fx["kf_percent"])
    rng = np.random.RandomState(0)
    sel = ss.select_keyframes(gcam, depth, fx["kf_cur_w2c"], fx["kf_w2cs"], k=2, pixels=len(idx), rng=rng, sample_idx=idx)
    assert len(sel) == 2 and set(int(i) for i in sel) <= set(fx["kf_selected"].tolist())
    # device-side sampling: a different draw of the same statistic (4000 samples -> tighter than the 400 of the fixture)
    pct2 = ss.keyframe_overlap(gcam, depth, fx["kf_cur_w2c"], fx["kf_w2cs"], pixels=4000, seed=5).cpu().numpy()
    assert np.all((pct2 > 0) == (fx["kf_percent"] > 0)) and np.abs(pct2 - fx["kf_percent"]).max() < 0.05


@pytest.mark.gpu
def test_gpu_iso_flat_matches_reference(orc, fx):
    """isogs_knn16 + isogs_iso_flat_loss against the reference's compute_iso_surface_loss_sampled /
    compute_flat_loss (fixture geo_*) and, for the neighbour table, against the brute-force oracle."""
    pkg, ss, torch, dev, gcam = _gpu(fx)
    params = {k: fx[f"geo_{k}"] for k in GAUSS if k != "rgb_colors"}
    params["rgb_colors"] = np.zeros_like(fx["geo_means3D"])
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    gmap.grad_flat.zero_()
    ws.small.zero_()
    qidx = torch.from_numpy(fx["geo_query_idx"]).cuda()
    for res in (None, 3, 40):                                   # coarse and fine grids give the same exact answer
        gmap.grad_flat.zero_()
        ws.small.zero_()
        ss.iso_flat_loss(ws, gmap, query_idx=qidx, res=res)
        torch.cuda.synchronize()
        _, _, _, nbr_ref = orc.iso_flat_loss(fx["geo_means3D"], fx["geo_log_scales"], fx["geo_unnorm_rotations"],
                                             fx["geo_logit_opacities"], fx["geo_query_idx"], with_grad=False)
        assert np.array_equal(ws.nbr.cpu().numpy(), nbr_ref), res
        geo = ws.geo.cpu().numpy()
        assert np.abs(geo - fx["geo_losses"]).max() < 1e-4 * np.abs(fx["geo_losses"]).max(), (geo, fx["geo_losses"])
        assert abs(float(ws.losses[0]) - (50 * fx["geo_losses"][0] + 2 * fx["geo_losses"][1])) < 1e-3
        for k in ("means3D", "log_scales", "unnorm_rotations", "logit_opacities"):
            assert _relmax(gmap.g[k].cpu().numpy().reshape(-1), fx[f"geo_grad_{k}"].reshape(-1)) < 1e-3, (res, k)


@pytest.mark.gpu
@pytest.mark.parametrize("n,q", [(20000, 512), (10, 10), (1, 1)])
def test_gpu_knn_exact(orc, fx, n, q):
    """Exactness of the grid search on clustered points (two dense blobs + a sparse halo), tiny inputs with fewer
    than 16 points (-1 padding) and the device-side query sampling."""
    import ctypes as C
    pkg, ss, torch, dev, gcam = _gpu(fx)
    rng = np.random.RandomState(n)
    pts = np.concatenate([0.05 * rng.randn(n // 2, 3), 2.0 + 0.3 * rng.randn(n - n // 2 - n // 10, 3),
                          10.0 * rng.rand(n // 10, 3)]).astype(np.float32)[:n]
    if n == 1:
        pts = np.zeros((1, 3), np.float32)
    params = dict(means3D=pts, rgb_colors=np.zeros_like(pts), unnorm_rotations=np.tile([1, 0, 0, 0], (n, 1)).astype(np.float32),
                  logit_opacities=np.zeros((n, 1), np.float32), log_scales=np.full((n, 3), -3.0, np.float32))
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(gcam, gmap.count, torch.device("cuda"))
    qi = rng.permutation(n)[:q].astype(np.int32)
    ss.iso_flat_loss(ws, gmap, query_idx=torch.from_numpy(qi).cuda(), with_grad=False)
    torch.cuda.synchronize()
    got = ws.nbr.cpu().numpy()
    d2 = ((pts[qi][:, None, :].astype(np.float64) - pts[None].astype(np.float64)) ** 2).sum(-1)
    k = min(16, n)
    ref_d = np.sort(d2, axis=1)[:, :k]
    got_d = np.take_along_axis(d2, got[:, :k].astype(np.int64), axis=1)
    assert (got[:, k:] == -1).all() and (got[:, :k] >= 0).all()
    assert np.allclose(got_d, ref_d, rtol=1e-5, atol=1e-12), "neighbour distances must be the k smallest, ascending"
    assert all(len(set(r[:k])) == k for r in got)
    if n >= 8192:                                                # sampled queries: valid indices, exact neighbours
        ss.iso_flat_loss(ws, gmap, seed=7, with_grad=False)
        torch.cuda.synchronize()
        sq = ws.query_idx.cpu().numpy()
        assert sq.shape[0] == 8192 and sq.min() >= 0 and sq.max() < n and len(np.unique(sq)) > 6000
        nb = ws.nbr.cpu().numpy()
        assert (nb[:, 0] == sq).all() or np.allclose(((pts[nb[:, 0]] - pts[sq]) ** 2).sum(-1), 0)
