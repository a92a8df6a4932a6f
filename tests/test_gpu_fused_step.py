"""The compiled host path (isogs_map_step / isogs_track_step: one C call per iteration) and its CUDA-graph replay
(SURVEY 8(f)-3), the sticky overflow word, the median scratch of the outlier mask and the uint8 / uint16 frame
unpacking -- all through the C ABI on a GPU."""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def t(a, dtype=torch.float32):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype)


@pytest.fixture(scope="module")
def ss(pkg):
    return importlib.import_module("isogs_slam_cpp_b200.slam_step")


def _scene(syn):
    params, frames, K = syn.make_map("tiny")
    return params, syn.SHAPES["tiny"][:6]


def _keyframes(syn, pkg, n=3):
    traj = syn.trajectory(8, step=3.0)
    out = []
    for k in range(2, 2 + n):
        c, d, _ = syn.render_frame("tiny", traj[k], seed=k)
        q, tr = syn.w2c_to_pose(traj[k])
        out.append(dict(rgb=t(c), depth=t(d[0]), pose=pkg.make_pose(q, tr), q=q, t=tr,
                        inv_nd=1.0 / max(1, int((d > 0).sum()))))
    return out


@pytest.mark.parametrize("ssim", [False, True])
def test_map_step_graph_replay_matches_eager(pkg, syn, ss, ssim):
    """10 mapping iterations over 3 keyframes: ONE captured graph replayed (the frame goes through the device-side
    slot, Adam's step count lives on the device) against the eager enqueue of the same kernels.  The first step
    sees identical parameters, so its render is bit-identical; later steps agree to the run-to-run spread of the
    float atomics in the blend backward."""
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    kfs = _keyframes(syn, pkg)
    runs = []
    for graph in (False, True, False):
        gmap = ss.GaussianMap(params, device="cuda")
        ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
        losses, first_render = [], None
        for it in range(10):
            k = kfs[it % len(kfs)]
            out = ss.map_step(ws, gmap, k["pose"], k["rgb"], k["depth"], k["inv_nd"], ssim=ssim, graph=graph)
            losses.append(out[:4].clone())
            if it == 0:
                first_render = ws.render.clone()
        torch.cuda.synchronize()
        if graph:
            assert len(ws._graphs) == 1, "one captured graph must serve every keyframe"
        assert int(gmap.adam_state.view(torch.int64)[0]) == 10, "device-side Adam step count"
        runs.append(dict(losses=torch.stack(losses).cpu().numpy(), render=first_render.cpu().numpy(),
                         p={k: gmap.view(k).cpu().numpy() for k in ("means3D", "rgb_colors", "logit_opacities", "log_scales")}))
    eager, graph, eager2 = runs
    assert np.array_equal(eager["render"], graph["render"]), "first-step render must be bit-identical"
    assert np.allclose(eager["losses"], graph["losses"], rtol=2e-4, atol=1e-7)
    lrs = ss.MAPPING_LRS
    for k in eager["p"]:
        spread = np.abs(eager["p"][k] - eager2["p"][k]).max()          # eager vs eager: the atomics' own spread
        d = np.abs(eager["p"][k] - graph["p"][k]).max()
        assert d <= max(3.0 * spread, 0.25 * lrs[k]), (k, d, spread)   # an L1 sign flip moves a step by O(lr)
        assert np.abs(graph["p"][k] - params[k].reshape(graph["p"][k].shape)).max() > 0, "the replay must have stepped"


def test_track_step_graph_replay_matches_eager(pkg, orc, syn, ss):
    """Tracking: the graph is captured once and replayed across iterations AND frames (PoseState.reset keeps the
    device buffers); against the eager path and the oracle's iteration."""
    import oracle_slam as osl
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam, ocam = pkg.make_camera(W, H, fx, fy, cx, cy), orc.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(8, step=3.0)
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
    rgb_buf = torch.empty((3, H, W), dtype=torch.float32, device="cuda")
    depth_buf = torch.empty((H, W), dtype=torch.float32, device="cuda")
    ps = None
    for f in (3, 4):
        c, d, _ = syn.render_frame("tiny", traj[f], seed=f)
        rgb_buf.copy_(t(c)); depth_buf.copy_(t(d[0]))                 # the frame lands in FIXED staging buffers
        q0, t0 = syn.w2c_to_pose(traj[f - 1])
        ps = ss.PoseState(q0, t0, "cuda") if ps is None else ps.reset(q0, t0)
        ts = osl.TrackState(q0, t0)
        for it in range(4):
            lo, gp = osl.track_iteration(orc, ocam, params, ts, np.ascontiguousarray(c), np.ascontiguousarray(d[0]))
            losses = ss.track_step(ws, gmap, ps, rgb_buf, depth_buf, graph=True)
            torch.cuda.synchronize()
            assert abs(float(losses[0]) - lo) <= 2e-4 * abs(lo), (f, it)
            assert np.abs(ps.pose.cpu().numpy() - ts.pose).max() <= 0.02 * 2e-3 + 1e-7
        bq, bt = ps.best_pose()
        assert np.abs(np.concatenate([bq, bt]) - ts.best).max() <= 1e-5
    assert len(ws._graphs) == 1, "one tracking graph across iterations and frames"


def test_overflow_is_sticky_and_turns_the_step_into_a_noop(pkg, syn, ss):
    """an overflow must be seen even when the host runs ahead, and must not corrupt the parameters."""
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    k = _keyframes(syn, pkg, 1)[0]
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
    ws._grow_isects(2048)                                  # far too small: every forward overflows
    ps = ss.PoseState(k["q"], k["t"], "cuda")
    pose0 = ps.pose.clone()
    import warnings
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        # enqueue without looking at the host mirror in between (check_overflow reads whatever has landed)
        L, st = pkg.lib(), C.c_void_p(torch.cuda.current_stream().cuda_stream)
        cws, cgs = ws.c_struct(), gmap.c_struct()
        cws.adam_state = gmap.adam_state.data_ptr()
        prm = ss._params(0, 3, 0.5, 1.0, 0.99, [1e-4, 2.5e-3, 1e-3, 0.05, 1e-3], 1e-15)
        fr = ws.frame(k["rgb"], k["depth"], k["pose"], k["inv_nd"])
        for _ in range(3):
            pkg._lib.check(L.isogs_map_step(C.byref(cam), fr.data_ptr(), C.byref(cws), C.byref(cgs), C.byref(prm), None, st))
        tprm = ss._params(0, 3, 0.5, 1.0, 0.99, [4e-4, 2e-3, 0, 0, 0], 1e-8)
        pkg._lib.check(L.isogs_track_step(C.byref(cam), fr.data_ptr(), C.byref(cws), C.byref(gmap.c_struct(False)),
                                          C.c_void_p(ps.pose.data_ptr()), C.c_void_p(ps.m.data_ptr()),
                                          C.c_void_p(ps.v.data_ptr()), C.c_void_p(ps.best.data_ptr()),
                                          C.c_void_p(ps.adam_state.data_ptr()), C.byref(tprm), None, st))
        ws.post_totals()
        torch.cuda.synchronize()
        tot = ws.totals.cpu().numpy()
        assert tot[1] == 1 and tot[0] > 2048, tot
        for name in ("means3D", "rgb_colors", "logit_opacities", "log_scales"):
            assert np.array_equal(gmap.view(name).cpu().numpy().reshape(-1), np.asarray(params[name]).reshape(-1)), name
        assert torch.equal(ps.pose, pose0) and float(ps.best[7]) == float(np.float32(1e20)), "pose / best candidate must be untouched"
        assert ws.check_overflow() is True                 # no event gating: the mirror carries the sticky word
        assert ws.isect_cap >= int(tot[0]) and int(ws.totals[1]) == 0
        assert any("overflowed" in str(x.message) for x in w)
    ss.map_step(ws, gmap, k["pose"], k["rgb"], k["depth"], k["inv_nd"])
    torch.cuda.synchronize()
    assert not ws.check_overflow(wait=True)
    assert np.abs(gmap.view("rgb_colors").cpu().numpy() - params["rgb_colors"]).max() > 0, "the step after growth must apply"


def test_outlier_mask_with_a_map_much_smaller_than_the_image(pkg, orc, syn, ss):
    """N << P -- the median sort must use its OWN temp storage (the intersection buffers of the forward
    are still needed by the backward).  Against the oracle's iteration with the same mask."""
    import oracle_slam as osl
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    sel = np.arange(0, params["means3D"].shape[0], 12)     # ~1 600 Gaussians for 19 200 pixels
    sub = {k: np.ascontiguousarray(v[sel]) for k, v in params.items()}
    sub["log_scales"] = sub["log_scales"] + np.float32(np.log(3.0))   # keep the frame covered
    cam, ocam = pkg.make_camera(W, H, fx, fy, cx, cy), orc.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(8, step=3.0)
    c, d, _ = syn.render_frame("tiny", traj[3], seed=3)
    q0, t0 = syn.w2c_to_pose(traj[2])
    gmap = ss.GaussianMap(sub, device="cuda")
    ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
    ws._grow_isects(12000)                                 # enough for the enlarged splats, still < P = 19 200
    assert ws.isect_cap < W * H
    vs_ptr, cap0 = ws.vs.data_ptr(), ws.isect_cap
    ps, ts = ss.PoseState(q0, t0, "cuda"), osl.TrackState(q0, t0)
    grgb, gdepth = t(c), t(d[0])
    for it in range(3):
        lo, gp = osl.track_iteration(orc, ocam, sub, ts, np.ascontiguousarray(c), np.ascontiguousarray(d[0]),
                                     ignore_outlier_depth=True)
        losses = ss.track_step(ws, gmap, ps, grgb, gdepth, ignore_outlier_depth=True)
        torch.cuda.synchronize()
        assert ws.vs.data_ptr() == vs_ptr and ws.isect_cap == cap0, "intersection buffers must not move mid-step"
        assert abs(float(losses[0]) - lo) <= 2e-4 * abs(lo), (it, float(losses[0]), lo)
        v = ws.v_pose.cpu().numpy()
        assert np.isfinite(v).all() and np.abs(v - gp).max() <= 1e-3 * np.abs(gp).max()


def test_unpack_rgbd_matches_the_reference_dataset_conversion(pkg):
    """basedataset.py:299-340: color / 255 -> [3,H,W]; depth / png_depth_scale."""
    W, H = 80, 56
    cam = pkg.make_camera(W, H, 60, 60, 39.5, 27.5)
    rng = np.random.RandomState(0)
    rgb8 = rng.randint(0, 256, (H, W, 3)).astype(np.uint8)
    d16 = rng.randint(0, 65536, (H, W)).astype(np.uint16)
    scale = 6553.5
    out_rgb = torch.empty((3, H, W), dtype=torch.float32, device="cuda")
    out_d = torch.empty((H, W), dtype=torch.float32, device="cuda")
    d_rgb8 = torch.from_numpy(rgb8).cuda()
    d_d16 = torch.from_numpy(d16.view(np.int16)).cuda()
    pkg._lib.check(pkg.lib().isogs_unpack_rgbd(C.byref(cam), C.c_void_p(d_rgb8.data_ptr()), C.c_void_p(d_d16.data_ptr()),
                                               1.0 / scale, C.c_void_p(out_rgb.data_ptr()), C.c_void_p(out_d.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    ref_rgb = (torch.from_numpy(rgb8).float() / 255.0).permute(2, 0, 1).numpy()
    assert np.array_equal(out_rgb.cpu().numpy(), ref_rgb)
    assert np.allclose(out_d.cpu().numpy(), d16.astype(np.float32) / np.float32(scale), rtol=2e-7, atol=0)


def test_dropin_and_fused_paths_share_one_screen_gradient_convention(pkg, syn, ss):
    """a12: means2D.grad of the drop-in rasterizer and the statistic the fused step accumulates are the SAME
    number (Inria NDC scaling v_xy * (W/2, H/2), the one grad_thresh = 2e-4 was tuned on), so both paths select the
    same clone / split sets."""
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    k = _keyframes(syn, pkg, 1)[0]
    gmap = ss.GaussianMap(params, device="cuda")
    ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
    N = gmap.count
    flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
    ss.render_forward(ws, gmap, pose=k["pose"])
    ss.loss_l1(ws, 1, k["rgb"], k["depth"], k["inv_nd"])
    ws.v_render[3:].zero_()                                   # colour render only (splatam.py:528-530)
    ss.render_backward(ws, gmap, pose=k["pose"])
    # fused path: statistics straight from v_rec (default scale)
    ss.accumulate_densify_stats(ws, gmap)
    fused = gmap.vars["means2D_gradient_accum"][:N, 0].clone()
    # drop-in path: the means2D.grad side channel (isogs_preprocess_bwd's v_means2d), accumulated as-is
    m2d = torch.empty((N, 3), dtype=torch.float32, device="cuda")
    p = lambda x: None if x is None else C.c_void_p(x.data_ptr())
    pkg._lib.check(pkg.lib().isogs_preprocess_bwd(C.byref(cam), C.byref(k["pose"]), None, N, p(gmap.p["means3D"]),
                                                  p(gmap.p["log_scales"]), 1, None, p(gmap.p["logit_opacities"]), 5, flags,
                                                  p(ws.radii), p(ws.v_rec), None, None, None, None, None, p(m2d), None, None,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    gmap.reset_stats()
    ss.accumulate_densify_stats(ws, gmap, v_xy=m2d)
    dropin = gmap.vars["means2D_gradient_accum"][:N, 0].clone()
    torch.cuda.synchronize()
    assert float(fused.max()) > 0
    assert torch.allclose(fused, dropin, rtol=1e-6, atol=0)
    thr = float(fused[fused > 0].median())
    assert torch.equal(fused >= thr, dropin >= thr), "the two paths must select the same Gaussians"


def test_fused_count_and_rect_binning_is_bit_identical(pkg, syn):
    """The fused step's binning -- per-tile counts accumulated by the projection kernel, scatter fed by the 16-byte
    (tile rect, depth bits) rows -- must produce exactly the lists of isogs_tile_bin (which are bit-exact against the
    oracle and the reference's own intersect kernels)."""
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
    means, ls = t(params["means3D"]), t(params["log_scales"])
    lo, rgb = t(params["logit_opacities"].reshape(-1)), t(params["rgb_colors"])
    q, tr = syn.w2c_to_pose(syn.trajectory(8, step=3.0)[4])
    pose = pkg.make_pose(q, tr)
    rec, radii, tiles, cum, M = pkg.raster.preprocess(cam, means, ls, None, lo, rgb, ch=5, flags=flags, pose=pose)
    ks, vs, ranges, _, _ = pkg.raster.tile_bin(cam, rec, radii, cum, M)
    N = means.shape[0]
    L, st = pkg.lib(), C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda x: None if x is None else C.c_void_p(x.data_ptr())
    rec2 = torch.empty_like(rec); radii2 = torch.empty_like(radii)
    rects = torch.empty((N, 4), dtype=torch.int32, device="cuda")
    _, sort_b = pkg._lib.temp_bytes(0, M + 5)
    tmp = torch.zeros((sort_b,), dtype=torch.uint8, device="cuda")          # counters (head of sort_temp) zeroed
    pkg._lib.check(L.isogs_preprocess_fwd(C.byref(cam), C.byref(pose), None, N, p(means), p(ls), 1, None, p(lo), p(rgb), 5,
                                          flags, p(rec2), p(radii2), None, None, None, 0, None, None, p(rects), p(tmp), st))
    bucket = torch.empty((M + 5,), dtype=torch.int64, device="cuda")
    ks2 = torch.empty((M + 5,), dtype=torch.int64, device="cuda")
    vs2 = torch.empty((M + 5,), dtype=torch.int32, device="cuda")
    rg2 = torch.empty_like(ranges)
    tot = torch.zeros((2,), dtype=torch.int32, device="cuda")
    pkg._lib.check(L.isogs_tile_bin_async(C.byref(cam), N, M + 5, p(rec2), p(radii2), p(bucket), p(ks2), p(vs2), p(rg2),
                                          p(tot), p(tmp), sort_b, p(rects), 1, st))
    torch.cuda.synchronize()
    assert torch.equal(rec2, rec) and torch.equal(radii2, radii)
    assert tot.tolist() == [M, 0]
    assert torch.equal(rg2, ranges) and torch.equal(vs2[:M], vs) and torch.equal(ks2[:M], ks)
    r = rects.cpu().numpy().view(np.uint32)
    n_tiles_per = ((r[:, 0] >> 16) - (r[:, 0] & 0xffff)) * ((r[:, 1] >> 16) - (r[:, 1] & 0xffff))
    assert np.array_equal(n_tiles_per.astype(np.int32), tiles.cpu().numpy())
    assert np.array_equal(r[:, 2][n_tiles_per > 0], rec.cpu().numpy()[:, 10].view(np.uint32)[n_tiles_per > 0])


def test_map_step_outlier_mask_and_bundle_adjust_modes_match_the_oracle(pkg, orc, syn, ss):
    """The two get_loss modes the shipped configs leave off (VERDICT r1 missing 8): ignore_outlier_depth_loss in
    mapping mode -- the depth term is the MEAN over the masked pixels, so its normaliser is computed on the device
    (splatam.py:545-549, 556-559) -- and do_ba=True, where the keyframe's pose gets a gradient and an Adam step next
    to the Gaussians (:505-510).  One iteration of each (and both together) against the oracle's iteration."""
    import oracle_slam as osl
    params, (W, H, fx, fy, cx, cy) = _scene(syn)
    cam, ocam = pkg.make_camera(W, H, fx, fy, cx, cy), orc.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(8, step=3.0)
    c, d, _ = syn.render_frame("tiny", traj[3], seed=3)
    c, d = np.ascontiguousarray(c), np.ascontiguousarray(d[0])
    d[40:60, 50:90] *= 0.5                               # a foreground object the map does not hold: depth outliers
    q, tr = syn.w2c_to_pose(traj[2])                     # the neighbouring frame's pose: a pose gradient exists
    inv_nd = 1.0 / max(1, int((d > 0).sum()))
    lrs = dict(ss.MAPPING_LRS, cam_unnorm_rots=4e-4, cam_trans=2e-3)
    grgb, gdepth = t(c), t(d)
    for outlier, ba in ((True, False), (False, True), (True, True)):
        gmap = ss.GaussianMap(params, device="cuda")
        ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
        state = osl.new_state(params)
        ts = osl.TrackState(q, tr) if ba else None
        ps = ss.PoseState(q, tr, "cuda") if ba else None
        og = {}
        lo = osl.map_iteration(orc, ocam, state, q, tr, c, d, grads_out=og, ignore_outlier_depth=outlier, ba=ts)
        losses = ss.map_step(ws, gmap, pkg.make_pose(q, tr), grgb, gdepth, inv_nd, lrs=lrs, ignore_outlier_depth=outlier,
                             ba=ps)
        torch.cuda.synchronize()
        assert not ws.check_overflow(wait=True)
        assert abs(float(losses[0]) - lo) <= 2e-4 * abs(lo), (outlier, ba, float(losses[0]), lo)
        for k in ("means3D", "rgb_colors", "logit_opacities", "log_scales"):
            got, want = gmap.g[k].cpu().numpy().reshape(-1), np.asarray(og[k]).reshape(-1)
            assert np.abs(got - want).max() <= 2e-3 * np.abs(want).max() + 1e-9, (outlier, ba, k)
        if ba:
            v = ws.v_pose.cpu().numpy()
            assert np.abs(v - og["pose"]).max() <= 1e-3 * np.abs(og["pose"]).max(), (v, og["pose"])
            got = ps.pose.cpu().numpy()
            assert np.abs(got - ts.pose).max() <= 0.02 * 4e-4, (got, ts.pose)      # one Adam step = lr * sign(g)
            assert np.abs(got - np.concatenate([q, tr])).max() > 1e-4, "the pose must have stepped"
        if outlier:   # the mask bites: the normaliser differs from 1 / n(valid depth)
            thr = orc.depth_outlier_threshold(orc.render_full(ocam, params["means3D"], params["log_scales"], None,
                                              params["logit_opacities"].reshape(-1), params["rgb_colors"], ch=5,
                                              flags=orc.F_LOG_SCALES | orc.F_LOGIT_OPAC, pose=orc.make_pose(q, tr))["render"][3], d)
            assert abs(float(ws.small[28]) - thr) <= 1e-5 * thr
            assert float(ws.outlier_scratch[0]) > inv_nd * 1.01
