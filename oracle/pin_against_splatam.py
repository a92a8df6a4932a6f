"""Pin the STEP around the rasterizer against the reference's own Python code.

Runs ONLY in the build container: it imports /root/reference/python_ordinary/scripts/splatam.py
and its helpers (utils/slam_helpers.py, utils/slam_external.py, utils/recon_helpers.py), which do
not exist on the GPU box.  Those files hard-code CUDA tensors and import a CUDA rasterizer, so

  * a TorchFunctionMode rewrites `.cuda()` / `device="cuda"` to the CPU,
  * the modules the reference imports but this image lacks (matplotlib, datasets.*, eval helpers)
    are stubbed -- none of them is on the step path,
  * `diff_gaussian_rasterization` is a stand-in whose forward/backward is the CPU ORACLE
    (oracle/isogs_oracle.cpp).  The rasterizer itself is pinned elsewhere (pin_against_gsplat.py,
    tests/test_gpu_vs_gsplat.py); what THIS script pins is everything the reference's Python does
    around it: get_pointcloud / initialize_params (splatam.py:296-399), setup_camera
    (recon_helpers.py:4-27), transform_to_frame + the two rendervar builders (slam_helpers.py:
    126-139, 218-252, 255-303), get_loss (splatam.py:494-760: masks, sum- vs mean-L1, weights),
    the autograd chain back to cam_unnorm_rots / cam_trans and to the Gaussian parameters, and
    torch.optim.Adam as initialize_optimizer configures it (splatam.py:400-406).

Outputs tests/golden/splatam_step_*.npz (inputs + the reference's losses, gradients and updated
parameters).  tests/test_oracle_golden.py repeats the oracle-side comparison anywhere;
tests/test_gpu_step.py compares track_step / map_step / add_new_gaussians on the GPU.

The mapping iterations use the reference's full colour term 0.8*L1 + 0.2*(1-SSIM) (splatam.py:568,
calc_ssim of slam_external.py:54-97) and the IsoGS terms 50*flat + 2*iso (:572-675,733-739; N < 8192, so every
centre is a query) -- i.e. the reference's complete mapping loss.  compute_iso_surface_loss_sampled / compute_flat_loss
are also recorded on their own on 600 random three-axis Gaussians with a 150-query subset (geo_*).  calc_ssim and its autograd gradient are also recorded on
their own (ssim_value / ssim_grad for x = rgb0, y = rgb1).

Usage:  python oracle/pin_against_splatam.py
"""
import os
import sys
import types
from typing import NamedTuple
from unittest import mock

import numpy as np
import torch
from torch.overrides import TorchFunctionMode

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

REF = "/root/reference/python_ordinary"


class CpuShim(TorchFunctionMode):
    """`.cuda()` -> identity, `device="cuda"` -> CPU, for the reference's hard-coded device strings."""

    def __torch_function__(self, func, types_, args=(), kwargs=None):
        kwargs = dict(kwargs or {})
        if getattr(func, "__name__", "") == "cuda" and args and torch.is_tensor(args[0]):
            return args[0]
        dev = kwargs.get("device")
        if dev is not None and str(dev).startswith("cuda"):
            kwargs["device"] = "cpu"
        return func(*args, **kwargs)


# ---------------------------------------------------------------------------------------------
# CPU stand-in for the `diff_gaussian_rasterization` module, backed by the oracle
# ---------------------------------------------------------------------------------------------
class GaussianRasterizationSettings(NamedTuple):
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


def oracle_camera(s):
    W, H = int(s.image_width), int(s.image_height)
    w2c = s.viewmatrix.detach().reshape(4, 4).t().double()
    P = s.projmatrix.detach().reshape(4, 4).t().double() @ torch.linalg.inv(w2c)
    fx, fy = W / (2.0 * float(s.tanfovx)), H / (2.0 * float(s.tanfovy))
    cx, cy = W * (1.0 + float(P[0, 2])) / 2.0, H * (1.0 + float(P[1, 2])) / 2.0
    return orc.make_camera(W, H, fx, fy, cx, cy, w2c.float().numpy(), near=0.01, far=100.0)


class _OracleRasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, colors, opacities, scales, rotations, cam):
        a = [t.detach().numpy().astype(np.float32) for t in (means3D, scales, rotations, opacities.reshape(-1), colors)]
        r = orc.render_full(cam, *a, ch=4, flags=0)
        ctx.cam, ctx.a, ctx.r, ctx.oshape = cam, a, r, opacities.shape
        radii = torch.from_numpy(r["radii"].max(axis=1).astype(np.int32))
        ctx.mark_non_differentiable(radii)
        ren = torch.from_numpy(r["render"])
        return ren[:3].clone(), radii, ren[3:4].clone()

    @staticmethod
    def backward(ctx, v_color, v_radii, v_depth):
        cam, a, r = ctx.cam, ctx.a, ctx.r
        H, W = cam.height, cam.width
        v_render = np.zeros((4, H, W), np.float32)
        if v_color is not None:
            v_render[:3] = v_color.numpy()
        if v_depth is not None:
            v_render[3:4] = v_depth.numpy()
        v_rec = orc.blend_bwd(cam, 4, r["rec"], r["vals"], r["ranges"], r["alpha"], r["last"], v_render)
        g = orc.preprocess_bwd(cam, *a, r["radii"], r["rec"], v_rec, ch=4, flags=0)
        t = torch.from_numpy
        # means2D.grad as the Inria rasterizer writes it: the screen-space gradient in NDC units,
        # v_xy * (W/2, H/2) -- the one convention of the repo (csrc/preprocess.cu, slam_step.py)
        v_m2d = torch.zeros((a[0].shape[0], 3))
        v_m2d[:, 0] = t(v_rec[:, 0].copy()) * (0.5 * W)
        v_m2d[:, 1] = t(v_rec[:, 1].copy()) * (0.5 * H)
        return (t(g["means"]), v_m2d, t(g["colors"]), t(g["opacities"]).reshape(ctx.oshape), t(g["scales"]),
                t(g["quats"]), None)


class GaussianRasterizer(torch.nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, colors_precomp=None, scales=None, rotations=None, **_):
        cam = oracle_camera(self.raster_settings)
        return _OracleRasterize.apply(means3D, means2D, colors_precomp, opacities, scales, rotations, cam)


def load_reference():
    """Import splatam.py with the off-path imports stubbed; returns the module."""
    stub = types.ModuleType("diff_gaussian_rasterization")
    stub.GaussianRasterizationSettings = GaussianRasterizationSettings
    stub.GaussianRasterizer = GaussianRasterizer
    sys.modules["diff_gaussian_rasterization"] = stub
    for name in ("matplotlib", "matplotlib.pyplot", "wandb", "datasets", "datasets.gradslam_datasets",
                 "utils.eval_helpers", "imageio", "natsort", "kornia", "open3d",
                 "pytorch_msssim", "lpips", "torchmetrics", "plyfile"):
        if name not in sys.modules:
            sys.modules[name] = mock.MagicMock(name=name)
    sys.path.insert(0, REF)
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_splatam", f"{REF}/scripts/splatam.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


# ---------------------------------------------------------------------------------------------
# synthetic RGB-D frames (analytic, so the fixture is reproducible without a dataset)
# ---------------------------------------------------------------------------------------------
W, H, FX, FY, CX, CY = 64, 48, 60.0, 60.0, 31.5, 23.5


def frame(shift):
    """Colour [3,H,W] in [0,1] and depth [1,H,W] (metres; a patch of invalid zeros)."""
    y, x = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    x, y = x + shift[0], y + shift[1]
    rgb = np.stack([0.5 + 0.4 * np.sin(x / 5.0) * np.cos(y / 9.0), 0.5 + 0.4 * np.cos(x / 7.0 + y / 11.0),
                    0.5 + 0.4 * np.sin(y / 4.0)]).astype(np.float32)
    d = (2.0 + 0.3 * np.sin(x / 11.0) + 0.2 * np.cos(y / 7.0) + shift[2]).astype(np.float32)
    d[10:15, 40:47] = 0.0
    if shift[2] != 0.0:
        d[30:36, 10:16] -= 0.8                  # a new foreground object: depth-error outliers for :551-554
    return torch.from_numpy(rgb), torch.from_numpy(d)[None]


def grads_of(params, names):
    return {k: (params[k].grad.detach().numpy().copy() if params[k].grad is not None else None) for k in names}


GAUSS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
MAP_LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=0.05, log_scales=1e-3,
               cam_unnorm_rots=0.0, cam_trans=0.0)
TRACK_LRS = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
                 cam_unnorm_rots=4e-4, cam_trans=2e-3)


def main():
    torch.manual_seed(0)
    out = {}
    with CpuShim():
        ref = load_reference()
        K = torch.tensor([[FX, 0, CX], [0, FY, CY], [0, 0, 1]], dtype=torch.float32)
        w2c0 = torch.eye(4)
        rgb0, d0 = frame((0.0, 0.0, 0.0))
        rgb1, d1 = frame((1.5, -1.3, 0.02))    # -1.3: no mirror symmetry in y, hence no exact depth ties
        cam = ref.setup_camera(W, H, K.numpy(), w2c0.numpy())

        # --- initialisation (initialize_first_timestep body, splatam.py:427-436) ---
        mask = (d0 > 0).reshape(-1)
        pt, msd = ref.get_pointcloud(rgb0, d0, K, w2c0, mask=mask, compute_mean_sq_dist=True,
                                     mean_sq_dist_method="projective")
        # "isotropic" is what the reference configs ask for; IsoGS turns it into three axes + N(0, 0.01) noise
        # (splatam.py:361-368), seeded by torch.manual_seed(0) above
        params, variables = ref.initialize_params(pt, 3, msd, "isotropic")
        variables["scene_radius"] = torch.max(d0) / 3.0
        out.update(rgb0=rgb0.numpy(), d0=d0.numpy(), rgb1=rgb1.numpy(), d1=d1.numpy(),
                   intr=np.array([W, H, FX, FY, CX, CY], np.float32))
        for k in GAUSS:
            out[f"init_{k}"] = params[k].detach().numpy().copy()

        # --- tracking: 3 iterations on frame 1 from a perturbed pose (splatam.py:1237-1290) ---
        with torch.no_grad():
            params["cam_unnorm_rots"][..., 1] = torch.tensor([[1.1, 0.011, -0.022, 0.0055]])
            params["cam_trans"][..., 1] = torch.tensor([[0.02, -0.01, 0.03]])
        out["track_pose0"] = np.concatenate([params["cam_unnorm_rots"][0, :, 1].detach().numpy(),
                                             params["cam_trans"][0, :, 1].detach().numpy()])
        curr = dict(cam=cam, im=rgb1, depth=d1, id=1, intrinsics=K, w2c=w2c0, iter_gt_w2c_list=[w2c0, w2c0])
        # one evaluation with ignore_outlier_depth_loss=True (splatam.py:551-554), no optimizer step
        loss, variables, losses = ref.get_loss(params, curr, variables, 1, dict(im=0.5, depth=1.0), True, 0.99, True,
                                               True, tracking=True)
        loss.backward()
        out["track_outlier_losses"] = np.array([float(loss), float(losses["im"]), float(losses["depth"])], np.float64)
        out["track_outlier_grad"] = np.concatenate([params["cam_unnorm_rots"].grad[0, :, 1].numpy(),
                                                    params["cam_trans"].grad[0, :, 1].numpy()])
        params["cam_unnorm_rots"].grad = None
        params["cam_trans"].grad = None
        opt = ref.initialize_optimizer(params, TRACK_LRS, tracking=True)
        tl, tg, tp = [], [], []
        for it in range(3):
            loss, variables, losses = ref.get_loss(params, curr, variables, 1, dict(im=0.5, depth=1.0), True, 0.99,
                                                   True, False, tracking=True)
            loss.backward()
            tl.append([float(loss), float(losses["im"]), float(losses["depth"])])
            tg.append(np.concatenate([params["cam_unnorm_rots"].grad[0, :, 1].numpy(),
                                      params["cam_trans"].grad[0, :, 1].numpy()]))
            opt.step()
            opt.zero_grad(set_to_none=True)
            tp.append(np.concatenate([params["cam_unnorm_rots"][0, :, 1].detach().numpy(),
                                      params["cam_trans"][0, :, 1].detach().numpy()]))
        out.update(track_losses=np.array(tl, np.float64), track_grads=np.array(tg, np.float32),
                   track_poses=np.array(tp, np.float32))

        # --- constant-velocity initialisation of frame 2 (splatam.py:844-863) ---
        ref.initialize_camera_pose(params, 2, forward_prop=True)
        out["init_pose2"] = np.concatenate([params["cam_unnorm_rots"][0, :, 2].detach().numpy(),
                                            params["cam_trans"][0, :, 2].detach().numpy()])

        # --- densification of frame 1 at the tracked pose (splatam.py:799-841) ---
        n0 = params["means3D"].shape[0]
        params, variables = ref.add_new_gaussians(params, variables, curr, 0.5, 1, "projective", "anisotropic")
        out["added"] = np.array([params["means3D"].shape[0] - n0], np.int64)
        for k in GAUSS:
            out[f"new_{k}"] = params[k].detach().numpy()[n0:].copy()

        # --- mapping: 2 iterations on frame 1, L1 terms only (splatam.py:1406-1476) ---
        x = rgb0.clone().requires_grad_(True)
        s_val = ref.calc_ssim(x, rgb1)
        s_val.backward()
        out.update(ssim_value=np.array([float(s_val)], np.float64), ssim_grad=x.grad.numpy().copy())
        for k in GAUSS:
            out[f"map_in_{k}"] = params[k].detach().numpy().copy()
        opt = ref.initialize_optimizer(params, MAP_LRS, tracking=False)
        weights = dict(im=0.5, depth=1.0)               # get_loss adds flat = 50, iso = 2 (splatam.py:733-739)
        ml = []
        for it in range(2):
            loss, variables, losses = ref.get_loss(params, curr, variables, 1, weights, True, 0.99, True, False,
                                                   mapping=True)
            loss.backward()
            ml.append([float(loss), float(losses["im"]), float(losses["depth"]), float(losses["flat"]),
                       float(losses["iso"])])
            if it == 0:
                for k, g in grads_of(params, GAUSS).items():
                    out[f"map_grad_{k}"] = g
                out["map_means2D_grad"] = variables["means2D"].grad.numpy().copy()
            opt.step()
            opt.zero_grad(set_to_none=True)
        out["map_losses"] = np.array(ml, np.float64)
        for k in GAUSS:
            out[f"map_out_{k}"] = params[k].detach().numpy().copy()

        # --- pruning (slam_external.py:162-185) ---
        with torch.no_grad():
            params["logit_opacities"][::7] = -6.0               # sigmoid < 0.005
            params["log_scales"][5::13] = 0.0                    # exp = 1 > 0.1 * scene_radius
        out["prune_in_logit_opacities"] = params["logit_opacities"].detach().numpy().copy()
        out["prune_in_log_scales"] = params["log_scales"].detach().numpy().copy()
        pd = dict(start_after=0, remove_big_after=0, stop_after=20, prune_every=20, removal_opacity_threshold=0.005,
                  final_removal_opacity_threshold=0.005, reset_opacities=False, reset_opacities_every=500)
        out["scene_radius"] = np.array([float(variables["scene_radius"])], np.float32)
        params, variables = ref.prune_gaussians(params, variables, opt, 0, pd)
        out["prune_out_means3D"] = params["means3D"].detach().numpy().copy()
        # --- gradient-driven densification (slam_external.py:100-104,191-243) on isotropic [N,1] scales, the
        # only layout the reference's `stds.repeat(n, 3)` supports; torch.normal's draw is captured so that the
        # device kernel can be fed the same standard normals ---
        import utils.slam_external as se
        g = torch.Generator().manual_seed(7)
        Nd = 300
        dp = dict(means3D=torch.randn(Nd, 3, generator=g), rgb_colors=torch.rand(Nd, 3, generator=g),
                  unnorm_rotations=torch.randn(Nd, 4, generator=g),
                  logit_opacities=3.0 * torch.randn(Nd, 1, generator=g),
                  log_scales=torch.log(0.005 + 0.075 * torch.rand(Nd, 1, generator=g)),
                  cam_unnorm_rots=torch.tensor([[[1.0], [0.0], [0.0], [0.0]]]), cam_trans=torch.zeros(1, 3, 1))
        dp = {k: torch.nn.Parameter(v.float().contiguous().requires_grad_(True)) for k, v in dp.items()}
        dopt = ref.initialize_optimizer(dp, MAP_LRS, tracking=False)
        for k in GAUSS:                                        # one Adam step so that the moments are non-trivial
            dp[k].grad = 1e-3 * torch.randn(dp[k].shape, generator=g)
        dopt.step()
        dopt.zero_grad(set_to_none=True)
        for k in GAUSS:
            out[f"dens_in_{k}"] = dp[k].detach().numpy().copy()
            st = dopt.state[dp[k]]
            out[f"dens_in_m_{k}"] = st["exp_avg"].numpy().copy()
            out[f"dens_in_v_{k}"] = st["exp_avg_sq"].numpy().copy()
        m2d = torch.zeros(Nd, 3, requires_grad=True)
        m2d.grad = torch.cat([8e-4 * torch.randn(Nd, 2, generator=g), torch.zeros(Nd, 1)], dim=1)
        seen = torch.rand(Nd, generator=g) < 0.8
        dvars = dict(means2D=m2d, seen=seen, scene_radius=torch.tensor(2.0),
                     means2D_gradient_accum=3e-4 * torch.rand(Nd, generator=g),
                     denom=torch.randint(0, 3, (Nd,), generator=g).float(), max_2D_radius=torch.zeros(Nd))
        # (no 'timestep' entry: the reference's densify does not extend it and remove_points then fails, :156-157)
        out.update(dens_v_xy=m2d.grad.numpy().copy(), dens_seen=seen.numpy().copy(),
                   dens_accum0=dvars["means2D_gradient_accum"].numpy().copy(), dens_denom0=dvars["denom"].numpy().copy())
        dvars = se.accumulate_mean2d_gradient(dvars)            # a first, separate accumulation (a12)
        out.update(dens_accum1=dvars["means2D_gradient_accum"].numpy().copy(), dens_denom1=dvars["denom"].numpy().copy())
        captured = {}
        real_normal = torch.normal

        def capture_normal(mean, std, **kw):
            r = real_normal(mean=mean, std=std, **kw)
            captured["z"] = (r / std).detach().numpy().copy()
            return r
        torch.normal = capture_normal
        ddict = dict(start_after=0, remove_big_after=0, stop_after=1000, densify_every=1, grad_thresh=0.0002,
                     num_to_split_into=2, removal_opacity_threshold=0.005, final_removal_opacity_threshold=0.005,
                     reset_opacities=False, reset_opacities_every=3000)
        try:
            dp, dvars = ref.densify(dp, dvars, dopt, 1, ddict)
        finally:
            torch.normal = real_normal
        out["dens_noise"] = captured["z"].astype(np.float32)
        for k in GAUSS:
            out[f"dens_out_{k}"] = dp[k].detach().numpy().copy()
        st = dopt.state[dp["means3D"]]
        out["dens_out_m_means3D"] = st["exp_avg"].numpy().copy()
        out["dens_out_accum"] = dvars["means2D_gradient_accum"].numpy().copy()
        # --- keyframe selection by overlap (utils/keyframe_selection.py:40-95); the sampled pixel indices and the
        # per-keyframe percentages are captured (torch.randint wrapped, module-level `sorted` shadowed) ---
        import builtins
        import utils.keyframe_selection as ks
        cap = {}
        real_randint = torch.randint

        def capture_randint(*a, **kw):
            r = real_randint(*a, **kw)
            cap["indices"] = r.clone()
            return r

        def capture_sorted(lst, **kw):
            cap["percent"] = [float(d["percent_inside"]) for d in lst]
            return builtins.sorted(lst, **kw)
        cur_w2c = torch.eye(4)
        cur_w2c[:3, :3] = ref.build_rotation(torch.nn.functional.normalize(torch.tensor([[1.0, 0.02, -0.03, 0.01]])))[0]
        cur_w2c[:3, 3] = torch.tensor([0.05, -0.02, 0.1])
        kfs = []
        for j in range(6):                                      # from nearly identical views to one looking away
            m = torch.eye(4)
            q = torch.nn.functional.normalize(torch.tensor([[1.0, 0.03 * j, 0.12 * j, -0.02 * j]]))
            m[:3, :3] = ref.build_rotation(q)[0]
            m[:3, 3] = torch.tensor([0.1 * j, -0.05 * j, 0.15 * j])
            kfs.append({"est_w2c": m})
        m = torch.eye(4)
        m[:3, :3] = torch.diag(torch.tensor([-1.0, 1.0, -1.0]))  # looking backwards: no overlap
        kfs.append({"est_w2c": m})
        torch.randint, ks.sorted = capture_randint, capture_sorted
        try:
            sel = ks.keyframe_selection_overlap(d1, cur_w2c, K, kfs, len(kfs), pixels=400)
        finally:
            torch.randint = real_randint
            del ks.sorted
        valid = torch.stack(torch.where(d1[0] > 0), dim=1)[cap["indices"]]
        out.update(kf_cur_w2c=cur_w2c.numpy().copy(), kf_w2cs=np.stack([k_["est_w2c"].numpy() for k_ in kfs]),
                   kf_sample_idx=(valid[:, 0] * W + valid[:, 1]).numpy().astype(np.int32),
                   kf_percent=np.array(cap["percent"], np.float32), kf_selected=np.array(sorted(int(i) for i in sel)))
        # --- the IsoGS geometric terms on their own: preamble of get_loss (splatam.py:578-640) + the reference's
        # compute_flat_loss / compute_iso_surface_loss_sampled, on a wavy sheet of 600 three-axis Gaussians ---
        g = torch.Generator().manual_seed(11)
        Ng = 600
        uv = 2.0 * torch.rand(Ng, 2, generator=g) - 1.0
        gm_ = torch.stack([uv[:, 0], uv[:, 1], 2.0 + 0.2 * torch.sin(3 * uv[:, 0]) * torch.cos(2 * uv[:, 1])], 1)
        gm_ = (gm_ + 0.01 * torch.randn(Ng, 3, generator=g)).requires_grad_(True)
        gls = (np.log(0.06) + 0.4 * torch.randn(Ng, 3, generator=g)).requires_grad_(True)
        gq_ = torch.randn(Ng, 4, generator=g).requires_grad_(True)
        glo = torch.randn(Ng, 1, generator=g).requires_grad_(True)
        gidx = torch.randperm(Ng, generator=g)[:150]
        Fn = torch.nn.functional
        scales = torch.clamp(torch.exp(gls), min=1e-5)
        Rm = ref.build_rotation(Fn.normalize(gq_))
        S_inv_sq = 1.0 / (scales ** 2 + 1e-8)
        inv_cov = torch.bmm(torch.bmm(Rm, torch.diag_embed(S_inv_sq)), Rm.transpose(1, 2))
        l_iso, dens = ref.compute_iso_surface_loss_sampled(query_points=gm_[gidx], means=gm_, inverse_covariances=inv_cov,
                                                           opacities=torch.sigmoid(glo), K=16, target_saturation=1.0,
                                                           chunk_size=128)
        l_flat = ref.compute_flat_loss(scales)
        (50.0 * l_flat + 2.0 * l_iso).backward()
        out.update(geo_means3D=gm_.detach().numpy().copy(), geo_log_scales=gls.detach().numpy().copy(),
                   geo_unnorm_rotations=gq_.detach().numpy().copy(), geo_logit_opacities=glo.detach().numpy().copy(),
                   geo_query_idx=gidx.numpy().astype(np.int32),
                   geo_losses=np.array([float(l_flat), float(l_iso), float(dens.mean())], np.float64),
                   geo_grad_means3D=gm_.grad.numpy().copy(), geo_grad_log_scales=gls.grad.numpy().copy(),
                   geo_grad_unnorm_rotations=gq_.grad.numpy().copy(), geo_grad_logit_opacities=glo.grad.numpy().copy())
    out["notes"] = np.array("tracking: get_loss(tracking=True, use_sil_for_loss=True, sil_thres=0.99, im=0.5, depth=1); "
                            "mapping: get_loss(mapping=True), colour term 0.8*L1 + 0.2*(1-SSIM), im=0.5, depth=1, flat=50, iso=2; "
                            "rasterizer = CPU oracle behind the reference's diff_gaussian_rasterization interface")
    path = os.path.join(ROOT, "tests", "golden", "splatam_step_64x48.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: getattr(v, "shape", None) for k, v in out.items() if k.startswith(("track", "map_l", "added"))})
    print("track losses", out["track_losses"], "\nmap losses", out["map_losses"], "added", out["added"])


if __name__ == "__main__":
    main()
