"""MEASUREMENT (build container only: imports /root/reference): SURVEY 8(d) CPU baselines (2) and (3), the reference's OWN
CPU-runnable code for the hot path, timed on this container's host cores and recorded under profiles/.

 (2) gsplat `_torch_impl._fully_fused_projection` + `_isect_tiles` + `_isect_offset_encode`
     (cpp/third_party/gsplat/gsplat/cuda/_torch_impl.py:286,378,477) on a 30 K-Gaussian subsample of the Replica-shaped
     bench map; `_isect_tiles` is a Python per-Gaussian loop, so the measured time is also extrapolated LINEARLY to the
     full 1 M map and labelled as an extrapolation.
 (3) BASELINE.json configs[0]: the reference's `get_pointcloud` + `initialize_params` (initialize_first_timestep body),
     `get_loss(mapping=True)` -> `loss.backward()` -> `torch.optim.Adam.step()` (python_ordinary/scripts/splatam.py:
     411-453,494-760,402-408) under the CPU device shim at 640x480 / one Gaussian per valid pixel, the un-vendored CUDA
     rasterizer replaced by the oracle-backed stand-in of oracle/pin_against_splatam.py.  Isotropic [N,1] scales as
     in BASELINE (flat / iso terms are 0 for them by splatam.py:599-601,673-675).

  python scripts/cpu_baselines_reference.py [profiles/r2_cpu_baselines_reference.json]
"""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402
import pin_against_gsplat as pg  # noqa: E402
import pin_against_splatam as ps  # noqa: E402

cores = os.cpu_count()
torch.set_num_threads(cores)
out = {"host": {"cores": cores, "where": "build container (no GPU); NOT the GPU box"}}

# ---- (2) projection + tile intersection through gsplat's pure-torch implementation -------------------------------
ti = pg.load_ref()
syn, params, kfs = bench.build_workload("replica", 1_000_000, 2)
W, H, fx, fy, cx, cy = syn.SHAPES["replica"][:6]
n_sub = 30_000
sel = np.sort(np.random.RandomState(0).choice(params["means3D"].shape[0], n_sub, replace=False))
means = torch.from_numpy(params["means3D"][sel])
scales = torch.from_numpy(np.exp(params["log_scales"][sel])).repeat(1, 3)
quats = torch.zeros((n_sub, 4)); quats[:, 0] = 1.0
covars, _ = ti._quat_scale_to_covar_preci(quats, scales, True, False, triu=False)
k = kfs[1]
q = np.asarray(k["q"], np.float64); q /= np.linalg.norm(q)
w, x, y, z = q
R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
              [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
              [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
vm = torch.eye(4); vm[:3, :3] = torch.from_numpy(R).float(); vm[:3, 3] = torch.from_numpy(np.asarray(k["t"], np.float32))
Ks = torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=torch.float32)
t0 = time.perf_counter()
radii, means2d, depths, conics, comp = ti._fully_fused_projection(means, covars, vm[None], Ks[None], W, H, 0.3, 0.01, 100.0)
t_proj = time.perf_counter() - t0
tw, th = (W + 15) // 16, (H + 15) // 16
t0 = time.perf_counter()
tiles_per_gauss, isect_ids, flatten_ids = ti._isect_tiles(means2d, radii, depths, 16, tw, th, sort=True)
t_isect = time.perf_counter() - t0
t0 = time.perf_counter()
offs = ti._isect_offset_encode(isect_ids, 1, tw, th)
t_off = time.perf_counter() - t0
tot = t_proj + t_isect + t_off
out["torch_impl_projection_isect"] = {
    "what": "gsplat _torch_impl: _fully_fused_projection + _isect_tiles + _isect_offset_encode, Replica-shaped 1200x680 view",
    "n_gaussians": n_sub, "n_isects": int(isect_ids.numel()),
    "seconds": {"projection": t_proj, "isect_tiles": t_isect, "isect_offset_encode": t_off, "total": tot},
    "extrapolated_to_1M_gaussians_seconds": tot * (1_000_000 / n_sub),
    "extrapolation": "LINEAR in the Gaussian count (the Python loop of _isect_tiles dominates); not a measurement",
}

# ---- (3) BASELINE configs[0] through the reference's own Python under the CPU shim --------------------------------
from oracle import oracle as orc  # noqa: E402
c0, d0, K = syn.render_frame("tum", syn.trajectory(2)[0], seed=0)
Wt, Ht, fxt, fyt, cxt, cyt = syn.SHAPES["tum"][:6]
with ps.CpuShim():
    ref = ps.load_reference()
    Kt = torch.from_numpy(K)
    w2c0 = torch.eye(4)
    rgb0, dep0 = torch.from_numpy(c0), torch.from_numpy(d0)
    cam = ref.setup_camera(Wt, Ht, K, w2c0.numpy())
    t0 = time.perf_counter()
    mask = (dep0 > 0).reshape(-1)
    pt, msd = ref.get_pointcloud(rgb0, dep0, Kt, w2c0, mask=mask, compute_mean_sq_dist=True, mean_sq_dist_method="projective")
    torch.manual_seed(0)
    prm, variables = ref.initialize_params(pt, 2, msd, "isotropic")
    # BASELINE benchmarks the isotropic [N,1] layout (slam_helpers.py:108-111): undo IsoGS's three-axis tiling
    prm["log_scales"] = torch.nn.Parameter(prm["log_scales"].detach()[:, :1].contiguous().requires_grad_(True))
    variables["scene_radius"] = torch.max(dep0) / 3.0
    t_init = time.perf_counter() - t0
    opt = ref.initialize_optimizer(prm, ps.MAP_LRS, tracking=False)
    curr = dict(cam=cam, im=rgb0, depth=dep0, id=0, intrinsics=Kt, w2c=w2c0, iter_gt_w2c_list=[w2c0])
    times = []
    for it in range(3):
        t0 = time.perf_counter()
        loss, variables, losses = ref.get_loss(prm, curr, variables, 0, dict(im=0.5, depth=1.0), True, 0.99, True, False,
                                               mapping=True)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        times.append(time.perf_counter() - t0)
out["configs0_reference_python_cpu"] = {
    "what": "BASELINE configs[0]: get_pointcloud + initialize_params, then get_loss(mapping=True) + backward + Adam.step through "
            "the reference's Python (CPU shim, rasterizer = oracle-backed stand-in), 640x480",
    "n_gaussians": int(prm["means3D"].shape[0]), "init_seconds": t_init, "step_seconds": times,
    "iters_per_s": 1.0 / float(np.median(times)), "oracle_threads": orc.lib().orc_num_threads(),
    "loss_first": float(loss),
}
dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r2_cpu_baselines_reference.json")
with open(dst, "w") as f:          # the reference prints to stdout while it runs: write the record to a file
    json.dump(out, f, indent=1)
print("wrote", dst, file=sys.stderr)
