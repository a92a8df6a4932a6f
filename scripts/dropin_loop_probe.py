"""What a user of the UNCHANGED reference loop gets: one mapping iteration written the reference's way
(transform_to_frame, two render-variable dicts, TWO `GaussianRasterizer` calls, masked L1 in torch ops, autograd
backward, torch.optim.Adam with seven parameter groups -- the flow of python_ordinary/scripts/splatam.py:494-570,
1412-1476) on top of the drop-in `diff_gaussian_rasterization`, timed beside the fused `slam_step.map_step`.
Prints one JSON line."""
import importlib
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import isogs_b200  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402

pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")


def settings(W, H, fx, fy, cx, cy, near=0.01, far=100.0):
    w2c = torch.eye(4, device="cuda")
    P = torch.tensor([[2 * fx / W, 0, -(W - 2 * cx) / W, 0], [0, 2 * fy / H, -(H - 2 * cy) / H, 0],
                      [0, 0, far / (far - near), -(far * near) / (far - near)], [0, 0, 1, 0]], device="cuda")
    vm = w2c.unsqueeze(0).transpose(1, 2)
    return GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=W / (2 * fx), tanfovy=H / (2 * fy),
                                         bg=torch.zeros(3, device="cuda"), scale_modifier=1.0, viewmatrix=vm,
                                         projmatrix=vm.bmm(P.unsqueeze(0).transpose(1, 2)), sh_degree=0,
                                         campos=torch.zeros(3, device="cuda"), prefiltered=False)


def rot_of(q):
    w, x, y, z = q.unbind(-1)
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(3, 3)


def _ev(marks, name):
    if marks is not None:
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        marks.append((name, e))


def iteration(p, cam, q_cam, t_cam, rgb, depth, opt, marks=None):
    _ev(marks, "begin")
    R = rot_of(F.normalize(q_cam, dim=0))
    pts = p["means3D"] @ R.T + t_cam
    common = dict(means3D=pts, rotations=F.normalize(p["unnorm_rotations"]), opacities=torch.sigmoid(p["logit_opacities"]),
                  scales=torch.exp(torch.tile(p["log_scales"], (1, 3))))
    m2d = torch.zeros_like(pts, requires_grad=True) + 0
    _ev(marks, "transform+activations")
    im, radius, _ = GaussianRasterizer(raster_settings=cam)(means2D=m2d, colors_precomp=p["rgb_colors"], **common)
    _ev(marks, "render colour fwd")
    z = pts[:, 2:3]
    ds, _, _ = GaussianRasterizer(raster_settings=cam)(means2D=torch.zeros_like(pts, requires_grad=True) + 0,
                                                       colors_precomp=torch.cat([z, torch.ones_like(z), z * z], 1), **common)
    _ev(marks, "render depth/sil fwd")
    d = ds[0:1]
    mask = (depth > 0) & ~torch.isnan(d)
    loss = 0.5 * torch.abs(im - rgb).mean() + torch.abs(depth - d)[mask].mean()
    _ev(marks, "loss")
    loss.backward()
    _ev(marks, "backward (2 renders + torch ops)")
    opt.step()
    opt.zero_grad(set_to_none=True)
    _ev(marks, "Adam")
    return loss


def main(steps=30):
    syn, params, kfs = bench.build_workload(bench.SHAPE, bench.N_GAUSS, bench.WINDOW)
    W, H, fx, fy, cx, cy = syn.SHAPES[bench.SHAPE][:6]
    dev = torch.device("cuda", 0)
    p = {k: torch.nn.Parameter(torch.from_numpy(np.ascontiguousarray(v)).float().to(dev)) for k, v in params.items()}
    lrs = ss.MAPPING_LRS
    opt = torch.optim.Adam([{"params": [v], "name": k, "lr": lrs[k]} for k, v in p.items()], lr=0.0, eps=1e-15)
    cam = settings(W, H, fx, fy, cx, cy)
    kd = [dict(q=torch.tensor(k["q"], device=dev), t=torch.tensor(k["t"], device=dev),
               rgb=torch.from_numpy(k["rgb"]).to(dev), depth=torch.from_numpy(k["depth"]).to(dev)[None]) for k in kfs]
    for i in range(5):
        k = kd[i % len(kd)]
        iteration(p, cam, k["q"], k["t"], k["rgb"], k["depth"], opt)
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        k = kd[i % len(kd)]
        iteration(p, cam, k["q"], k["t"], k["rgb"], k["depth"], opt)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    stage = {}
    for i in range(5):
        marks = []
        k = kd[i % len(kd)]
        iteration(p, cam, k["q"], k["t"], k["rgb"], k["depth"], opt, marks)
        torch.cuda.synchronize()
        for (n0, a), (n1, b) in zip(marks[:-1], marks[1:]):
            stage[n1] = stage.get(n1, 0.0) + a.elapsed_time(b) / 5
    print(json.dumps(dict(workload=f"{bench.SHAPE} {W}x{H}, N={bench.N_GAUSS}", dropin_reference_style_iters_per_s=1e3 / ms,
                          ms_per_iter=ms, stage_ms={k: round(v, 3) for k, v in stage.items()}, peak_mem_gb=torch.cuda.max_memory_allocated() / 2 ** 30,
                          note="two GaussianRasterizer calls + torch-op loss + autograd + torch.optim.Adam per iteration "
                               "(the reference loop's structure); compare bench.py `value` for the fused map_step")))


if __name__ == "__main__":
    main()
