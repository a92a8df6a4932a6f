"""Times the REFERENCE's own rasterization kernels (gsplat 1.5.3 built for sm_100a by
oracle/build_ref.sh, --use_fast_math as the reference builds them) next to the B200-native blend kernels on
the bench workload: same projected records, same tile offsets, same sorted ids, 5 channels
(rgb, depth, depth^2), CUDA events, 20 launches after 5 warm-ups.  Writes one JSON line.
TEST / MEASUREMENT INFRASTRUCTURE: nothing here is on the product path."""
import importlib
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import isogs_b200  # noqa: E402

pkg = isogs_b200.pkg
spec = importlib.util.spec_from_file_location("gsplat_ref", os.path.join(ROOT, "oracle", "_ref", "gsplat_ref.so"))
gsref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gsref)

syn, params, kfs = bench.build_workload(bench.SHAPE, bench.N_GAUSS, 2)
W, H, fx, fy, cx, cy = syn.SHAPES[bench.SHAPE][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
dev = "cuda:0"
flags = pkg._lib.F_LOG_SCALES | pkg._lib.F_LOGIT_OPAC
tt = lambda a: torch.from_numpy(a).to(dev)
k = kfs[1]
pose = pkg.make_pose(k["q"], k["t"])
means, ls = tt(params["means3D"]), tt(params["log_scales"])
lo, rgb = tt(params["logit_opacities"].reshape(-1)), tt(params["rgb_colors"])
ch = 5
render, alpha, ctx = pkg.raster.forward(cam, means, ls, None, lo, rgb, ch=ch, flags=flags, pose=pose)
rec = ctx.rec
m2d = rec[:, 0:2].contiguous()[None]
conics = torch.stack([2.0 * rec[:, 2], rec[:, 3], 2.0 * rec[:, 4]], -1).contiguous()[None]
opac = rec[:, 6].contiguous()[None]
colors = rec[:, 7:7 + ch].contiguous()[None]
th, tw = (H + 15) // 16, (W + 15) // 16
offsets = ctx.ranges[:-1].reshape(1, th, tw).contiguous()
ids = ctx.vals.contiguous()
v_render = torch.randn((ch, H, W), device=dev)
v_render[4] = 0
v_hwc = v_render.permute(1, 2, 0).contiguous()[None]
v_alpha = torch.zeros((1, H, W, 1), device=dev)


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


r_ref, a_ref, l_ref = gsref.rast_fwd(m2d, conics, colors, opac, W, H, offsets, ids)
ms = dict(
    gsplat_fwd=timeit(lambda: gsref.rast_fwd(m2d, conics, colors, opac, W, H, offsets, ids)),
    gsplat_bwd=timeit(lambda: gsref.rast_bwd(m2d, conics, colors, opac, W, H, offsets, ids, a_ref, l_ref, v_hwc, v_alpha)),
    ours_fwd=timeit(lambda: pkg.raster.blend_forward(cam, ch, ctx.rec, ctx.vals, ctx.ranges)),
    ours_bwd=timeit(lambda: pkg.raster.blend_backward(ctx, v_render)),
)
err = float((render - r_ref[0].permute(2, 0, 1)).abs().max())
print(json.dumps({"what": "blend kernels on the bench workload: reference gsplat 1.5.3 kernels (sm_100a, --use_fast_math) vs "
                          "B200-native, same inputs", "N": int(means.shape[0]), "M": int(ctx.M), "W": W, "H": H, "channels": ch,
                  "ms": {a: round(b, 4) for a, b in ms.items()},
                  "note": "gsplat timings include its output allocation + zero-fill inside the op (at::empty / at::zeros_like, "
                          "as in Rasterization.cpp:56-64,162-169); ours include torch.empty / zeros of the same outputs",
                  "speedup_fwd": round(ms["gsplat_fwd"] / ms["ours_fwd"], 2),
                  "speedup_bwd": round(ms["gsplat_bwd"] / ms["ours_bwd"], 2), "max_abs_image_diff": err}))
