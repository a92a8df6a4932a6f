"""Pin the two get_loss modes the shipped configs leave off against the reference's OWN Python code:

  * get_loss(mapping=True, ignore_outlier_depth_loss=True)  (splatam.py:545-559: the depth term is the MEAN over the
    pixels with |err| < 10 * median(err));
  * get_loss(mapping=True, do_ba=True)                       (splatam.py:505-510: transform_to_frame with
    gaussians_grad=True AND camera_grad=True -- the keyframe's pose gets a gradient next to the Gaussians).

Same recipe as oracle/pin_against_splatam.py (runs ONLY in the build container: it imports
/root/reference/python_ordinary/scripts/splatam.py on the CPU with the oracle standing in for the CUDA rasterizer),
on the SAME inputs: the Gaussians the mapping fixture starts from (map_in_* of tests/golden/splatam_step_64x48.npz),
frame 1 of that fixture (it contains a foreground object, i.e. depth outliers) and its tracked pose.  One evaluation
of the reference's complete mapping loss (0.8*L1 + 0.2*(1-SSIM) colour term, 50*flat + 2*iso) per mode, no optimizer
step.  Output: tests/golden/splatam_modes_64x48.npz.

Usage:  python oracle/pin_modes_splatam.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import pin_against_splatam as pin  # noqa: E402

GAUSS = pin.GAUSS


def main():
    fx = dict(np.load(os.path.join(ROOT, "tests", "golden", "splatam_step_64x48.npz")))
    W, H, FX, FY, CX, CY = [float(v) for v in fx["intr"]]
    out = {}
    torch.manual_seed(0)
    with pin.CpuShim():
        ref = pin.load_reference()
        K = torch.tensor([[FX, 0, CX], [0, FY, CY], [0, 0, 1]], dtype=torch.float32)
        w2c0 = torch.eye(4)
        cam = ref.setup_camera(int(W), int(H), K.numpy(), w2c0.numpy())
        rgb1, d1 = torch.from_numpy(fx["rgb1"]), torch.from_numpy(fx["d1"])
        N = fx["map_in_means3D"].shape[0]
        # a parameter / variables set of the right shapes from the reference's own initialiser, then the fixture's values
        pt = torch.cat([torch.from_numpy(fx["map_in_means3D"]), torch.from_numpy(fx["map_in_rgb_colors"])], 1)
        params, variables = ref.initialize_params(pt, 3, torch.ones(N), "isotropic")
        variables["scene_radius"] = torch.tensor(float(fx["scene_radius"][0]))
        pose = fx["track_poses"][2]
        with torch.no_grad():
            for k in GAUSS:
                params[k].copy_(torch.from_numpy(fx[f"map_in_{k}"]).reshape(params[k].shape))
            params["cam_unnorm_rots"][0, :, 1] = torch.from_numpy(pose[:4])
            params["cam_trans"][0, :, 1] = torch.from_numpy(pose[4:])
        curr = dict(cam=cam, im=rgb1, depth=d1, id=1, intrinsics=K, w2c=w2c0, iter_gt_w2c_list=[w2c0, w2c0])
        weights = dict(im=0.5, depth=1.0)
        for name, kw in (("outlier", dict(mapping=True)), ("ba", dict(mapping=True, do_ba=True))):
            for p in params.values():
                p.grad = None
            loss, variables, losses = ref.get_loss(params, curr, variables, 1, weights, True, 0.99, True,
                                                   name == "outlier", **kw)
            loss.backward()
            out[f"{name}_losses"] = np.array([float(loss), float(losses["im"]), float(losses["depth"]),
                                              float(losses["flat"]), float(losses["iso"])], np.float64)
            for k, g in pin.grads_of(params, GAUSS).items():
                out[f"{name}_grad_{k}"] = g
            if name == "ba":
                out["ba_grad_pose"] = np.concatenate([params["cam_unnorm_rots"].grad[0, :, 1].numpy(),
                                                      params["cam_trans"].grad[0, :, 1].numpy()])
            else:
                assert params["cam_unnorm_rots"].grad is None or float(params["cam_unnorm_rots"].grad.abs().max()) == 0.0
    out["pose"] = pose.astype(np.float32)
    out["notes"] = np.array("inputs: map_in_* / rgb1 / d1 / intr / track_poses[2] of splatam_step_64x48.npz; "
                            "outlier: get_loss(mapping=True, ignore_outlier_depth_loss=True); "
                            "ba: get_loss(mapping=True, do_ba=True); weights im=0.5 depth=1, flat=50, iso=2")
    path = os.path.join(ROOT, "tests", "golden", "splatam_modes_64x48.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if "losses" in k or k == "ba_grad_pose"})
    print(out["outlier_losses"], out["ba_losses"], out["ba_grad_pose"])


if __name__ == "__main__":
    main()
