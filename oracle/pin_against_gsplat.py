"""Pin the CPU oracle against the reference's own CPU-runnable transcription.

Runs ONLY in the build container (it imports /root/reference/.../gsplat/cuda/_torch_impl.py,
which does not exist on the GPU box).  It
  1. evaluates gsplat's pure-torch `_quat_scale_to_covar_preci`, `_fully_fused_projection`,
     `_isect_tiles`, `_isect_offset_encode` (reference file: cpp/third_party/gsplat/gsplat/cuda/
     _torch_impl.py:45,286,378,477) on seeded inputs,
  2. checks the oracle against them (projection within fp32 tolerance, keys/ranges exact), and
  3. writes inputs + reference outputs to tests/golden/*.npz so tests/test_oracle_golden.py can
     repeat the comparison anywhere.

Usage:  python oracle/pin_against_gsplat.py
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

REF = "/root/reference/cpp/third_party/gsplat"


def load_ref():
    spec = importlib.util.spec_from_file_location("gs_torch_impl", f"{REF}/gsplat/cuda/_torch_impl.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def garden_case(n=3000, seed=42):
    """gsplat's own test recipe (tests/test_basic.py:21-60, gsplat/_helper.py:9-55) on a subsample."""
    d = np.load(f"{REF}/assets/test_garden.npz")
    means = torch.from_numpy(d["means3d"]).float()
    sel = ((means.abs() < 2).all(-1)).nonzero().squeeze(-1)
    g = torch.Generator().manual_seed(seed)
    sel = sel[torch.randperm(sel.numel(), generator=g)[:n]]
    means = means[sel].contiguous()
    colors = torch.from_numpy(d["colors"][sel.numpy()] / 255.0).float()
    N = means.shape[0]
    scales = torch.rand((N, 3), generator=g) * 0.02
    quats = torch.randn((N, 4), generator=g)
    quats = quats / quats.norm(dim=-1, keepdim=True)
    opac = torch.rand((N,), generator=g)
    viewmat = torch.from_numpy(d["viewmats"][0]).float()
    K = torch.from_numpy(d["Ks"][0]).float()
    W, H = int(d["width"]), int(d["height"])
    return dict(name="gsplat_garden_3k", means=means, scales=scales, quats=quats, opac=opac, colors=colors,
                viewmat=viewmat, K=K, W=W, H=H)


def splatam_iso_case(n=5000, seed=0):
    """SplaTAM-shaped: points back-projected from pixels, s = z / f (splatam.py:298-348), view = I."""
    g = torch.Generator().manual_seed(seed)
    W, H, f = 320, 240, 260.0
    u = torch.rand(n, generator=g) * (W + 40) - 20
    v = torch.rand(n, generator=g) * (H + 40) - 20
    z = 0.5 + torch.rand(n, generator=g) * 5.0
    means = torch.stack([(u - 159.5) / f * z, (v - 119.5) / f * z, z], -1)
    s = (z / f) * (0.5 + 3.0 * torch.rand(n, generator=g))
    scales = s[:, None].repeat(1, 3)
    quats = torch.zeros(n, 4)
    quats[:, 0] = 1
    opac = torch.rand(n, generator=g)
    colors = torch.rand(n, 3, generator=g)
    viewmat = torch.eye(4)
    K = torch.tensor([[f, 0, 159.5], [0, f, 119.5], [0, 0, 1]])
    return dict(name="splatam_iso_5k", means=means, scales=scales, quats=quats, opac=opac, colors=colors,
                viewmat=viewmat, K=K, W=W, H=H)


def run_reference(m, c):
    means = c["means"].clone().requires_grad_(True)
    quats = c["quats"].clone().requires_grad_(True)
    scales = c["scales"].clone().requires_grad_(True)
    covars, _ = m._quat_scale_to_covar_preci(quats, scales, True, False, triu=False)
    radii, means2d, depths, conics, _ = m._fully_fused_projection(
        means, covars, c["viewmat"][None], c["K"][None], c["W"], c["H"], eps2d=0.3, near_plane=0.01,
        far_plane=100.0)
    # projection backward (the pattern of gsplat tests/test_basic.py:230-268): random cotangents
    g = torch.Generator().manual_seed(7)
    v_means2d = torch.randn(means2d.shape, generator=g)
    v_depths = torch.randn(depths.shape, generator=g)
    v_conics = torch.randn(conics.shape, generator=g)
    sel = (radii > 0).all(-1)
    loss = ((means2d * v_means2d).sum(-1) * sel).sum() + (depths * v_depths * sel).sum() + \
        ((conics * v_conics).sum(-1) * sel).sum()
    gm, gq, gs = torch.autograd.grad(loss, [means, quats, scales])
    means2d, depths, conics = means2d.detach(), depths.detach(), conics.detach()
    bwd = dict(v_means2d=v_means2d[0].numpy(), v_depths=v_depths[0].numpy(), v_conics=v_conics[0].numpy(),
               g_means=gm.numpy(), g_quats=gq.numpy(), g_scales=gs.numpy())
    tw, th = (c["W"] + 15) // 16, (c["H"] + 15) // 16
    tiles, isect_ids, flatten_ids = m._isect_tiles(means2d, radii, depths, 16, tw, th, sort=True)
    offsets = m._isect_offset_encode(isect_ids, 1, tw, th)
    return dict(radii=radii[0].numpy(), means2d=means2d[0].numpy(), depths=depths[0].numpy(),
                conics=conics[0].numpy(), tiles=tiles[0].numpy(), isect_ids=isect_ids.numpy(),
                flatten_ids=flatten_ids.numpy(), offsets=offsets.reshape(-1).numpy(), **bwd)


def check_case(inp, ref, verbose=True):
    """Shared by this script and tests/test_oracle_golden.py.  inp/ref: dicts of numpy arrays."""
    W, H = int(inp["W"]), int(inp["H"])
    K = inp["K"]
    cam = orc.make_camera(W, H, K[0, 0], K[1, 1], K[0, 2], K[1, 2], inp["viewmat"])
    iso = bool(inp.get("iso", False))
    scales = inp["scales"][:, :1] if iso else inp["scales"]
    rec, radii, tiles = orc.preprocess_fwd(cam, inp["means"], scales, None if iso else inp["quats"], inp["opac"],
                                           inp["colors"], ch=4, flags=orc.F_NO_OPAC_EXTENT)
    # gsplat's torch impl does not cull on opacity; compare where both are visible
    vis_o = radii[:, 0] > 0
    vis_r = (ref["radii"] > 0).all(-1)
    low_opac = inp["opac"] < 1.0 / 255.0
    assert (vis_o | low_opac == vis_r | low_opac).mean() > 0.999, "visibility sets differ"
    both = vis_o & vis_r
    out = {}
    out["means2d"] = np.abs(rec[both, :2] - ref["means2d"][both]).max()
    out["depths"] = np.abs(rec[both, 10] - ref["depths"][both]).max()
    con = np.stack([2 * rec[:, 2], rec[:, 3], 2 * rec[:, 4]], -1)
    rel = np.abs(con[both] - ref["conics"][both]) / (np.abs(ref["conics"][both]).max(-1, keepdims=True) + 1e-12)
    out["conics_rel"] = rel.max()
    out["radii_mismatch"] = int((radii[both] != ref["radii"][both]).any(-1).sum())
    assert out["means2d"] < 2e-3 and out["depths"] < 1e-5 and out["conics_rel"] < 2e-3, out
    assert out["radii_mismatch"] <= max(1, both.sum() // 1000), out
    # keys / ranges: feed the REFERENCE's projected values through the oracle's binning -> exact
    rec2 = np.zeros_like(rec)
    rec2[:, :2] = ref["means2d"]
    rec2[:, 10] = ref["depths"]
    rad2 = ref["radii"].astype(np.int32).copy()
    rad2[~vis_r] = 0
    tw, th = (W + 15) // 16, (H + 15) // 16
    tiles2 = np.zeros(rec.shape[0], np.int32)
    # oracle recomputes rects; tile counts come from the reference
    tiles2[:] = ref["tiles"]
    ks, vs, ranges = orc.tile_bin(cam, rec2, rad2, tiles2)
    assert ks.shape[0] == ref["isect_ids"].shape[0]
    assert (ks.astype(np.int64) == ref["isect_ids"]).all(), "sorted keys differ"
    # torch.sort is not stable: compare values as (key, value) lexicographic order
    o1 = np.lexsort((vs, ks))
    o2 = np.lexsort((ref["flatten_ids"], ref["isect_ids"]))
    assert (vs[o1] == ref["flatten_ids"][o2]).all(), "flatten ids differ"
    assert (ranges[:-1] == ref["offsets"]).all() and ranges[-1] == ks.shape[0], "tile ranges differ"
    # projection backward vs torch autograd of the reference transcription
    v_rec = np.zeros_like(rec)
    v_rec[:, 0:2] = ref["v_means2d"]
    v_rec[:, 2:5] = ref["v_conics"]
    v_rec[:, 9] = ref["v_depths"]
    gr = orc.preprocess_bwd(cam, inp["means"], scales, None if iso else inp["quats"], inp["opac"], inp["colors"],
                            radii, rec, v_rec, ch=4, flags=orc.F_NO_OPAC_EXTENT)

    def relerr(a, b):
        return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-20))
    out["g_means_rel"] = relerr(gr["means"][both], ref["g_means"][both])
    if iso:
        out["g_scales_rel"] = relerr(gr["scales"][both, 0], ref["g_scales"][both].sum(-1))
    else:
        out["g_scales_rel"] = relerr(gr["scales"][both], ref["g_scales"][both])
        out["g_quats_rel"] = relerr(gr["quats"][both], ref["g_quats"][both])
    assert all(v < 1e-3 for k, v in out.items() if k.startswith("g_")), out
    out["n_isects"] = int(ks.shape[0])
    if verbose:
        print(inp.get("name", ""), out)
    return out


def main():
    m = load_ref()
    os.makedirs(f"{ROOT}/tests/golden", exist_ok=True)
    for case, iso in ((garden_case(), False), (splatam_iso_case(), True)):
        ref = run_reference(m, case)
        inp = {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in case.items() if k != "name"}
        inp["iso"] = iso
        inp["name"] = case["name"]
        check_case(inp, ref)
        np.savez_compressed(f"{ROOT}/tests/golden/{case['name']}.npz",
                            **{f"in_{k}": v for k, v in inp.items() if k != "name"},
                            **{f"ref_{k}": v for k, v in ref.items()})
        print("wrote", case["name"])


if __name__ == "__main__":
    main()
