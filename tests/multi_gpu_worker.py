"""TEST INFRASTRUCTURE: one rank of the keyframe-sharded mapping parity run (launched by tests/test_gpu_multi.py with
torch.distributed.run, one process per GPU).  Every rank renders a DIFFERENT keyframe per iteration against the
replicated map; the exchange step is NCCL, the peer-memory all-reduce, or the fused reduce + Adam kernel.  Rank 0
also runs the CPU oracle on the same batched schedule (same keyframe draws, averaged gradient, one Adam
step) and checks every mode against it; all ranks check that the replicas are bit-identical."""
import importlib
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import isogs_b200  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pkg = isogs_b200.pkg
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    syn = importlib.import_module("isogs_slam_cpp_b200.synthetic")
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    params, _, _ = syn.make_map("tiny")
    W, H, fx, fy, cx, cy = syn.SHAPES["tiny"][:6]
    cam = pkg.make_camera(W, H, fx, fy, cx, cy)
    traj = syn.trajectory(16, step=2.0)
    kfs = []
    for k in range(2, 2 + max(world, 4)):
        c, d, _ = syn.render_frame("tiny", traj[k], seed=k)
        q, t = syn.w2c_to_pose(traj[k])
        kfs.append(dict(rgb=np.ascontiguousarray(c), depth=np.ascontiguousarray(d[0]), q=q, t=t,
                        inv_nd=1.0 / max(1, int((d > 0).sum()))))
    tt = lambda a: torch.from_numpy(a).to(dev)
    for k in kfs:
        k["rgb_d"], k["depth_d"], k["pose"] = tt(k["rgb"]), tt(k["depth"]), pkg.make_pose(k["q"], k["t"])
    names = ("means3D", "rgb_colors", "logit_opacities", "log_scales")
    results = {}
    nvls_note = None
    for mode in ("nccl", "p2p", "fused", "nvls"):
        ar, alloc = dp.allreduce_mean_, None
        if mode == "p2p":
            ar = dp.PeerAllReduce(dev); alloc = ar.alloc
        elif mode == "fused":
            ar = dp.PeerReduceAdam(dev); alloc = ar.alloc
        elif mode == "nvls":
            ar = dp.NvlsReduceAdam(dev); alloc = ar.alloc
            try:
                ar.alloc(1024)                     # COLLECTIVE probe: raises on every rank when there is no multicast
            except RuntimeError as e:
                nvls_note = str(e)
                continue
        gmap = ss.GaussianMap(params, device=dev, grad_alloc=alloc)
        if mode in ("fused", "nvls"):
            ar.adopt(gmap)
        ws = ss.StepWorkspace(cam, gmap.count, dev)
        losses = []
        for it in range(iters):
            k = kfs[dp.keyframe_for(it, rank, world, len(kfs))]
            out = ss.map_step(ws, gmap, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"], allreduce=ar)
            losses.append(float(out[0].item()))
        torch.cuda.synchronize()
        if hasattr(ar, "check_status"):
            ar.check_status()
        same = dp.replicas_identical([gmap.view(n) for n in names])
        results[mode] = dict(p={n: gmap.view(n).cpu().numpy().copy() for n in names}, losses=losses, same=bool(same))
        if mode in ("fused", "nvls"):
            # a re-sliced count (prune) needs complete moments everywhere: after sync_moments the moments of every
            # rank equal those of a full (unsharded) Adam
            ar.sync_moments(gmap)
            m_same = dp.replicas_identical([gmap.m[n][:gmap.count] for n in names] + [gmap.v[n][:gmap.count] for n in names])
            results[mode]["moments_identical_after_sync"] = bool(m_same)
            results[mode]["m"] = {n: gmap.m[n][:gmap.count].cpu().numpy().copy() for n in names}
            gmap.p, gmap.grad_flat, gmap.g = {}, None, {}
            ar.drop_params()
        if hasattr(ar, "release"):
            gmap.grad_flat, gmap.g = None, {}
            ar.release()
        dist.barrier()
    ok, report = True, {}
    if rank == 0:
        import oracle_slam as osl
        from oracle import oracle as orc
        ocam = orc.make_camera(W, H, fx, fy, cx, cy)
        st = osl.new_state(params)
        for it in range(iters):
            mine = kfs[dp.keyframe_for(it, 0, world, len(kfs))]
            extra = []
            for r in range(1, world):
                k = kfs[dp.keyframe_for(it, r, world, len(kfs))]
                extra.append(osl.gradients_only(orc, ocam, st["p"], k["q"], k["t"], k["rgb"], k["depth"]))
            osl.map_iteration(orc, ocam, st, mine["q"], mine["t"], mine["rgb"], mine["depth"], extra_grads=extra)
        lrs = ss.MAPPING_LRS
        for mode, res in results.items():
            worst = 0.0
            for n in names:
                ref = st["p"][n].reshape(res["p"][n].shape)
                # Adam normalises a step to ~lr per element: compare in units of a step
                worst = max(worst, float(np.abs(res["p"][n] - ref).max() / lrs[n]))
            report[mode] = dict(max_param_error_in_steps=worst, replicas_identical=res["same"])
            ok = ok and res["same"] and worst <= 0.05 * iters
        # all three exchange variants must agree with each other far more tightly than with the oracle
        for n in names:
            d = float(np.abs(results["fused"]["p"][n] - results["p2p"]["p"][n]).max() / lrs[n])
            report.setdefault("fused_vs_p2p_in_steps", {})[n] = d
            ok = ok and d <= 0.02 * iters
        report["nvls"] = report.get("nvls", dict(unavailable=nvls_note))
        if "nvls" in results:
            ok = ok and results["nvls"]["moments_identical_after_sync"]
            for n in names:
                d = float(np.abs(results["nvls"]["p"][n] - results["fused"]["p"][n]).max() / lrs[n])
                report.setdefault("nvls_vs_fused_in_steps", {})[n] = d
                ok = ok and d <= 0.02 * iters
        report["moments_identical_after_sync"] = results["fused"]["moments_identical_after_sync"]
        ok = ok and results["fused"]["moments_identical_after_sync"]
        m_ref = st["m"]["rgb_colors"].reshape(results["fused"]["m"]["rgb_colors"].shape)
        report["moment_error_rel"] = float(np.abs(results["fused"]["m"]["rgb_colors"] - m_ref).max() / (np.abs(m_ref).max() + 1e-30))
        ok = ok and report["moment_error_rel"] <= 2e-3
        print("MULTI_GPU_PARITY " + json.dumps(dict(world=world, iters=iters, ok=bool(ok), **report)), flush=True)
    flag = torch.tensor([1 if (ok and all(r["same"] for r in results.values())) else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
