"""Pure cost of the exchange step of keyframe-sharded mapping (MEASUREMENT INFRASTRUCTURE): two exchanges back to
back -- the first aligns the ranks (it ends on every rank when the last owner's slice has landed), so the second is
timed without rank skew.  torchrun --nproc-per-node N scripts/exchange_probe.py [n_gaussians ...]"""
import importlib
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isogs_b200  # noqa: E402,F401


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    sizes = [int(a) for a in sys.argv[1:]] or [1000000, 250000]
    out = {}
    for n in sizes:
        rng = np.random.RandomState(0)
        params = dict(means3D=rng.randn(n, 3).astype(np.float32), rgb_colors=rng.rand(n, 3).astype(np.float32),
                      unnorm_rotations=np.tile(np.array([[1, 0, 0, 0]], np.float32), (n, 1)),
                      logit_opacities=np.zeros((n, 1), np.float32), log_scales=np.full((n, 1), -4.0, np.float32))
        for mode in ("fused", "nvls", "p2p", "nccl"):
            ar, alloc = dp.allreduce_mean_, None
            try:
                if mode == "fused":
                    ar = dp.PeerReduceAdam(dev); alloc = ar.alloc
                elif mode == "nvls":
                    ar = dp.NvlsReduceAdam(dev); alloc = ar.alloc; ar.alloc(1024)
                elif mode == "p2p":
                    ar = dp.PeerAllReduce(dev); alloc = ar.alloc
            except RuntimeError as e:
                out[f"{mode}_{n}"] = str(e)
                continue
            gmap = ss.GaussianMap(params, device=dev, grad_alloc=alloc)
            if mode in ("fused", "nvls"):
                ar.adopt(gmap)
            gmap.grad_flat.normal_()
            gmap.adam_state.view(torch.int64)[0] = 1
            gmap.adam_state[1:3] = 1.0

            def exch():
                if mode in ("fused", "nvls"):
                    ar.fused_adam(gmap, ss.MAPPING_LRS)
                else:
                    ar(gmap.grad_flat[:gmap.grad_used])
                    ss.adam_gaussians(gmap, ss.MAPPING_LRS)
            for _ in range(5):
                exch()
            torch.cuda.synchronize()
            ts = []
            for _ in range(30):
                dist.barrier()
                e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                e[0].record(); exch(); e[1].record(); exch(); e[2].record()
                torch.cuda.synchronize()
                ts.append(e[1].elapsed_time(e[2]))
            t = torch.tensor([float(np.median(ts))], device=dev)
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            out[f"{mode}_{n}"] = dict(bytes=int(gmap.grad_used) * 4, ms_per_rank=[round(float(x), 4) for x in allt])
            if hasattr(ar, "check_status"):
                ar.check_status()
            if mode in ("fused", "nvls"):
                gmap.p = {}
                ar.drop_params()
            if hasattr(ar, "release"):
                gmap.grad_flat, gmap.g = None, {}
                ar.release()
            dist.barrier()
    if rank == 0:
        print("EXCHANGE_PROBE " + json.dumps(dict(world=world, **out)), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
