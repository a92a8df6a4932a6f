"""Debug helper (needs `make -C isogs-slam-cpp_b200 clean all STATS=1`): prints the work counters of
the blend kernels on the bench workload -- batches, quad iterations, active / valid lanes."""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import isogs_b200  # noqa: E402

pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn, params, kfs = bench.build_workload(bench.SHAPE, bench.N_GAUSS, 2)
W, H, fx, fy, cx, cy = syn.SHAPES[bench.SHAPE][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
dev = torch.device("cuda:0")
gmap = ss.GaussianMap(params, device=dev)
ws = ss.StepWorkspace(cam, gmap.count, dev)
k = kfs[1]
pose = pkg.make_pose(k["q"], k["t"])
L = pkg.lib()
buf = (C.c_ulonglong * 16)()
L.isogs_debug_stats(buf, 1)
ss.map_step(ws, gmap, pose, torch.from_numpy(k["rgb"]).to(dev), torch.from_numpy(k["depth"]).to(dev), k["inv_nd"])
torch.cuda.synchronize()
L.isogs_debug_stats(buf, 0)
s = list(buf)
print("N", gmap.count, "M", ws.M)
print(f"fwd: batches {s[0]}  iterations {s[1]}  iters/batch {s[1]/max(1,s[0]):.2f}  active lanes/iter {s[2]/max(1,s[1]):.1f}"
      f"  valid lanes/iter {s[3]/max(1,s[1]):.1f}  valid total {s[3]}")
print(f"bwd: batches {s[8]}  iterations {s[9]}  iters/batch {s[9]/max(1,s[8]):.2f}  active lanes/iter {s[10]/max(1,s[9]):.1f}"
      f"  valid lanes/iter {s[11]/max(1,s[9]):.1f}  reducing iterations {s[12]}  valid total {s[11]}")
