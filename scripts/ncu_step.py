"""A handful of eager mapping iterations at a bench shape, for ncu (MEASUREMENT script).
  python scripts/ncu_step.py [shape] [steps]"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, isogs_b200
shape = sys.argv[1] if len(sys.argv) > 1 else "replica"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
n = {"tum": 307200, "scannetpp": 3000000}.get(shape, 1000000)
pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn, params, kfs = bench.build_workload(shape, n, 8)
W, H, fx, fy, cx, cy = syn.SHAPES[shape][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
dev = torch.device("cuda:0")
gmap = ss.GaussianMap(params, device=dev)
ws = ss.StepWorkspace(cam, gmap.count, dev)
for k in kfs:
    k["pose"] = pkg.make_pose(k["q"], k["t"]); k["rgb_d"] = torch.from_numpy(k["rgb"]).to(dev); k["depth_d"] = torch.from_numpy(k["depth"]).to(dev)
for i in range(steps):
    k = kfs[1]
    ss.map_step(ws, gmap, k["pose"], k["rgb_d"], k["depth_d"], k["inv_nd"])
torch.cuda.synchronize()
print("ok", ws.M)
