"""Why do 20-step blocks of some legs read slower than a continuous run?  MEASUREMENT script."""
import importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, isogs_b200
pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn, params, kfs = bench.build_workload("replica", 1000000, 8)
W, H, fx, fy, cx, cy = syn.SHAPES["replica"][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
dev = torch.device("cuda:0")
gmap = ss.GaussianMap(params, device=dev)
ws = ss.StepWorkspace(cam, gmap.count, dev)
for k in kfs:
    k["pose"] = pkg.make_pose(k["q"], k["t"]); k["rgb_d"] = torch.from_numpy(k["rgb"]).to(dev); k["depth_d"] = torch.from_numpy(k["depth"]).to(dev)
ps = ss.PoseState(kfs[0]["q"], kfs[0]["t"], dev)
def blocks(fn, n_blocks=15, steps=20):
    out = []
    for b in range(n_blocks):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps): fn(b * steps + i)
        e1.record(); torch.cuda.synchronize()
        out.append(round(e0.elapsed_time(e1) / steps, 4))
    return out
track = lambda i: ss.track_step(ws, gmap, ps, kfs[1]["rgb_d"], kfs[1]["depth_d"], graph=True)
track_rot = lambda i: ss.track_step(ws, gmap, ps, kfs[i % 8]["rgb_d"], kfs[i % 8]["depth_d"], graph=True)
mapk1 = lambda i: ss.map_step(ws, gmap, kfs[1]["pose"], kfs[1]["rgb_d"], kfs[1]["depth_d"], kfs[1]["inv_nd"], graph=True)
maprot = lambda i: ss.map_step(ws, gmap, kfs[i % 8]["pose"], kfs[i % 8]["rgb_d"], kfs[i % 8]["depth_d"], kfs[i % 8]["inv_nd"], graph=True)
res = {}
for name, fn in (("map_rotating_kf", maprot), ("map_kf1_only", mapk1), ("track_kf1_only", track), ("track_rotating_frames", track_rot)):
    for i in range(5): fn(i)
    res[name] = blocks(fn)
    ss.render_forward(ws, gmap, pose=kfs[1]["pose"]); torch.cuda.synchronize()
    res[name + "_M_kf1"] = ws.M
Ms = []
for k in kfs:
    ss.render_forward(ws, gmap, pose=k["pose"]); torch.cuda.synchronize(); Ms.append(ws.M)
res["M_per_keyframe"] = Ms
print(json.dumps(res))
