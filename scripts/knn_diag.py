import sys, importlib, numpy as np, torch
sys.path.insert(0, ".")
import bench, isogs_b200
pkg = isogs_b200.pkg
ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
syn, params, kfs = bench.build_workload(bench.SHAPE, bench.N_GAUSS, 2)
W, H, fx, fy, cx, cy = syn.SHAPES[bench.SHAPE][:6]
cam = pkg.make_camera(W, H, fx, fy, cx, cy)
p3 = dict(params); p3["log_scales"] = np.tile(np.asarray(params["log_scales"]).reshape(-1, 1), (1, 3)).astype(np.float32)
gmap = ss.GaussianMap(p3, device="cuda"); ws = ss.StepWorkspace(cam, gmap.count, torch.device("cuda"))
m = np.asarray(params["means3D"]); print("extent", m.min(0), m.max(0))
for res in (None, 160, 256, 400):
    ws._geo_key = None
    ss.iso_flat_loss(ws, gmap, seed=1, res=res, with_grad=False); torch.cuda.synchronize()
    g = ws.knn_grid[:8].cpu().numpy(); h = g[3:4].view(np.float32)[0]; nx, ny, nz = g[5:8]
    cells = int(nx) * int(ny) * int(nz)
    st = ws.knn_grid[16:16 + cells + 1].cpu().numpy(); occ = np.diff(st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10): ss.iso_flat_loss(ws, gmap, seed=i, res=res, with_grad=False)
    e1.record(); torch.cuda.synchronize()
    print("res", res, "h", h, "dims", nx, ny, nz, "occupied", (occ > 0).sum(), "max occ", occ.max(), "mean occ of occupied", occ[occ > 0].mean(),
          "p99", np.percentile(occ[occ > 0], 99), "ms/call", e0.elapsed_time(e1) / 10)
