"""Host-side logic of the keyframe-sharded data-parallel mapping, on CPU with the
gloo backend and world_size 2.  Gradients come from the CPU oracle (test infrastructure); what is under
test is the product's plumbing: keyframe schedule, the single flat gradient buffer of GaussianMap, the
one-collective average, and that replicas stay bit-identical after identical Adam steps."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WINDOW, ITERS = 4, 3


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _workload():
    import importlib
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import isogs_b200  # noqa: F401
    syn = importlib.import_module("isogs_slam_cpp_b200.synthetic")
    shape = (64, 48, 52.0, 52.0, 31.5, 23.5, 0.0, 0.0)
    traj = syn.trajectory(WINDOW + 1, step=3.0)
    c0, d0, K = syn.render_frame(shape, traj[0])
    pts, cols, msd = syn.backproject(c0, d0, K, traj[0])
    params = syn.init_params(pts, cols, msd, isotropic=True)
    kfs = []
    for k in range(1, WINDOW + 1):
        c, d, _ = syn.render_frame(shape, traj[k])
        q, t = syn.w2c_to_pose(traj[k])
        kfs.append(dict(rgb=np.ascontiguousarray(c), depth=np.ascontiguousarray(d[0]), q=q, t=t))
    return syn, shape, params, kfs


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import importlib
        syn, shape, params, kfs = _workload()
        from oracle import oracle as orc
        import oracle_slam as osl
        ss = importlib.import_module("isogs_slam_cpp_b200.slam_step")
        dp = importlib.import_module("isogs_slam_cpp_b200.dp")
        ocam = orc.make_camera(*shape[:6])
        gmap = ss.GaussianMap(params, device="cpu")           # the product's replicated map + flat gradient buffer
        assert gmap.grad_names == ["means3D", "rgb_colors", "logit_opacities", "log_scales"]
        assert gmap.grad_used == gmap.count * (3 + 3 + 1 + 1)
        state = osl.new_state(params)                         # oracle-side Adam state of THIS replica
        seen = []
        for it in range(ITERS):
            slot = dp.keyframe_for(it, rank, world, WINDOW)
            seen.append(slot)
            kf = kfs[slot]
            cur = {k: state["p"][k] for k in params}
            g = osl.gradients_only(orc, ocam, cur, kf["q"], kf["t"], kf["rgb"], kf["depth"])
            for name in gmap.grad_names:                      # projection backward writes into the flat buffer
                gmap.g[name].copy_(torch.from_numpy(np.asarray(g[name], np.float32).reshape(gmap.g[name].shape)))
            flat = gmap.grad_flat[:gmap.grad_used]
            before = flat.clone()
            dp.allreduce_mean_(flat)                          # ONE collective for every parameter group
            assert flat.data_ptr() == gmap.g["means3D"].data_ptr(), "views must alias the flat buffer"
            if world > 1:
                assert not torch.equal(before, flat)
            state["step"] += 1
            for name in gmap.grad_names:
                lr = osl.MAP_LRS[name]
                orc.adam(state["p"][name].reshape(-1), gmap.g[name].numpy().reshape(-1), state["m"][name].reshape(-1),
                         state["v"][name].reshape(-1), lr, eps=1e-15, step=state["step"])
            tens = [torch.from_numpy(state["p"][n]) for n in gmap.grad_names]
            assert dp.replicas_identical(tens), f"replicas diverged at iteration {it}"
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), seen=np.array(seen),
                 **{n: state["p"][n] for n in gmap.grad_names})
    finally:
        dist.destroy_process_group()


def test_keyframe_schedule():
    import importlib
    sys.path.insert(0, ROOT)
    import isogs_b200  # noqa: F401
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    for world in (1, 2, 4, 8):
        for window in (8, 24):
            for it in range(6):
                slots = [dp.keyframe_for(it, r, world, window) for r in range(world)]
                assert len(set(slots)) == min(world, window), "ranks must render different keyframes"
            per_rank = {dp.keyframe_for(it, 0, world, window) for it in range(window)}
            assert len(per_rank) == window // np.gcd(world, window)
    assert dp.allreduce_mean_(torch.ones(4)) is not None      # world 1: no process group needed


@pytest.mark.timeout(300)
def test_peer_allreduce_slices_partition_the_buffer():
    """Slice ownership of the two-shot peer all-reduce (csrc/peer.cu): float4 granularity, disjoint, complete."""
    import importlib

    import isogs_b200  # noqa: F401
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    for world in (1, 2, 3, 4, 5, 8, 16):
        for numel in (0, 4, 8, 40, 1000, 4096 + 8, 8 * 1000 * 1000):
            prev = 0
            for r in range(world):
                b, e = dp.slice_bounds(numel, world, r)
                assert b == prev and e >= b and b % 4 == 0 and e % 4 == 0
                prev = e
            assert prev == numel


def test_dp_mapping_world2_gloo(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert list(r0["seen"]) == [0, 2, 0] and list(r1["seen"]) == [1, 3, 1]
    names = ["means3D", "rgb_colors", "logit_opacities", "log_scales"]
    for n in names:
        assert np.array_equal(r0[n], r1[n]), "replicas must be bit-identical"
    # reference: the same batched schedule executed in ONE process (averaged gradient)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    syn, shape, params, kfs = _workload()
    from oracle import oracle as orc
    import oracle_slam as osl
    ocam = orc.make_camera(*shape[:6])
    st = osl.new_state(params)
    for it in range(ITERS):
        a, b = kfs[(it * 2) % WINDOW], kfs[(it * 2 + 1) % WINDOW]
        gb = osl.gradients_only(orc, ocam, {k: st["p"][k] for k in params}, b["q"], b["t"], b["rgb"], b["depth"])
        osl.map_iteration(orc, ocam, st, a["q"], a["t"], a["rgb"], a["depth"], extra_grads=[gb])
    for n in names:
        step = np.abs(st["p"][n] - np.asarray(params[n], np.float32)).max()
        assert np.abs(st["p"][n] - r0[n]).max() <= 1e-3 * step + 1e-9, n
