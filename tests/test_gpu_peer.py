"""The NVLink peer all-reduce of keyframe-sharded mapping (csrc/peer.cu).

One-GPU part (what the driver's 1-GPU box runs): the kernel's protocol with the "ranks" being buffers and
streams of ONE device -- flags, slice ownership, rank-order sum, identical bits everywhere, repeated epochs.
Two-GPU part (skipped with fewer than 2 devices): real CUDA IPC between two processes, against NCCL.
"""
import ctypes as C
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import isogs_b200
    return isogs_b200.pkg


def _expected(bufs):
    """rank-order fp32 sum, then * (1/world) -- the arithmetic of reduce_slice."""
    acc = np.zeros_like(bufs[0])
    for b in bufs:
        acc = (acc + b).astype(np.float32)
    return (acc * np.float32(1.0 / len(bufs))).astype(np.float32)


@pytest.mark.parametrize("world,numel", [(1, 1024), (2, 4096 + 8), (3, 1000), (4, 40), (8, 8 * 4096 + 4), (5, 4)])
def test_protocol_on_one_device(pkg, world, numel):
    L = pkg.lib()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(world * 1000 + numel)
    host = [rng.standard_normal(numel).astype(np.float32) for _ in range(world)]
    bufs = [torch.from_numpy(h).to(dev) for h in host]
    flags = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(world)]
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    bp = (C.c_void_p * world)(*[b.data_ptr() for b in bufs])
    fp = (C.c_void_p * world)(*[f.data_ptr() for f in flags])
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):
        want = _expected(host)
        for r in range(world):            # all "ranks" must be co-resident: a handful of CTAs each
            pkg._lib.check(L.isogs_allreduce_avg_p2p(world, r, bp, fp, numel, epoch, 4, status.data_ptr(), streams[r].cuda_stream))
            pkg._lib.check(L.isogs_peer_wait(world, flags[r].data_ptr(), epoch, status.data_ptr(), streams[r].cuda_stream))
        torch.cuda.synchronize()
        assert int(status.item()) == 0
        for r in range(world):
            got = bufs[r].cpu().numpy()
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (world, numel, epoch, r)
        # next epoch: new "gradients" = reduced value + a rank-dependent perturbation
        host = [(want + np.float32(0.25 * (r + 1))).astype(np.float32) for r in range(world)]
        for r in range(world):
            bufs[r].copy_(torch.from_numpy(host[r]))
        torch.cuda.synchronize()


def test_argument_checks(pkg):
    L = pkg.lib()
    dev = torch.device("cuda", 0)
    b = torch.zeros(16, device=dev)
    f = torch.zeros(64, dtype=torch.int32, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    bp, fp = (C.c_void_p * 1)(b.data_ptr()), (C.c_void_p * 1)(f.data_ptr())
    assert L.isogs_allreduce_avg_p2p(1, 0, bp, fp, 6, 1, 0, st.data_ptr(), None) != 0          # numel % 4
    assert L.isogs_allreduce_avg_p2p(17, 0, bp, fp, 8, 1, 0, st.data_ptr(), None) != 0         # world > 16
    assert L.isogs_allreduce_avg_p2p(1, 1, bp, fp, 8, 1, 0, st.data_ptr(), None) != 0          # rank >= world
    bp_bad = (C.c_void_p * 1)(b.data_ptr() + 4)
    assert L.isogs_allreduce_avg_p2p(1, 0, bp_bad, fp, 8, 1, 0, st.data_ptr(), None) != 0      # misaligned
    assert L.isogs_allreduce_avg_p2p(1, 0, bp, fp, 0, 1, 0, st.data_ptr(), None) == 0          # empty is fine
    torch.cuda.synchronize()


def test_missing_peer_times_out_instead_of_hanging(pkg):
    """world = 2 but only rank 0 runs: both spins give up after 2 s and report through `status`."""
    L = pkg.lib()
    dev = torch.device("cuda", 0)
    bufs = [torch.ones(64, device=dev) for _ in range(2)]
    flags = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(2)]
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    bp = (C.c_void_p * 2)(*[b.data_ptr() for b in bufs])
    fp = (C.c_void_p * 2)(*[f.data_ptr() for f in flags])
    pkg._lib.check(L.isogs_allreduce_avg_p2p(2, 0, bp, fp, 64, 1, 1, status.data_ptr(), None))
    pkg._lib.check(L.isogs_peer_wait(2, flags[0].data_ptr(), 1, status.data_ptr(), None))
    torch.cuda.synchronize()
    assert int(status.item()) != 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _two_gpu_worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    import isogs_b200  # noqa: F401
    dp = importlib.import_module("isogs_slam_cpp_b200.dp")
    ar = dp.PeerAllReduce(dev)
    n = 1 << 20
    buf = ar.alloc(n)
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    ok = True
    for it in range(4):
        x = torch.randn(n, generator=g).to(dev)
        ref = x.clone()
        dist.all_reduce(ref, op=dist.ReduceOp.SUM)
        ref *= 1.0 / world
        buf.copy_(x)
        ar(buf[: n - 4 * it])
        ar.check_status()
        ok &= bool(torch.allclose(buf[: n - 4 * it], ref[: n - 4 * it], rtol=0, atol=1e-6))
        ok &= dp.replicas_identical([buf[: n - 4 * it]])
    ar.release()
    if rank == 0:
        open(out, "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_processes_over_ipc_match_nccl(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "result.txt")
    mp.spawn(_two_gpu_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == "ok"
