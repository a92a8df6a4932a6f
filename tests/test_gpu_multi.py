"""Keyframe-sharded mapping on >= 2 real GPUs: the product's sharded map_step with each exchange variant
(NCCL all-reduce, peer-memory all-reduce, fused reduce + Adam over peer memory) against the CPU oracle executing the
same batched schedule, replicas bit-identical.  Skipped on a single-GPU box; run with `gpurun --gpus 2`."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_map_step_matches_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, OMP_NUM_THREADS="8")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29610 + world), os.path.join(ROOT, "tests", "multi_gpu_worker.py"), "3"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    line = [l for l in r.stdout.splitlines() if l.startswith("MULTI_GPU_PARITY")]
    print(line[-1] if line else r.stdout[-2000:] + r.stderr[-3000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert line and '"ok": true' in line[-1]
