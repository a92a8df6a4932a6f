# Multi-GPU check on an N-GPU box: sharded parity test, then the bench with each exchange variant.
# Usage: gpurun --gpus N -- bash scripts/scale_check.sh <tag> N
TAG=$1; N=$2
mkdir -p gpurun_out
[ -n "$SKIP_PYTEST" ] || timeout 900 python -m pytest tests/test_gpu_multi.py -q -s > gpurun_out/${TAG}_multi_pytest.log 2>&1; grep -E "MULTI_GPU_PARITY|passed|failed|skipped" gpurun_out/${TAG}_multi_pytest.log | cut -c1-700
for mode in ${MODES:-nvls fused p2p nccl}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 20 --warmup 5 --allreduce $mode > gpurun_out/${TAG}_bench_${N}gpu_$mode.json 2> gpurun_out/${TAG}_bench_${N}gpu_$mode.err || tail -5 gpurun_out/${TAG}_bench_${N}gpu_$mode.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench_${N}gpu_$mode.json"))
    print("$mode", round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), d["stage_ms"], d.get("rank_stats"), d["config"]["replicas_identical_after_run"])
except Exception as e:
    print("$mode: no json", e)
PY
done
