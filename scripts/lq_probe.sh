# A/B of the lane-per-quad blend backward (ISOGS_BWD_LQ) against the quad-of-lanes kernel: parity tests, then bench.
TAG=${1:-lq}
ISOGS_BWD_LQ=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_step.py tests/test_gpu_vs_gsplat.py tests/test_gpu_fullsize.py -q -m gpu 2>&1 | tail -15 > gpurun_out/${TAG}_pytest_lq1.log
cat gpurun_out/${TAG}_pytest_lq1.log
for cfg in "0 4" "1 4" "1 8" "2 4"; do
  set -- $cfg
  ISOGS_BWD_LQ=$1 ISOGS_NW_BWD=$2 timeout 300 python bench.py --no-cpu --steps 50 --warmup 5 > gpurun_out/${TAG}_bench_lq$1_nw$2.json 2> gpurun_out/${TAG}_bench_lq$1_nw$2.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_lq$1_nw$2.json"))
print("LQ=$1 NW=$2", round(d["value"],1), "iters/s  track", round(d["tracking_iters_per_s"],1), d["stage_ms"])
PY
done
