# A/B of the lane-per-quad blend kernels (ISOGS_FWD_LQ / ISOGS_BWD_LQ) against the quad-of-lanes kernels:
# parity tests, bench per configuration "FWD_LQ BWD_LQ NW_FWD NW_BWD", optional ncu captures.
TAG=${1:-lq}; shift
ISOGS_LQ_WARPS=${LQ_TEST_WARPS:-2} ISOGS_FWD_LQ=1 ISOGS_BWD_LQ=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_step.py tests/test_gpu_vs_gsplat.py tests/test_gpu_fullsize.py -q -m gpu 2>&1 | tail -15 > gpurun_out/${TAG}_pytest_lq.log
cat gpurun_out/${TAG}_pytest_lq.log
for cfg in "$@"; do
  set -- $cfg
  n=lq$1$2_nw$3$4_w${5:-2}
  ISOGS_LQ_WARPS=${5:-2} ISOGS_FWD_LQ=$1 ISOGS_BWD_LQ=$2 ISOGS_NW_FWD=$3 ISOGS_NW_BWD=$4 timeout 300 python bench.py --no-cpu --steps 50 --warmup 5 > gpurun_out/${TAG}_bench_$n.json 2> gpurun_out/${TAG}_bench_$n.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_$n.json"))
print("FWD_LQ=$1 BWD_LQ=$2 NW=$3/$4 W=${5:-2}", round(d["value"],1), "iters/s  track", round(d["tracking_iters_per_s"],1), "raster", round(d["raster_fwd_bwd_ms"],4), d["stage_ms"])
PY
done
if [ -n "$LQ_NCU" ]; then
  for k in blend_bwd_lq_kernel blend_fwd_lq_kernel; do
    ISOGS_FWD_LQ=1 ISOGS_BWD_LQ=1 ISOGS_NW_FWD=${LQ_NWF:-4} ISOGS_NW_BWD=${LQ_NWB:-8} timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/${TAG}_$k python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu_$k.log 2>&1
  done
fi
