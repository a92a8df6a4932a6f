# A/B on the GPU box: bench lines (no CPU leg) for a list of environment settings.  Usage: bash scripts/gpu_ab.sh <tag> "ENV1=.. ENV2=.." "ENV.." ...
TAG=$1; shift
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  echo "== [$i] $envs"
  env $envs timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu $BENCH_ARGS > gpurun_out/${TAG}_ab$i.json 2> gpurun_out/${TAG}_ab$i.err || tail -5 gpurun_out/${TAG}_ab$i.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_ab$i.json"))
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in d.items() if k in ("value","ms_per_step","tracking_iters_per_s","mapping_with_ssim_iters_per_s","mapping_full_loss_3axis_iters_per_s","raster_fwd_bwd_ms")}, "e2e", round(d["e2e"]["value"],1), d["stage_ms"], d["config"]["n_isects"])
except Exception as e:
    print("no json", e)
PY
done
