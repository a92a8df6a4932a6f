# Round profile on the GPU box: GPU tests, the bench line + reference arm, the ncu launch list and ONE --set full
# capture of every kernel of one mapping iteration.  Usage: bash scripts/profile_round.sh <tag>
TAG=${1:-r2x}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2> gpurun_out/${TAG}_bench_reference_arm.err
for shape in tum scannetpp; do
  ISOGS_BENCH_SHAPE=$shape timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/${TAG}_bench_$shape.json 2> gpurun_out/${TAG}_bench_$shape.err
done
# launch list of 6 eager iterations (every launch with its device time; cold-cache, serialised: compare shares)
python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_ncu_launch.log 2>&1
# one full capture of each of the 11 kernels of the 5th iteration (4 x 11 launches skipped)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"^(step_begin|preprocess_|tile_|blend_|loss_l1|adam_)" -s 44 -c 11 -f -o gpurun_out/${TAG}_step python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/${TAG}_step.ncu-rep
timeout 600 python scripts/time_to_loss.py --ms "${STEP_MS:-1:0.813,2:0.864,8:0.918}" > gpurun_out/${TAG}_time_to_loss.log 2>&1; grep TIME_TO_LOSS gpurun_out/${TAG}_time_to_loss.log | cut -c1-1500 || tail -5 gpurun_out/${TAG}_time_to_loss.log
python - <<PY
import json
for n in ("bench", "bench_tum", "bench_scannetpp"):
    try:
        d = json.load(open("gpurun_out/${TAG}_%s.json" % n))
        print(n, round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms  e2e", round(d["e2e"]["value"], 1), "track", d.get("tracking_iters_per_s"), d["stage_ms"], d.get("parity"))
    except Exception as e:
        print(n, "no json", e)
PY
