# Round profile: GPU tests, bench line, ncu launch list, one --set full capture per kernel of interest.
# Usage (on the GPU box): bash scripts/profile_round.sh <tag> "<kernel regexes>"
TAG=${1:-r1x}
KERNELS=${2:-"blend_bwd_kernel blend_fwd_kernel tile_sort_reg_kernel preprocess_bwd"}
set -x
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -5 > gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu_launch.log 2>&1
for k in $KERNELS; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/${TAG}_$k python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu_$k.log 2>&1
done
cat gpurun_out/${TAG}_pytest.log; cut -c1-400 gpurun_out/${TAG}_bench.json
