# One GPU-box round: GPU tests, the default bench line, the reference arm.  Usage: bash scripts/gpu_round.sh <tag> [pytest args]
TAG=${1:-r2x}
shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.csv 2>&1
nproc > gpurun_out/${TAG}_nproc.txt
timeout 1500 python -m pytest tests -q -m gpu "$@" > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -25 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -5 gpurun_out/${TAG}_bench.err; cut -c1-3000 gpurun_out/${TAG}_bench.json
