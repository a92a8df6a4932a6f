# Multi-GPU check on an N-GPU box: mapping bench with the peer-memory gradient average and with NCCL, the
# two-process IPC test.  Usage (on the GPU box): bash scripts/scale_check.sh <tag> <n_gpus>
TAG=${1:-r1x}; N=${2:-8}
run() { # $1 = ranks, $2 = out suffix, rest = extra bench flags
  local n=$1 suf=$2; shift 2
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 50 --warmup 5 "$@" > gpurun_out/${TAG}_bench_${suf}.json 2> gpurun_out/${TAG}_bench_${suf}.err
  echo "rc=$? ${suf}: $(cut -c1-330 gpurun_out/${TAG}_bench_${suf}.json)"
}
run $N ${N}gpu
run $N ${N}gpu_nccl --allreduce nccl
if [ -n "$SCALE_PEER_TEST" ]; then timeout 200 python -m pytest tests/test_gpu_peer.py -q -m gpu 2>&1 | tail -3 > gpurun_out/${TAG}_pytest_peer.log; cat gpurun_out/${TAG}_pytest_peer.log; fi
if [ $N -gt 4 ]; then run 4 4gpu; fi
if [ $N -gt 2 ]; then run 2 2gpu; fi
