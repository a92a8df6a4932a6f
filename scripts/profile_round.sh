set -x
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -5 > gpurun_out/r1h_pytest.log
timeout 600 python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches.csv python bench.py --steps 4 --warmup 3 > gpurun_out/r1h_ncu_launch.log 2>&1
for k in blend_bwd_kernel blend_fwd_kernel tile_sort_reg_kernel preprocess_bwd; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/r1h_$k python bench.py --steps 3 --warmup 3 > gpurun_out/r1h_ncu_$k.log 2>&1
done
ls -la gpurun_out | tail -20
cat gpurun_out/r1h_pytest.log; cat gpurun_out/r1h_bench.json
