# ncu --set full of one blend fwd + one blend bwd launch, for the tile-shared (ts) and the lane-per-quad (lq) kernels
TAG=$1
ISOGS_TS=1 python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_plain_ts.log 2>&1 && \
ISOGS_TS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 8 -c 2 -f -o gpurun_out/${TAG}_ts python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_ncu_ts.log 2>&1
ISOGS_TS=0 python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_plain_lq.log 2>&1 && \
ISOGS_TS=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 8 -c 2 -f -o gpurun_out/${TAG}_lq python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_ncu_lq.log 2>&1
ls -la gpurun_out/${TAG}_*.ncu-rep
