# One full ncu capture of the two blend kernels of the 5th eager mapping iteration (the plain run first: ncu only
# after the same command exited 0 without it).  Usage: bash scripts/ncu_blend.sh <tag>; then
#   python scripts/ncu_summary_all.py gpurun_out/<tag>_blend.ncu-rep profiles/<tag>
#   python scripts/ncu_regions.py gpurun_out/<tag>_blend.ncu-rep
TAG=${1:-r2x}
mkdir -p gpurun_out
python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 8 -c 2 -f -o gpurun_out/${TAG}_blend python scripts/ncu_step.py replica 6 > gpurun_out/${TAG}_ncu_blend.log 2>&1
ls -la gpurun_out/${TAG}_blend.ncu-rep
