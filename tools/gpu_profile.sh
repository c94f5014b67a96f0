#!/bin/bash
# ncu launch list + one full capture of the flagship kernel (1 GPU).  Usage: tools/gpu_profile.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --n 200000 --steps 2 --warmup 1 --no-cpu --no-parity"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:mf_kernel -s 1 -c 1 -o gpurun_out/${TAG}_mf_full -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full rc=$?"
ls -la gpurun_out | tail -12
