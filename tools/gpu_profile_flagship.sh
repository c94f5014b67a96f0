#!/bin/bash
# r01g (1 GPU): tests, smoke, default bench + reference arm, launch list + full ncu capture of the flagship kernel,
# DRAM traffic of one flagship launch at N = 1M
mkdir -p gpurun_out
T=${1:-r01g}
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/${T}_pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_smoke.log
echo "== bench"; timeout 900 python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench.err; echo "rc=$?"; cat gpurun_out/${T}_bench_1gpu.json
echo "== ref"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "rc=$?"
CMD="python bench.py --n 200000 --steps 2 --warmup 1 --no-cpu --no-parity"
echo "== launch list"; ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu_launch.log 2>&1; echo "rc=$?"
echo "== full"; ncu --set full --clock-control none --import-source on -k regex:sym_kernel -s 1 -c 1 -o gpurun_out/${T}_sym_full -f $CMD > gpurun_out/${T}_ncu_full.log 2>&1; echo "rc=$?"
echo "== traffic"; timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:sym_kernel -c 2 --csv --log-file gpurun_out/${T}_traffic_n1m.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-parity --no-solve > gpurun_out/${T}_traffic.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_traffic_n1m.csv
