#!/bin/bash
# guard test + ncu metrics of the C3 kernels (pair-symmetric defaults)
mkdir -p gpurun_out
TAG=${1:-r01k}
echo "== pytest gpu (all)"; timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
echo "== ncu c3"; timeout 900 ncu --metrics sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread --clock-control none -k regex:sym_kernel --csv --log-file gpurun_out/${TAG}_c3_kernels_ncu.csv python tools/bench_c3.py > gpurun_out/${TAG}_c3_ncu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/${TAG}_c3_ncu.log
