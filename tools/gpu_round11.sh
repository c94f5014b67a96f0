#!/bin/bash
# fused pair-symmetric kernel: tests + C3
mkdir -p gpurun_out
TAG=${1:-r01k}
echo "== pytest gpu (all)"; timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
echo "== c3"; timeout 600 python tools/bench_c3.py > gpurun_out/${TAG}_c3.json 2> gpurun_out/${TAG}_c3.err; echo "rc=$?"; cat gpurun_out/${TAG}_c3.json; tail -3 gpurun_out/${TAG}_c3.err
