#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_gpu.log
echo "== c3"; timeout 900 python tools/bench_c3.py 2> gpurun_out/c3.err | tee gpurun_out/c3.json; tail -3 gpurun_out/c3.err
bash tools/gpu_profile.sh r01b
