#!/bin/bash
mkdir -p gpurun_out
echo "== pytest rigid"; timeout 900 python -m pytest tests/test_rigid_gpu.py -x -q > gpurun_out/r8_pytest_rigid.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/r8_pytest_rigid.log
echo "== c5 400 bodies 1 gpu"; timeout 900 python tools/bench_c5.py --bodies 400 2> gpurun_out/r8_c5_400.err | tee gpurun_out/r8_c5_400.json; tail -3 gpurun_out/r8_c5_400.err
echo "== c5 1600 bodies 1 gpu"; timeout 900 python tools/bench_c5.py --bodies 1600 2> gpurun_out/r8_c5_1600.err | tee gpurun_out/r8_c5_1600.json; tail -3 gpurun_out/r8_c5_1600.err
