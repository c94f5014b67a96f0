#!/bin/bash
mkdir -p gpurun_out
echo "== pytest st"; timeout 900 python -m pytest tests/test_source_target_gpu.py -x -q > gpurun_out/r7_pytest_st.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/r7_pytest_st.log
echo "== pytest all gpu"; timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r7_pytest_gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r7_pytest_gpu.log
echo "== st rates"; timeout 600 python tools/bench_st.py 200000 | tee gpurun_out/r7_st_rates.json
