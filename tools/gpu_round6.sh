#!/bin/bash
# validation after container re-creation: GPU tests, smoke, default bench, reference arm
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/r6_gpu.txt
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r6_pytest_gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r6_pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/r6_smoke.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r6_smoke.log
echo "== bench"; timeout 900 python bench.py > gpurun_out/r6_bench.json 2> gpurun_out/r6_bench.err; echo "rc=$?"; cat gpurun_out/r6_bench.json
echo "== ref"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r6_ref.json 2> gpurun_out/r6_ref.err; echo "rc=$?"; cat gpurun_out/r6_ref.json
