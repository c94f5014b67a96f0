#!/bin/bash
# First GPU contact: smoke -> tests -> bench -> ncu launch list.  Everything under timeouts.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke (TMA)"; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/smoke.log
echo "== smoke (plain loads)"; RMB_FLAGS=2 timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_notma.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/smoke_notma.log
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_gpu.log
echo "== bench small"; timeout 300 python bench.py --n 200000 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_200k.json 2> gpurun_out/bench_200k.err; echo "rc=$?"; cat gpurun_out/bench_200k.json; tail -3 gpurun_out/bench_200k.err
echo "== bench 1M"; timeout 900 python bench.py > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err; echo "rc=$?"; cat gpurun_out/bench_1m.json; tail -3 gpurun_out/bench_1m.err
