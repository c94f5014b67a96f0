#!/bin/bash
# round-end validation on one GPU: tests, smoke, default bench, reference arm
mkdir -p gpurun_out
T=${1:-r01n}
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/${T}_pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_smoke.log
echo "== bench"; timeout 900 python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench.err; echo "rc=$?"; cut -c1-900 gpurun_out/${T}_bench_1gpu.json
echo "== ref"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "rc=$?"; cut -c1-200 gpurun_out/${T}_ref.json
echo "== c3"; timeout 300 python tools/bench_c3.py > gpurun_out/${T}_c3.json 2> gpurun_out/${T}_c3.err; echo "rc=$?"; cat gpurun_out/${T}_c3.json
