#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu.log
echo "== rates"; timeout 600 python tools/tune.py 400000:UF:1:0:open 400000:UF:0:0:open 400000:UT:1:0:open 400000:WF:1:0:open 400000:WT:1:0:open 400000:UF_UT:1:0:open 200000:UF:1:0:xy 400000:UF:1:1:open 400000:WT:1:1:open 15000:UF:1:0:open 1000:UF:1:0:open 2>&1 | tail -1 | tee gpurun_out/rates.json
echo "== bench 1M"; timeout 900 python bench.py > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err; echo "rc=$?"; cat gpurun_out/bench_1m.json; tail -3 gpurun_out/bench_1m.err
