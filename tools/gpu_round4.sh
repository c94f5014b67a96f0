#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_gpu.log
echo "== c2 solve"; timeout 600 python -c "
import bench, json
for nb in (100, 1000, 4000):
    print(json.dumps(bench.gmres_mobility_solve(nb)))
" 2>&1 | tail -5 | tee gpurun_out/c2_solve.jsonl
