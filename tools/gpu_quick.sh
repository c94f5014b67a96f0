#!/bin/bash
# quick check after a kernel change: GPU tests + short flagship bench + per-block rates (tools/tune.py if present)
mkdir -p gpurun_out
TAG=${1:-q}
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
echo "== bench"; timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu --no-solve > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "rc=$?"; cat gpurun_out/${TAG}_bench.json
