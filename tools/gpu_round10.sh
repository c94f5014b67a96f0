#!/bin/bash
# periodic pair-symmetric kernel: tests, rates, C3
mkdir -p gpurun_out
TAG=${1:-r01i}
echo "== new tests"; timeout 900 python -m pytest tests/test_mobility_gpu.py -m gpu -x -q -k "symmetric or sharded or selection" > gpurun_out/${TAG}_pytest_sym.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${TAG}_pytest_sym.log
echo "== rates"; timeout 600 python tools/time_sym.py 100000 400000 > gpurun_out/${TAG}_sym_rates.json 2> gpurun_out/${TAG}_sym_rates.err; echo "rc=$?"; cat gpurun_out/${TAG}_sym_rates.json; tail -3 gpurun_out/${TAG}_sym_rates.err
echo "== c3"; timeout 600 python tools/bench_c3.py > gpurun_out/${TAG}_c3.json 2> gpurun_out/${TAG}_c3.err; echo "rc=$?"; cat gpurun_out/${TAG}_c3.json; tail -3 gpurun_out/${TAG}_c3.err
echo "== pytest gpu (all)"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
