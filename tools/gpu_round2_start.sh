#!/bin/bash
# First GPU call of a new round (one B200, ~12 min): the tests added last in round 1 (each was run on a B200 on its
# own; the Blake-limit invariant only on the oracle), the full GPU suite, smoke, bench + reference arm, and the
# symmetric/row crossover of the free-space rotation blocks at small N.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round2_start.sh r02a'
mkdir -p gpurun_out
T=${1:-r02a}
echo "== new rigid tests"; timeout 120 python -m pytest tests/test_zzz_invariants_more.py tests/test_zzz_schemes_more_gpu.py -m gpu -q 2>&1 | tail -3
echo "== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/${T}_pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_smoke.log
echo "== bench"; timeout 900 python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench.err; echo "rc=$?"; cut -c1-900 gpurun_out/${T}_bench_1gpu.json
echo "== ref"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "rc=$?"; cut -c1-200 gpurun_out/${T}_ref.json
echo "== free-space symmetric vs row at small N (threshold RMB_SYMMETRIC_MIN_N = 16384)"
for n in 16384 32768 65536; do timeout 120 python tools/time_sym_free.py $n $n 2>&1 | tail -1; done | tee gpurun_out/${T}_sym_free_small_n.jsonl
