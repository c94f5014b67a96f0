#!/bin/bash
# r01e: full default bench, reference arm, all-block rates, ST rates, DRAM traffic of the flagship launch at N = 1M
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r01e_pytest_gpu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/r01e_pytest_gpu.log
echo "== bench"; timeout 900 python bench.py > gpurun_out/r01e_bench.json 2> gpurun_out/r01e_bench.err; echo "rc=$?"; cat gpurun_out/r01e_bench.json
echo "== rates"; timeout 600 python tools/tune.py 400000:UF:1:0:open 400000:UF:0:0:open 400000:UT:1:0:open 400000:WF:1:0:open 400000:WT:1:0:open 400000:UF_UT:1:0:open 200000:UF:1:0:xy 400000:UF:1:1:open 400000:WT:1:1:open 15000:UF:1:0:open 2>/dev/null | tail -1 | tee gpurun_out/r01e_rates.json
echo "== st"; timeout 600 python tools/bench_st.py 200000 | tee gpurun_out/r01e_st_rates.json
echo "== traffic"; timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:mf_kernel -c 2 --csv --log-file gpurun_out/r01e_traffic_n1m.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-parity --no-solve > gpurun_out/r01e_traffic.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r01e_traffic_n1m.csv
