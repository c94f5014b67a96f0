#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r01f_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r01f_pytest_gpu.log
echo "== multi check"; timeout 600 $TR --nproc-per-node 2 --master-port 29531 tools/check_multi.py 2>gpurun_out/r01f_check2.err | tee gpurun_out/r01f_check_2gpu.json
echo "== bench 1"; timeout 900 python bench.py --steps 3 --warmup 3 --no-solve > gpurun_out/r01f_bench_1gpu.json 2> gpurun_out/r01f_bench_1gpu.err; echo "rc=$?"; cat gpurun_out/r01f_bench_1gpu.json
echo "== bench 2"; timeout 900 $TR --nproc-per-node 2 --master-port 29532 bench.py --gpus 2 --steps 3 --warmup 3 2> gpurun_out/r01f_bench_2gpu.err | tee gpurun_out/r01f_bench_2gpu.json
