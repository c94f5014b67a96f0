#!/bin/bash
# 8-GPU box (r01g kernels): scaling bench 1/2/4/8, C5 Brownian step at 8 ranks, multi-rank parity check at 8 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== check 8"; timeout 600 $TR --nproc-per-node 8 --master-port 29541 tools/check_multi.py 2>gpurun_out/r01g_check8.err | tee gpurun_out/r01g_check_8gpu.json
echo "== c5 full, 8 ranks"
timeout 1200 $TR --nproc-per-node 8 --master-port 29542 tools/bench_c5.py 2> gpurun_out/r01g_c5_8gpu.err | tee gpurun_out/r01g_c5_8gpu.json; tail -2 gpurun_out/r01g_c5_8gpu.err
for n in 8 4 2; do
  echo "== bench $n"
  timeout 900 $TR --nproc-per-node $n --master-port 2955$n bench.py --gpus $n --steps 3 --warmup 3 2> gpurun_out/r01g_bench_${n}gpu.err | tee gpurun_out/r01g_bench_${n}gpu.json | cut -c1-260
done
echo "== bench 1 (same box)"; timeout 900 python bench.py --steps 3 --warmup 3 --no-solve --no-cpu 2> gpurun_out/r01g_bench_1gpu_box8.err | tee gpurun_out/r01g_bench_1gpu_box8.json | cut -c1-260
