#!/bin/bash
# 8-GPU box: C5 full size, C5 consistency at 2 ranks, bench.py at 2/4/8 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > gpurun_out/r9_gpus.txt
echo "== rigid test"; timeout 600 python -m pytest tests/test_rigid_gpu.py -x -q 2>&1 | tail -2
echo "== c5 1600 bodies, 2 ranks (compare with r8_c5_1600.json)"
timeout 600 $TR --nproc-per-node 2 --master-port 29511 tools/bench_c5.py --bodies 1600 2> gpurun_out/r9_c5_1600_2gpu.err | tee gpurun_out/r9_c5_1600_2gpu.json
echo "== c5 full, 8 ranks"
timeout 1200 $TR --nproc-per-node 8 --master-port 29512 tools/bench_c5.py 2> gpurun_out/r9_c5_8gpu.err | tee gpurun_out/r9_c5_8gpu.json; tail -3 gpurun_out/r9_c5_8gpu.err
for n in 8 4 2; do
  echo "== bench $n"
  timeout 900 $TR --nproc-per-node $n --master-port 2952$n bench.py --gpus $n --steps 3 --warmup 3 2> gpurun_out/r9_bench_${n}gpu.err | tee gpurun_out/r9_bench_${n}gpu.json
done
