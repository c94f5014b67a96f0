#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_mobility_gpu.py -q -k "symmetric or default_selection" 2>&1 | tail -1
for n in 8 4 2; do
  echo "== bench $n"
  timeout 900 $TR --nproc-per-node $n --master-port 2956$n bench.py --gpus $n --steps 3 --warmup 3 2> gpurun_out/r01h_bench_${n}gpu.err | tee gpurun_out/r01h_bench_${n}gpu.json | cut -c1-200
done
