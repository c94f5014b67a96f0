#!/bin/bash
# multi-GPU check + bench on the N GPUs of this box.  Usage: tools/gpu_scaling.sh <tag> <n_gpus>
mkdir -p gpurun_out
T=${1:-r01m}; n=${2:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== check $n"; timeout 600 $TR --nproc-per-node $n --master-port 2946$n tools/check_multi.py 2> gpurun_out/${T}_check_${n}gpu.err | tee gpurun_out/${T}_multi_rank_parity_${n}gpu.json | cut -c1-1200
echo "== bench $n"; timeout 900 $TR --nproc-per-node $n --master-port 2956$n bench.py --gpus $n --steps 3 --warmup 3 2> gpurun_out/${T}_bench_${n}gpu.err | tee gpurun_out/${T}_bench_${n}gpu.json | cut -c1-300
