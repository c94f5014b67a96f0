#!/bin/bash
# run tools/tune.py over every library in build_tune/
mkdir -p gpurun_out
for so in build_tune/*.so; do
  RMB_LIB=$PWD/$so timeout 300 python tools/tune.py "$@" 2>&1 | tail -1
done | tee gpurun_out/tune.jsonl
