#!/bin/bash
# One GPU session on a gpurun box (1 GPU): tools/gpu_session.sh <tag> <stage> [<stage> ...]
# stages: tests smoke bench ref sched per rates ncu     (outputs under gpurun_out/<tag>_*)
T=$1; shift
mkdir -p gpurun_out
NCU_FULL="ncu --set full --clock-control none --import-source on"
for stage in "$@"; do
case $stage in
tests) echo "== pytest gpu"; timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_pytest_gpu.log;;
smoke) echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_smoke.log;;
bench) echo "== bench"; timeout 1200 python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench.err; echo "rc=$?"; cut -c1-1500 gpurun_out/${T}_bench_1gpu.json; tail -3 gpurun_out/${T}_bench.err;;
ref) echo "== ref"; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "rc=$?"; cut -c1-900 gpurun_out/${T}_ref.json; tail -3 gpurun_out/${T}_ref.err;;
sched) echo "== schedule variants: one rank's share of an 8-rank job, and the whole product, N = 1M"
  for so in default build_tune/chunk8_32.so build_tune/chunk2_8.so build_tune/items32.so; do
    [ $so = default ] && unset RMB_LIB || export RMB_LIB=$PWD/$so
    timeout 300 python tools/time_rank_share.py 8 1000000 2>&1 | tail -1
    timeout 300 python tools/run_case.py UF_wall_open_f64 --n 1000000 --reps 2 --warm 1 2>&1 | tail -1
    timeout 300 python tools/run_case.py UF_wall_open_f64 --n 100000 --reps 5 --warm 2 2>&1 | tail -1
  done | tee gpurun_out/${T}_sched.jsonl; unset RMB_LIB;;
per) echo "== pseudo-periodic variants (9 boxes)"
  for so in default build_tune/per_u1.so build_tune/per_u4.so build_tune/per_u4_b4.so build_tune/per_u2_b4.so; do
    [ $so = default ] && unset RMB_LIB || export RMB_LIB=$PWD/$so
    timeout 300 python tools/run_case.py UF_wall_xy_f64,UF_wall_xy_f64_nofar,UF_wall_xy_f64_row,UF_free_xy_f64,UT_wall_xy_f64,WF_wall_xy_f64,WT_wall_xy_f64,UFUT_wall_xy_f64 --n 200000 2>&1 | tail -1
  done | tee gpurun_out/${T}_per.jsonl; unset RMB_LIB;;
rates) echo "== rates"
  timeout 600 python tools/run_case.py UF_wall_open_f64,UF_wall_open_f64_row,UF_wall_open_f32,UT_wall_open_f64,WF_wall_open_f64,WT_wall_open_f64,UFUT_wall_open_f64,UF_free_open_f64,st_wall_open_f64,st_free_open_f64,st_fs_open_f64,st_wall_open_f32 --n 400000 2>&1 | tail -1 | tee gpurun_out/${T}_rates_dilute.json
  timeout 600 python tools/run_case.py UF_wall_open_f64,UF_wall_open_f64_row,UF_wall_open_f32,WT_wall_open_f64,st_wall_open_f64 --n 200000 --cloud dense 2>&1 | tail -1 | tee gpurun_out/${T}_rates_dense.json
  timeout 600 python tools/run_case.py UF_wall_open_f64,UF_wall_open_f64_row,UF_wall_open_f32,WT_wall_open_f64,st_wall_open_f64 --n 200000 --cloud c1box 2>&1 | tail -1 | tee gpurun_out/${T}_rates_c1box.json;;
ncu) echo "== ncu full captures"
  for spec in ${NCU_SPECS:-"UF_wall_open_f64:200000:sym_kernel:flagship" "UF_wall_xy_f64:100000:sym_kernel:uf_xy"}; do
    IFS=: read c n k name <<< "$spec"
    timeout 600 $NCU_FULL -k regex:$k -s 1 -c 1 -o /tmp/${T}_${name}_full -f python tools/run_case.py $c --n $n --reps 1 --warm 1 > gpurun_out/${T}_${name}_ncu.log 2>&1; echo "$name rc=$?"
    # the .ncu-rep files are ~15 MB each and gpurun_out is capped at 64 MiB: keep the raw page as CSV (and the source page of the flagship)
    ncu -i /tmp/${T}_${name}_full.ncu-rep --page raw --csv > gpurun_out/${T}_${name}_raw.csv 2>/dev/null
    [ $name = flagship ] && ncu -i /tmp/${T}_${name}_full.ncu-rep --page source --csv > gpurun_out/${T}_${name}_source.csv 2>/dev/null
  done
  echo "== launch list"; ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --n 200000 --steps 2 --warmup 1 --no-cpu --no-extra --profiling > gpurun_out/${T}_ncu_launch.log 2>&1; echo "rc=$?";;
multi) G=${GPUS:-8}; echo "== $G GPUs: bench (with c5 at 8), per-rank breakdown, reference arm under torchrun"
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511"
  timeout 900 $TR bench.py --gpus $G --steps 5 --warmup 3 > gpurun_out/${T}_bench_${G}gpu.json 2> gpurun_out/${T}_bench_${G}gpu.err; echo "rc=$?"; cut -c1-600 gpurun_out/${T}_bench_${G}gpu.json; tail -3 gpurun_out/${T}_bench_${G}gpu.err
  timeout 300 $TR tools/rank_breakdown.py > gpurun_out/${T}_rank_breakdown_${G}gpu.json 2> gpurun_out/${T}_rank_breakdown.err; echo "rc=$?"; cat gpurun_out/${T}_rank_breakdown_${G}gpu.json
  timeout 300 $TR bench.py --impl reference --gpus $G --steps 1 --warmup 0 --no-solve > gpurun_out/${T}_ref_${G}gpu.json 2> gpurun_out/${T}_ref_${G}gpu.err; echo "rc=$?"; cut -c1-400 gpurun_out/${T}_ref_${G}gpu.json;;
c2) echo "== config-2 solve breakdown; row vs symmetric kernel around the 16,384-blob threshold"
  timeout 300 python tools/time_c2.py 2>&1 | tail -1 | tee gpurun_out/${T}_c2_breakdown.json
  for n in 8000 12000 15000 20000 30000; do timeout 120 python tools/run_case.py UF_wall_open_f64_row,UF_wall_open_f64_sym --n $n --reps 20 --warm 3 2>&1 | tail -1; done | tee gpurun_out/${T}_row_vs_sym_small.jsonl;;
smalln) echo "== small-N latency: this build vs the r01 source split (>= 1024 sources per CTA)"
  for so in default build_tune/split1024.so; do
    [ $so = default ] && unset RMB_LIB || export RMB_LIB=$PWD/$so
    timeout 300 python tools/time_small_n.py 1000 2000 5000 15000 50000 2>&1 | tail -1
  done | tee gpurun_out/${T}_small_n.jsonl; unset RMB_LIB;;
traffic) echo "== DRAM traffic of one flagship launch at N = 1M"
  timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:sym_kernel -s 1 -c 1 --csv --log-file gpurun_out/${T}_traffic_n1m.csv python tools/run_case.py UF_wall_open_f64 --n 1000000 --reps 1 --warm 1 > gpurun_out/${T}_traffic.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/${T}_traffic_n1m.csv;;
esac
done
ls -la gpurun_out | tail -30
