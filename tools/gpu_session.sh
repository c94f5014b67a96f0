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
  for spec in "UF_wall_open_f64:200000:sym_kernel:flagship" "UF_wall_xy_f64:100000:sym_kernel:uf_xy" "st_wall_open_f64:100000:st_kernel:st_wall" "UF_wall_open_f32:200000:mf_kernel:uf_f32" "UF_wall_open_f64_row:200000:mf_kernel:uf_row"; do
    IFS=: read c n k name <<< "$spec"
    timeout 600 $NCU_FULL -k regex:$k -s 1 -c 1 -o /tmp/${T}_${name}_full -f python tools/run_case.py $c --n $n --reps 1 --warm 1 > gpurun_out/${T}_${name}_ncu.log 2>&1; echo "$name rc=$?"
    # the .ncu-rep files are ~15 MB each and gpurun_out is capped at 64 MiB: keep the raw page as CSV (and the source page of the flagship)
    ncu -i /tmp/${T}_${name}_full.ncu-rep --page raw --csv > gpurun_out/${T}_${name}_raw.csv 2>/dev/null
    [ $name = flagship ] && ncu -i /tmp/${T}_${name}_full.ncu-rep --page source --csv > gpurun_out/${T}_${name}_source.csv 2>/dev/null
  done
  echo "== launch list"; ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --n 200000 --steps 2 --warmup 1 --no-cpu > gpurun_out/${T}_ncu_launch.log 2>&1; echo "rc=$?";;
ab2) echo "== A/B of decode variants"
  for so in build_tune/r01.so build_tune/c0_nopad.so build_tune/c1_pad3.so build_tune/c2_smem.so build_tune/c4_nopair.so build_tune/c2_smem_nopad.so build_tune/c4_nopair_nopad.so build_tune/r01_w200.so; do
    export RMB_LIB=$PWD/$so
    timeout 300 python tools/run_case.py UF_wall_open_f64,WT_wall_open_f64,UF_wall_xy_f64 --n 400000 --reps 3 2>&1 | tail -1
    timeout 300 python tools/time_rank_share.py 8 1000000 2>&1 | tail -1
  done | tee gpurun_out/${T}_ab2.jsonl; unset RMB_LIB;;
ab) echo "== A/B r01 library vs this build (same box)"
  nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit,clocks_event_reasons.active --format=csv
  for so in build_tune/r01.so default build_tune/r01.so default; do
    [ $so = default ] && unset RMB_LIB || export RMB_LIB=$PWD/$so
    timeout 300 python tools/run_case.py UF_wall_open_f64,UF_wall_open_f64_row,WT_wall_open_f64 --n 400000 --reps 4 2>&1 | tail -1
    timeout 300 python tools/run_case.py UF_wall_open_f64 --n 1000000 --reps 2 --warm 1 2>&1 | tail -1
  done | tee gpurun_out/${T}_ab.jsonl; unset RMB_LIB
  echo "== light ncu metrics, both libraries"
  for so in build_tune/r01.so default; do
    [ $so = default ] && unset RMB_LIB || export RMB_LIB=$PWD/$so
    ncu --metrics sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum --clock-control none -k regex:sym_kernel -s 1 -c 1 --csv --log-file gpurun_out/${T}_ab_ncu_$(basename $so .so).csv python tools/run_case.py UF_wall_open_f64 --n 400000 --reps 1 --warm 1 > /dev/null 2>&1; echo "rc=$?"
  done; unset RMB_LIB
  (nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/${T}_smi.csv &) ; python tools/run_case.py UF_wall_open_f64 --n 1000000 --reps 3 --warm 1 | tail -1; pkill -n nvidia-smi; tail -8 gpurun_out/${T}_smi.csv;;
esac
done
ls -la gpurun_out | tail -30
