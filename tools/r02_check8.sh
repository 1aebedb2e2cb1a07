#!/bin/bash
# round-2 GPU check 8: does the ROUND-1 library still run the dense level in 3.1 ms on today's box?
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv > gpurun_out/c8_gpu.txt
for v in R01 BASE ALL R01 BASE; do
  lib=tools/ab/libpgq_$v.so; opt=""; hl=""
  [ $v = BASE ] && lib=duckdb-pgq.old_b200/libpgq_b200.so
  [ $v = ALL ] && opt="dyn=0" && hl=1
  echo "== $v $opt" >> gpurun_out/c8_ab.log
  PGQ_TRACE_HOSTLOOP=$hl PGQ_B200_LIB=$PWD/$lib timeout 300 python tools/trace_levels.py 24 512 $opt 2>&1 | grep "iter [1-8]\|level_kernel_ms\|rror" | sed -E 's/next:.*//; s/.*level_kernel_ms.: ([0-9.]+).*/total \1/' >> gpurun_out/c8_ab.log
done
cat gpurun_out/c8_ab.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,sm__inst_executed.sum,lts__t_sector_hit_rate.pct
PGQ_B200_LIB=$PWD/tools/ab/libpgq_R01.so timeout 600 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/c8_ncu_r01.csv python tools/trace_levels.py 24 512 > gpurun_out/c8_ncu_r01.log 2>&1
timeout 600 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/c8_ncu_new.csv python tools/trace_levels.py 24 512 > gpurun_out/c8_ncu_new.log 2>&1
