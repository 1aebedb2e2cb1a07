#!/bin/bash
# round-2 profile capture, launch lists only (see tools/r02_profiles.sh for the --set full part)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
CMD1="python bench.py --loop 0 --pairs 2048 --steps 1 --warmup 1 --no-cpu-baseline --no-csr --no-paths"
timeout 600 $CMD1 > gpurun_out/p_bench_plain.json 2> gpurun_out/p_bench_plain.err; echo "plain rc=$?"
timeout 1500 ncu --metrics $M --clock-control none -c 4000 --csv --log-file gpurun_out/p_launches.csv $CMD1 > gpurun_out/p_ncu1.log 2>&1; echo "ncu1 rc=$?"
python tools/ncu_traffic.py gpurun_out/p_launches.csv gpurun_out/r02_traffic.json "ncu --metrics $M --clock-control none -c 4000 --csv $CMD1" 24 16 512 | tail -22
CMD2="python tools/csr_only.py 22 1 noprep"
timeout 300 $CMD2 > gpurun_out/p_csr_plain.log 2>&1; echo "csr plain rc=$?"; tail -2 gpurun_out/p_csr_plain.log
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/p_csr_launches.csv $CMD2 > gpurun_out/p_ncu2.log 2>&1; echo "ncu2 rc=$?"
python tools/ncu_csr_traffic.py gpurun_out/p_csr_launches.csv gpurun_out/r02_csr_traffic.json "ncu --metrics $M --clock-control none -c 400 --csv $CMD2" 22 16 | tail -12
