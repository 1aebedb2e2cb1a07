#!/bin/bash
# round-2 profile capture (1 GPU): launch lists with DRAM bytes for the MS-BFS level kernels (host-driven loop: ncu
# does not see kernels launched from conditional graph nodes) and for the CSR build, plus one --set full capture of the
# dominant level kernels.  Every profiled command first runs to completion WITHOUT ncu.
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
# dominant level kernels of one 512-lane batch, full set
CMD3="python tools/trace_levels.py 24 512"
PGQ_TRACE_HOSTLOOP=1 timeout 300 $CMD3 > gpurun_out/p_trace_plain.log 2>&1; echo "trace plain rc=$?"; tail -3 gpurun_out/p_trace_plain.log | cut -c1-200
PGQ_TRACE_HOSTLOOP=1 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:pull_ --launch-skip 30 --launch-count 24 -o gpurun_out/p_pull_full $CMD3 > gpurun_out/p_ncu3.log 2>&1; echo "ncu3 rc=$?"
ncu -i gpurun_out/p_pull_full.ncu-rep --page raw --csv > gpurun_out/p_pull_full_raw.csv 2>/dev/null
ncu -i gpurun_out/p_pull_full.ncu-rep --page source --csv --print-source sass > gpurun_out/p_pull_full_sass.csv 2>/dev/null
rm -f gpurun_out/p_pull_full.ncu-rep   # gpurun_out travels back only below 64 MiB: keep the exported pages
gzip -f gpurun_out/p_pull_full_sass.csv
du -sh gpurun_out
ls -la gpurun_out | tail -12
