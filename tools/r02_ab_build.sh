#!/bin/bash
# builds library variants for kernel A/B runs: gpurun_out is not shipped, so they go to tools/ab/ (git-ignored)
# usage: tools/r02_ab_build.sh NAME "-DFLAG ..." [NAME "-D..."]...
set -e
cd "$(dirname "$0")/../duckdb-pgq.old_b200"
mkdir -p ../tools/ab
ARCH="-gencode arch=compute_100a,code=sm_100a"
FL="$ARCH -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr"
OBJS=$(ls build/*.o | grep -v msbfs.o)
names=()
while [ $# -ge 2 ]; do
  nvcc $FL $2 -Xptxas -v -c csrc/msbfs.cu -o ../tools/ab/msbfs_$1.o 2> ../tools/ab/msbfs_$1.ptxas.log &
  names+=($1); shift 2
done
wait
for v in "${names[@]}"; do
  nvcc $ARCH -shared -o ../tools/ab/libpgq_$v.so $OBJS ../tools/ab/msbfs_$v.o -lcudart_static -lpthread -ldl -lrt
  rm ../tools/ab/msbfs_$v.o
  for k in pull_list_kernelILi512ELi2ELb0ELb0 pull_warp_kernelILi512ELb0ELb0 pull_seg_kernelILi512ELi8ELb0; do
    echo "$v $k: $(grep -A1 $k ../tools/ab/msbfs_$v.ptxas.log | grep Used | head -1 | cut -c1-90)"
  done
done
ls -la ../tools/ab
