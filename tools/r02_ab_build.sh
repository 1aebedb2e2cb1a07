#!/bin/bash
# builds library variants for the kernel-cost bisect: gpurun_out is not shipped, so they go to tools/ab/
set -e
cd "$(dirname "$0")/../duckdb-pgq.old_b200"
mkdir -p ../tools/ab
ARCH="-gencode arch=compute_100a,code=sm_100a"
FL="$ARCH -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr"
OBJS=$(ls build/*.o | grep -v msbfs.o)
for v in NODYN ND ALL; do
  D="-DPGQ_AB_$v"
  [ $v = ALL ] && D="-DPGQ_AB_NODONE -DPGQ_AB_NODYN -DPGQ_AB_DENSE"; [ $v = ND ] && D="-DPGQ_AB_NODONE -DPGQ_AB_DENSE"
  nvcc $FL $D -c csrc/msbfs.cu -o ../tools/ab/msbfs_$v.o 2>/dev/null &
done
wait
for v in NODYN ND ALL; do
  nvcc $ARCH -shared -o ../tools/ab/libpgq_$v.so $OBJS ../tools/ab/msbfs_$v.o -lcudart_static -lpthread -ldl -lrt
  rm ../tools/ab/msbfs_$v.o
done
ls -la ../tools/ab
