#!/bin/bash
# round-2 GPU check 36 (2 GPUs): the reference arm under torchrun, as the driver launches it
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T0=$(date +%s)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 --ref-seconds 30 > gpurun_out/c36_ref_n2.json 2> gpurun_out/c36_ref_n2.err
echo "ref n2 rc=$? wall=$(( $(date +%s) - T0 )) s"; cat gpurun_out/c36_ref_n2.json | cut -c1-900; tail -2 gpurun_out/c36_ref_n2.err | cut -c1-200
