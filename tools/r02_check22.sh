#!/bin/bash
# round-2 GPU check 22 (8 GPUs): the driver's scaling command at N=8, wall-clocked, and the reference arm under torchrun
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L | head -8
T0=$(date +%s)
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/c22_bench_n8.json 2> gpurun_out/c22_bench_n8.err
echo "bench n8 rc=$? wall=$(( $(date +%s) - T0 )) s"
tail -5 gpurun_out/c22_bench_n8.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/c22_bench_n8.json'))
print("N8 GTEPS", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'])
print('strong', d['strong']['ms'], d['strong']['speedup_vs_1'])
print('one_process', {k:d['one_process'].get(k) for k in ('ms','one_device_ms','speedup_vs_1_device','parity_vs_one_device','error')})
p=d['part']
for k in ('allgather','fused','fused_512'): print(k, p.get(k))
print({k:v for k,v in p.items() if k not in ('allgather','fused','fused_512')})
print(d['clocks'])
PY
