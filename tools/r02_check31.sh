#!/bin/bash
# round-2 GPU check 31 (4 GPUs): the driver's scaling command at N=4
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T0=$(date +%s)
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/c31_bench_n4.json 2> gpurun_out/c31_bench_n4.err
echo "bench n4 rc=$? wall=$(( $(date +%s) - T0 )) s"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c31_bench_n4.json'))
print("N4 GTEPS", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']))
print('strong', d['strong']['ms'], d['strong']['speedup_vs_1'])
print('one_process', {k:d['one_process'].get(k) for k in ('ms','one_device_ms','speedup_vs_1_device','parity_vs_one_device','error')})
p=d['part']
for k in ('allgather','fused','fused_512'): print(k, {x:p[k][x] for x in ('ms_per_step','pairs_per_s','value')} if k in p else None)
print({k:v for k,v in p.items() if k in ('parity_vs_single_gpu','fused_equals_allgather','error','workload')})
PY
