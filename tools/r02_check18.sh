#!/bin/bash
# round-2 GPU check 18 (2 GPUs): one_process behind a host barrier, strong / part sections
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/c18_bench_n2.json 2> gpurun_out/c18_bench_n2.err
echo "bench n2 rc=$?"
tail -5 gpurun_out/c18_bench_n2.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/c18_bench_n2.json'))
print("N2 GTEPS", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']))
print('strong', d['strong']['ms'], d['strong']['speedup_vs_1'])
print('one_process', {k:d['one_process'].get(k) for k in ('ms','one_device_ms','speedup_vs_1_device','parity_vs_one_device','error')})
p=d['part']
for k in ('allgather','fused','fused_512'): print(k, p.get(k))
print(p.get('parity_vs_single_gpu'), p.get('fused_equals_allgather'), p.get('error'))
PY
timeout 600 python -m pytest tests/test_gpu_parity.py -q -p no:faulthandler -k "multi_gpu_context" 2>&1 | tail -3
