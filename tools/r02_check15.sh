#!/bin/bash
# round-2 GPU check 15 (2 GPUs): loop-graph heuristic, one_process with a per-batch wait per worker
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/graph_build_cost.py > gpurun_out/c15_graph_cost.log 2>&1; cat gpurun_out/c15_graph_cost.log
timeout 900 python -m pytest tests/test_gpu_parity.py -q -p no:faulthandler -k "multi_gpu_context or loop_graph_is_built or device_loop_has_no" > gpurun_out/c15_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/c15_pytest.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c15_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/c15_smoke.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/c15_bench_n2.json 2> gpurun_out/c15_bench_n2.err
echo "bench n2 rc=$?"
tail -5 gpurun_out/c15_bench_n2.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/c15_bench_n2.json'))
print("N2 GTEPS", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']))
print('strong', d['strong']['ms'], d['strong']['speedup_vs_1'])
print('one_process', {k:d['one_process'].get(k) for k in ('ms','one_device_ms','speedup_vs_1_device','parity_vs_one_device','error')})
p=d['part']
for k in ('allgather','fused','fused_512'): print(k, p.get(k))
print(p.get('parity_vs_single_gpu'), p.get('fused_equals_allgather'), p.get('error'))
PY
