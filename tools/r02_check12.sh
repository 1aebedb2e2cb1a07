#!/bin/bash
# round-2 GPU check 12 (2 GPUs): one-process multi-GPU context test, bench --gpus 2 (strong / one_process / part sections)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PGQ_B200_BACKTRACE=1
nvidia-smi -L > gpurun_out/c12_gpus.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -q -p no:faulthandler -k "multi_gpu_context or sharded_query" > gpurun_out/c12_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/c12_pytest.log
tail -15 gpurun_out/c12_pytest.log | cut -c1-250
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/c12_bench_n2.json 2> gpurun_out/c12_bench_n2.err
echo "bench n2 rc=$?"
tail -15 gpurun_out/c12_bench_n2.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/c12_bench_n2.json'))
print("N2 GTEPS", round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']))
for k in ('strong','one_process','part'):
    print(k, json.dumps(d.get(k))[:1500])
print(d['parity'])
PY
