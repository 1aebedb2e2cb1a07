#!/bin/bash
# round-2 final capture (1 GPU): traffic launch lists, then the default bench line and the reference arm
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
bash tools/r02_profiles_a.sh
cp gpurun_out/r02_traffic.json gpurun_out/r02_csr_traffic.json profiles/   # so that the bench line below carries roofline.traffic
timeout 1700 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/final_bench_n1.err | cut -c1-300
timeout 600 python bench.py --impl reference --ref-seconds 60 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/final_bench_n1.json'))
print({k:d[k] for k in ('metric','value','unit','n_gpus','steps','warmup','ms_per_step','dtype','gpu_launches','vs_baseline')})
print('e2e', d['e2e']); print('e2e_chunked', d.get('e2e_chunked'))
print('roofline', {k:v for k,v in d['roofline'].items() if k not in ('note','model')})
print('reference_model', d.get('reference_model'))
print('cpu_baseline', d['cpu_baseline'])
print('clocks', d['clocks'])
print('parity', d['parity'])
print('csr', {k:v for k,v in d['csr_build'].items() if k in ('value','unit','ms_per_build','roofline')})
p=d['paths']; print('C4', {k:(p[k].get('pairs_per_s') if isinstance(p[k],dict) else None) for k in ('shortestpath','iterativelength','cheapest_path_length')})
print('C1', {k:v for k,v in d['config1'].items() if k!='workload'})
r=json.load(open('gpurun_out/final_bench_ref.json')); print('REF', {k:r.get(k) for k in ('impl','value','unit','ms_per_step','cpu_baseline','unavailable')})
PY
