#!/bin/bash
# round-2 GPU check 17b: cheapest_path_length probe (timings + launch list of one batch)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -p no:faulthandler -k "cheapest" 2>&1 | tail -3
timeout 300 python tools/sssp_probe.py 64 sssp_slots=1 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 256 sssp_slots=1 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=1 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=2 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=4 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=4 sssp_narrow=0 2>&1 | tail -1
PROBE_ONCE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:sssp -c 200 --csv --log-file gpurun_out/c17_launches.csv python tools/sssp_probe.py 64 sssp_slots=1 > gpurun_out/c17_ncu1.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.DictReader(l for l in open('gpurun_out/c17_launches.csv') if not l.startswith('=='))]
agg=collections.OrderedDict()
for r in rows:
    k=r['Kernel Name'].split('(')[0][:40]
    agg.setdefault(k,[]).append(float(r['Metric Value'])/1e3)
for k,v in agg.items(): print(k, len(v), "total %.1f us"%sum(v), " ".join("%.0f"%x for x in v[:16]))
PY
