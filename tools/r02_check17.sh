#!/bin/bash
# round-2 GPU check 17: where a cheapest_path_length round spends its time (ncu on one 64-lane batch)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/sssp_probe.py 64 sssp_slots=1 2>&1 | tail -3
timeout 300 python tools/sssp_probe.py 256 sssp_slots=1 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 256 sssp_slots=4 2>&1 | tail -1
timeout 300 python tools/sssp_probe.py 1024 sssp_slots=2 2>&1 | tail -1
PROBE_ONCE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:sssp -c 200 --csv --log-file gpurun_out/c17_launches.csv python tools/sssp_probe.py 64 sssp_slots=1 > gpurun_out/c17_ncu1.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.DictReader(l for l in open('gpurun_out/c17_launches.csv') if not l.startswith('=='))]
for r in rows[:120]:
    print(r['ID'], r['Kernel Name'][:40], r['Metric Value'], r['Metric Unit'], r.get('Grid Size'))
PY
PROBE_ONCE=1 timeout 900 ncu --set full --clock-control none -k regex:sssp16_relax_kernel --launch-skip 4 --launch-count 3 -o gpurun_out/c17_relax python tools/sssp_probe.py 64 sssp_slots=1 > gpurun_out/c17_ncu2.log 2>&1
ncu -i gpurun_out/c17_relax.ncu-rep --page raw --csv > gpurun_out/c17_relax_raw.csv 2>/dev/null
ls -la gpurun_out/ | tail -5
