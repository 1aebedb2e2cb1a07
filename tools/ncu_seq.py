"""Dev helper: print the kernel sequence of an `ncu --csv` launch list with per-launch metrics."""
import csv, sys, collections
path = sys.argv[1]
start = sys.argv[2] if len(sys.argv) > 2 else None
lines = [l for l in open(path) if not l.startswith('==')]
rows = collections.OrderedDict()
for x in csv.DictReader(lines):
    k = int(x['ID'])
    r = rows.setdefault(k, {'name': x['Kernel Name']})
    v = float(x['Metric Value'].replace(',', ''))
    u = x['Metric Unit']
    m = x['Metric Name']
    if m == 'gpu__time_duration.sum':
        v = v / 1e3 if u == 'us' else v / 1e6 if u == 'ns' else v * 1e3 if u == 's' else v
    if 'bytes' in m:
        v = v * {'byte': 1e-9, 'Kbyte': 1e-6, 'Mbyte': 1e-3, 'Gbyte': 1.0}.get(u, 1.0)
    r[m] = v
seq = list(rows.values())
if start:
    idx = [i for i, r in enumerate(seq) if start in r['name']]
    seq = seq[idx[-1]:]
def short(n):
    n = n.replace('void ', '').replace('pgq::', '')
    return n.split('(pgq')[0].split('(const')[0].split('(unsigned')[0][:52]
for r in seq:
    print("%-52s %8.3f ms  rd %6.2f GB wr %6.2f GB  L2hit %5.1f  warps %5.1f" % (
        short(r['name']), r.get('gpu__time_duration.sum', 0), r.get('dram__bytes_read.sum', 0),
        r.get('dram__bytes_write.sum', 0), r.get('lts__t_sector_hit_rate.pct', 0),
        r.get('sm__warps_active.avg.pct_of_peak_sustained_active', 0)))
