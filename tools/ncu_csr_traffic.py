"""Summarise an `ncu --csv` launch list of `tools/csr_only.py S 1 noprep` into profiles/r02_csr_traffic.json: DRAM bytes
of one forward CSR build (bench.py's csr_build.roofline.traffic).

usage: python tools/ncu_csr_traffic.py gpurun_out/csr_launches.csv profiles/r02_csr_traffic.json "<command>" scale edgefactor
"""
import collections
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import src_hash  # noqa: E402


def main():
    path, out, cmd, scale, ef = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
    rows = collections.OrderedDict()
    for x in csv.DictReader(l for l in open(path) if not l.startswith("==")):
        r = rows.setdefault(int(x["ID"]), {"name": x["Kernel Name"]})
        v = float(x["Metric Value"].replace(",", ""))
        u, m = x["Metric Unit"], x["Metric Name"]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)
        if "bytes" in m:
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
        r[m] = v
    kern = []
    tot_b = tot_ms = 0.0
    for r in rows.values():
        name = r["name"].replace("void ", "").replace("pgq::", "").split("(")[0]
        if name.startswith(("gen_", "rmat_", "probe_")):
            continue
        ms = r.get("gpu__time_duration.sum", 0.0)
        by = r.get("dram__bytes_read.sum", 0.0) + r.get("dram__bytes_write.sum", 0.0)
        kern.append({"name": name, "ms": round(ms, 4), "dram_GB": round(by / 1e9, 4)})
        tot_b += by
        tot_ms += ms
    summary = {"command": cmd, "src_hash": src_hash("csr"), "scale": scale, "edgefactor": ef,
               "dram_bytes_per_build": tot_b, "kernel_ms_under_ncu": tot_ms, "kernels": kern}
    json.dump(summary, open(out, "w"), indent=1)
    print(json.dumps({k: v for k, v in summary.items() if k != "kernels"}, indent=1))
    for k in kern:
        print("%-52s %8.3f ms %8.3f GB" % (k["name"][:52], k["ms"], k["dram_GB"]))


if __name__ == "__main__":
    main()
