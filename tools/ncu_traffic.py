"""Summarise an `ncu --csv` launch list of bench.py into profiles/: per-kernel time share and the
DRAM bytes the MS-BFS level kernels move per 512-lane batch (bench.py's roofline.traffic).

usage: python tools/ncu_traffic.py gpurun_out/launches.csv profiles/r01_traffic.json "<command that was profiled>"
"""
import collections
import csv
import json
import sys

LEVEL = ("pull_list_kernel", "pull_seg_kernel", "pull_tiny_kernel", "pull_warp_kernel", "pull_hub_kernel", "hub_finalize_kernel", "push_kernel",
         "clear_by_bitmap_kernel", "compact_bitmap_kernel", "frontier_stats_kernel")


def main():
    path, out, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
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
    agg = collections.OrderedDict()
    batches = 0
    lvl_bytes = lvl_ms = tot_ms = 0.0
    for r in rows.values():
        name = r["name"].replace("void ", "").replace("pgq::", "").split("(")[0]
        base = name.split("<")[0]
        ms = r.get("gpu__time_duration.sum", 0.0)
        by = r.get("dram__bytes_read.sum", 0.0) + r.get("dram__bytes_write.sum", 0.0)
        a = agg.setdefault(name, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += ms
        a[2] += by
        tot_ms += ms
        if base == "bfs_init_kernel":
            batches += 1
        if base in LEVEL:
            lvl_bytes += by
            lvl_ms += ms
    summary = {
        "command": cmd,
        "note": "ncu --clock-control none launch list; per-launch times are cold-cache and serialised, "
                "shares (not absolutes) are comparable with bench.py's CUDA-event numbers",
        "batches_profiled": batches,
        "level_kernel_dram_bytes_per_batch": lvl_bytes / max(batches, 1),
        "level_kernel_ms_per_batch_under_ncu": lvl_ms / max(batches, 1),
        "level_kernel_share_of_all_kernel_time": lvl_ms / tot_ms if tot_ms else None,
        "kernels": [{"name": k, "launches": v[0], "ms": round(v[1], 3), "share": round(v[1] / tot_ms, 4),
                     "dram_GB": round(v[2] / 1e9, 3)}
                    for k, v in sorted(agg.items(), key=lambda t: -t[1][1])],
    }
    json.dump(summary, open(out, "w"), indent=1)
    print(json.dumps({k: summary[k] for k in summary if k != "kernels"}, indent=1))
    for k in summary["kernels"][:12]:
        print("%-48s n=%5d %9.3f ms %5.1f%% %8.3f GB" % (k["name"][:48], k["launches"], k["ms"], 100 * k["share"],
                                                        k["dram_GB"]))


if __name__ == "__main__":
    main()
