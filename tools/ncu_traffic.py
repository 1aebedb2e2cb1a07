"""Summarise an `ncu --csv` launch list of bench.py into profiles/: per-kernel time share and the
DRAM bytes the MS-BFS level kernels move per 512-lane batch (bench.py's roofline.traffic).

usage: python tools/ncu_traffic.py gpurun_out/launches.csv profiles/r02_traffic.json "<command that was profiled>" \
           [scale edgefactor lanes]
The summary records the hash of the kernel sources and the workload keys it was taken at; bench.py refuses to apply
it to any other build or workload (roofline.traffic: null + the reason).
"""
import collections
import csv
import json
import sys

LEVEL = ("pull_list_kernel", "pull_seg_kernel", "pull_tiny_kernel", "pull_warp_kernel", "pull_hub_kernel", "hub_finalize_kernel", "push_kernel",
         "clear_by_bitmap_kernel", "level_clear_kernel", "scrub_kernel", "zero_next_kernel", "compact_bitmap_kernel",
         "frontier_stats_kernel", "bfs_snapshot_items_kernel")
CSR = ("rw_global_hist_kernel", "rw_gdig_scan_kernel", "rw_scatter_kernel", "rw_hist_kernel", "rs_hist_kernel", "rs_scatter_kernel",
       "offsets_from_sorted_kernel", "scan_")


def main():
    path, out, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
    keys = [int(x) for x in sys.argv[4:7]]
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import src_hash
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
    tot_ms = 0.0
    for r in rows.values():
        name = r["name"].replace("void ", "").replace("pgq::", "").split("(")[0]
        ms = r.get("gpu__time_duration.sum", 0.0)
        by = r.get("dram__bytes_read.sum", 0.0) + r.get("dram__bytes_write.sum", 0.0)
        a = agg.setdefault(name, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += ms
        a[2] += by
        tot_ms += ms
    # Batches of several workers interleave in ncu's (serialised) launch order, so batches are NOT cut out of the list
    # by position (round 2's first version did, and divided by too few batches: 31.7-41.5 GB per batch from the same
    # per-kernel totals).  Instead: the instrumented pass launches the STATS=1 instantiations of the level kernels (last
    # template argument), every other pass the STATS=0 ones, and every pass answers the same pairs, so
    #   instrumented share = launches(STATS=1) / launches(STATS=0 or 1)   over the flagged kernels,
    #   plain batches = bfs_init_kernel launches x (1 - share),
    #   plain bytes = bytes of the STATS=0 instantiations + (1 - share) x bytes of the level kernels without the flag.
    FLAGGED = ("pull_list_kernel", "pull_seg_kernel", "pull_tiny_kernel", "pull_warp_kernel", "pull_hub_kernel", "push_kernel")
    n_init = agg.get("bfs_init_kernel", [0])[0]
    l0 = l1 = 0
    b0 = ms0 = bu = msu = 0.0
    for name, (cnt, ms, by) in agg.items():
        base = name.split("<")[0]
        if base not in LEVEL:
            continue
        if base in FLAGGED and "<" in name:
            flag = name.rstrip(">").split(",")[-1].strip()
            if flag == "1":
                l1 += cnt
            else:
                l0 += cnt
                b0 += by
                ms0 += ms
        else:
            bu += by
            msu += ms
    share = l1 / (l0 + l1) if l0 + l1 else 0.0
    batches = int(round(n_init * (1.0 - share)))
    lvl_bytes = b0 + (1.0 - share) * bu
    lvl_ms = ms0 + (1.0 - share) * msu
    n_check = agg.get("bfs_check_kernel", [0])[0]
    summary = {
        "command": cmd,
        "src_hash": src_hash("msbfs"),
        "scale": keys[0] if len(keys) > 0 else None, "edgefactor": keys[1] if len(keys) > 1 else None,
        "lanes": keys[2] if len(keys) > 2 else None,
        "note": "ncu --clock-control none launch list; per-launch times are cold-cache and serialised, "
                "shares (not absolutes) are comparable with bench.py's CUDA-event numbers",
        "batches_profiled": batches, "batches_instrumented": n_init - batches,
        "levels_per_batch": n_check / max(n_init, 1),
        "level_kernel_dram_bytes_per_batch": lvl_bytes / max(batches, 1),
        "level_kernel_ms_per_batch_under_ncu": lvl_ms / max(batches, 1),
        "level_kernel_share_of_all_kernel_time": lvl_ms / tot_ms if tot_ms else None,
        "how": "the profiled command must run with --loop 0: ncu does not see kernels launched from conditional graph "
               "nodes (the device-side level loop); the host-driven loop launches the same kernels",
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
