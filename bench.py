#!/usr/bin/env python
"""bench.py — MS-BFS GTEPS / pair-queries/s and CSR-build GE/s on B200 (BASELINE.json metric).

Step  = one pass of iterativelength over the whole pair set of the workload (config 3 of
        BASELINE.json: R-MAT scale 24 ef16, 65 536 (src,dst) pairs, 512-lane batches).
value = traversed edges per second (TEPS numerator of SURVEY.md §8(d), counted on the device in
        an untimed instrumented pass — it is a deterministic function of the inputs) with the
        CSR and the pairs already resident in HBM.
e2e   = the same metric through the host-buffer C-ABI call (pgq_iterativelength): pairs H2D and
        results D2H inside the timed region.
N>1   : one process per GPU (torchrun), CSR replicated, every rank answers its own pair set
        (source-batch sharding, no data-path collective) -> "scaling": "weak".
--impl reference : the CPU restatement (oracle/) in the reference's execution shape
        (<=1024-row calls, 512-lane std::bitset batches) on the host cores, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED_BASE = 0x5eed0000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=int, default=24)
    ap.add_argument("--edgefactor", type=int, default=16)
    ap.add_argument("--pairs", type=int, default=65536)
    ap.add_argument("--csr-scale", type=int, default=22)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-csr", action="store_true")
    ap.add_argument("--no-paths", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0)
    ap.add_argument("--ref-seconds", type=float, default=240.0,
                    help="--impl reference: wall-clock box for the timed CPU passes")
    ap.add_argument("--loop", type=int, default=1,
                    help="1 = device-side level loop (CUDA graph), 0 = host-driven loop (one sync per level)")
    ap.add_argument("--streams", type=int, default=0,
                    help="batches of one call in flight at once (0 = library default)")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, nm in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_pairs(V, v_off, n, seed):
    """src,dst uniform over vertices with out-degree > 0; no NULL / trivial rows."""
    deg = np.diff(v_off[:V + 1])
    cand = np.nonzero(deg > 0)[0].astype(np.int64)
    rng = np.random.default_rng(seed)
    src = cand[rng.integers(0, len(cand), n)]
    dst = cand[rng.integers(0, len(cand), n)]
    same = src == dst
    dst[same] = cand[(np.searchsorted(cand, dst[same]) + 1) % len(cand)]
    return src, dst


def lib_hash():
    """sha256 (16 hex digits) of the built CUDA library: measured-traffic profiles are tied to it."""
    import hashlib
    import pgq_b200
    try:
        return hashlib.sha256(open(pgq_b200.LIB_PATH, "rb").read()).hexdigest()[:16]
    except Exception:
        return None


SRC_SETS = {  # the sources a profiled kernel family is compiled from (common.cuh holds host-side structs only)
    "msbfs": ("msbfs_kernels.cuh", "msbfs.cu"),
    "csr": ("csr_build.cu", "radix_sort.cu", "radix_sort_wide.cu", "scan.cu"),
}


def src_hash(which="msbfs"):
    """sha256 (16 hex digits) over the sources of one kernel family: survives a rebuild of the same code on another
    box, and an edit to an unrelated kernel (say sssp.cu) does not make the MS-BFS traffic profile stale."""
    import hashlib
    import re
    h = hashlib.sha256()
    for f in SRC_SETS[which]:
        text = open(os.path.join(ROOT, "duckdb-pgq.old_b200", "csrc", f), "r", errors="replace").read()
        text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)  # comments and layout do not change the kernels
        text = re.sub(r"//[^\n]*", " ", text)
        h.update(" ".join(text.split()).encode())
    return h.hexdigest()[:16]


def load_traffic(name, want):
    """Measured DRAM bytes from a committed ncu summary (profiles/<name>), or (None, why).  The summary
    must have been taken on the same kernel sources and the same workload keys (`want`), else it is stale and
    NOT applied (a constant file silently going stale)."""
    path = os.path.join(ROOT, "profiles", name)
    try:
        prof = json.load(open(path))
    except Exception:
        return None, "no profiles/%s" % name
    which = "csr" if "csr" in name else "msbfs"
    if prof.get("src_hash") != src_hash(which):
        return None, "profiles/%s was taken on other kernel sources (src_hash %s != %s)" % (
            name, prof.get("src_hash"), src_hash(which))
    for k, v in want.items():
        if prof.get(k) != v:
            return None, "profiles/%s was taken at %s=%s, this run has %s" % (name, k, prof.get(k), v)
    return prof, "profiles/%s (%s)" % (name, prof.get("command", ""))


def csr_parity(ctx, csr_id, V, E, ov, oe):
    """memcmp of the GPU-built CSR (reference layout, int64) with the oracle's arrays built on the host from
    host-generated rows.  Downloads `e` in slices of the host array to bound memory."""
    import pgq_b200
    gv = np.empty(V + 2, dtype=np.int64)
    ge = np.empty(max(E, 1), dtype=np.int64)
    ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, csr_id, gv.ctypes.data_as(pgq_b200.I64P),
                                      ge.ctypes.data_as(pgq_b200.I64P), None))
    same = bool(np.array_equal(gv, ov) and np.array_equal(ge[:E], oe))
    return "identical" if same else "MISMATCH"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_sample_threads(V, requested):
    """threads for the CPU leg: bounded by cores and by 3 bitmaps of V*64 B per thread."""
    cores = host_cores()
    t = requested or min(cores, 16)
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
        per = 3 * V * 64 + (64 << 20)
        t = max(1, min(t, int(avail * 0.5 // per)))
    except Exception:
        pass
    return t, cores


def run_cpu_sample(V, v, e, src, dst, nthreads, rows_per_thread=512):
    """The reference's execution shape on host cores: each worker answers one <=1024-row call
    (here rows_per_thread rows = 512-lane batches).  Returns (seconds, rows, edges_lane)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    orc.lib()
    res = [None] * nthreads

    def work(t):
        lo = t * rows_per_thread
        res[t] = orc.iterativelength(V, v, e, src[lo:lo + rows_per_thread], dst[lo:lo + rows_per_thread],
                                     stats=True)

    ths = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    secs = time.perf_counter() - t0
    edges = sum(r[2]["edges_lane"] for r in res)
    return secs, nthreads * rows_per_thread, edges, res


def dist_setup(n_gpus):
    """torchrun plumbing (NCCL over NVLink is only used for the barrier / max-over-ranks)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        # NCCL announces its version on stdout when the communicator comes up; stdout carries exactly one
        # JSON line, so that goes to stderr
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            t = torch.zeros(1, device="cuda:%d" % local)
            dist.all_reduce(t)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    return rank, world, local, dist


def barrier_max(dist, local, value):
    if dist is None:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device="cuda:%d" % local)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sum(dist, local, value):
    if dist is None:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device="cuda:%d" % local)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def bench_csr(ctx, scale, ef, steps, warmup, peak_gbs, with_parity=True):
    """BASELINE config 2: CSR build from the edge table, R-MAT scale 22 ef16."""
    import pgq_b200
    V, E = 1 << scale, ef << scale
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
    parity = None
    if with_parity:
        # config-size parity (untimed): the oracle builds the CSR on the host from HOST-generated rows (its own
        # generator, bit-identical stream) and the GPU's arrays must memcmp-equal it
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        hs, hd = orc.gen_rmat(SEED_BASE + scale, scale, E=E, nthreads=min(host_cores(), 64))
        ov, oe, _ = orc.csr_build(V, hs, hd, want_eids=False)
        del hs, hd
        ctx.csr_build_device(100, V, E, d_src, d_dst)
        parity = csr_parity(ctx, 100, V, E, ov, oe)
        ctx.csr_delete(100)
        del ov, oe
    kern_ms = []
    for i in range(warmup + steps):
        ctx.flush_l2()
        ctx.csr_build_device(100, V, E, d_src, d_dst)
        if i >= warmup:
            kern_ms.append(ctx.stats()["csr_build_ms"])
        ctx.csr_delete(100)
    # e2e: edge table in pinned HOST memory -> H2D inside the timed region
    h_src, p1 = pgq_b200.pinned_empty(E, np.int64)
    h_dst, p2 = pgq_b200.pinned_empty(E, np.int64)
    ctx.d2h(h_src, d_src)
    ctx.d2h(h_dst, d_dst)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    e2e_ms = []
    for i in range(2 + steps):
        t0 = time.perf_counter()
        ctx.csr_build(100, V, h_src, h_dst)
        dt = (time.perf_counter() - t0) * 1e3
        if i >= 2:
            e2e_ms.append(dt)
        ctx.csr_delete(100)
    pgq_b200.pinned_free(p1)
    pgq_b200.pinned_free(p2)
    ms = float(np.mean(kern_ms))
    algo = 32 * E + 24 * V
    prof, prof_src = load_traffic("r02_csr_traffic.json", {"scale": scale, "edgefactor": ef})
    traffic = float(prof["dram_bytes_per_build"]) if prof else None
    return {"workload": "csr_build rmat scale %d ef%d (V=%d E=%d)" % (scale, ef, V, E),
            "parity": parity, "parity_what": "v[V+2] and e[E] memcmp-equal the oracle's CSR built on the host from "
                                             "host-generated rows" if parity else None,
            "value": E / (ms * 1e-3) / 1e9, "unit": "GE/s", "ms": ms,
            "e2e": {"value": E / (float(np.mean(e2e_ms)) * 1e-3) / 1e9, "unit": "GE/s",
                    "ms": float(np.mean(e2e_ms)), "h2d_bytes_per_step": 16 * E, "d2h_bytes_per_step": 0},
            "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak_gbs, "unit": "GB/s",
                         "frac": algo / (ms * 1e-3) / 1e9 / peak_gbs, "traffic": traffic,
                         "traffic_source": prof_src,
                         "dram_frac": traffic / (ms * 1e-3) / 1e9 / peak_gbs if traffic else None,
                         "algo_bytes": algo, "kernels": "stable radix sort by source: global bucket histogram + two 11-bit passes (tile-reordered scatter, decoupled look-back) + offsets from sorted keys (whole build)"}}


def knows_params(V):
    """Affine spread of the knows generator (same arithmetic as oracle/pgq_oracle_util.c:160-167)."""
    import math
    m = 2654435761 % V
    if m < 2:
        m = 1
    while math.gcd(m, V) != 1:
        m += 1
    return m, 0x9E3779B9 % V


def oracle_paths_mt(V, v, e, eids, src, dst, nthreads, rows_per_thread=512):
    """orc.shortestpath on nthreads host threads, one <=512-row call each (ctypes releases the GIL).
    -> list of (lo, hi, (off, len, valid, vals))."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    orc.lib()
    res = [None] * nthreads

    def work(t):
        lo = t * rows_per_thread
        hi = min(len(src), lo + rows_per_thread)
        res[t] = (lo, hi, orc.shortestpath(V, v, e, eids, src[lo:hi], dst[lo:hi], mode=0))

    ths = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    return res


def bench_cheapest(ctx, csr_id, V, E, src, dst, nw, peak_gbs, with_parity):
    """SURVEY.md section 8(f) f4 on the C4 graph: cheapest_path_length over BIGINT weights 1..100 per knows row (and
    the same values as DOUBLE), host buffers through the C ABI.  Reported: the default path (16-bit distance rows,
    destination bound, 4 lane batches in flight), the same on 8-byte rows, round 1's shape (8-byte rows, no bound,
    one batch at a time), and a byte model from the instrumented kernels."""
    w = np.random.default_rng(4243).integers(1, 101, size=E).astype(np.int64)
    ctx.csr_set_weights(csr_id, w)

    def timed(reps=2):
        ctx.cheapest_path_length(csr_id, src[:nw], dst[:nw])
        ctx.reset_stats()
        t0 = time.perf_counter()
        for _ in range(reps):
            res = ctx.cheapest_path_length(csr_id, src[:nw], dst[:nw])
        secs = (time.perf_counter() - t0) / reps
        st = ctx.stats()
        return res, secs, {k: int(st[k] // reps) for k in ("batches", "levels", "sssp_narrow_batches",
                                                           "sssp_redo_batches", "kernel_launches")}
    res, secs, st = timed()
    line = {"pairs": nw, "pairs_per_s": nw / secs, "ms": secs * 1e3, "batches": st["batches"],
            "rounds": st["levels"], "narrow_batches": st["sssp_narrow_batches"], "redo_batches": st["sssp_redo_batches"],
            "found": int(np.sum(res[1])), "weights": "int64 uniform 1..100",
            "what": "16-bit distance rows relaxed with VIADDMNMX.U16x2, 8 threads per row (exact: non-negative int64 "
                    "weights, a batch whose distances pass 0x3FFF is redone on 8-byte rows), destination bound, hub "
                    "in-lists in 4096-edge chunks, 4 lane batches in flight, control block read every 4 rounds"}
    # byte model from the instrumented instantiation (one extra pass, not timed)
    ctx.set_option("stats", 1)
    ctx.reset_stats()
    ctx.cheapest_path_length(csr_id, src[:nw], dst[:nw])
    ss = ctx.stats()
    ctx.set_option("stats", 0)
    hot, scanned, rounds = int(ss["k_gathered"]), int(ss["k_scanned"]), int(ss["levels"])
    # per source row read: the 128-byte row + its 2-byte weight; per in-edge scanned: the 4-byte source id; per round:
    # the in-degree offsets (4 B x V); per batch: the row fill (128 B x V).  The changed-bitmap tests hit L1/L2.
    model = hot * (128 + 2) + scanned * 4 + rounds * 4 * V + st["batches"] * 128 * V
    l2_peak = 6300 * 1.92  # GB/s: LTS cap of ~6300 B/clk (B300_MICROARCH.md, L2 table) at the 1.92 GHz boost clock
    line["roofline"] = {"bound": "l2", "achieved": model / secs / 1e9, "peak": l2_peak, "unit": "GB/s",
                        "frac": model / secs / 1e9 / l2_peak, "traffic": None,
                        "hbm_peak": peak_gbs, "frac_of_hbm_peak": model / secs / 1e9 / peak_gbs if peak_gbs else None,
                        "model_bytes_per_call": model, "source_rows_read": hot, "in_edges_scanned": scanned,
                        "note": "whole-call wall time (4 streams overlap, so no per-kernel events); the 16-bit rows of "
                                "this graph (V x 128 B = %.0f MB per batch in flight) largely sit in the 126 MB L2 "
                                "(ncu: 86 %% L2 hit rate, 0.4 GB of DRAM reads against 5 GB of row reads in a dense "
                                "round), so the bound is the L2-to-SM path, not HBM" % (V * 128 / 1e6)}
    try:
        ctx.set_option("sssp_narrow", 0)
        r2, secs2, st2 = timed(1)
        line["rows_8_byte"] = {"pairs_per_s": nw / secs2, "ms": secs2 * 1e3, "rounds": st2["levels"],
                               "same_answers": bool(np.array_equal(r2[0], res[0]) and np.array_equal(r2[1], res[1]))}
        ctx.set_option("sssp_narrow", 2)
        r4, secs4, st4 = timed(1)
        line["rows_32_bit"] = {"pairs_per_s": nw / secs4, "ms": secs4 * 1e3, "rounds": st4["levels"],
                               "same_answers": bool(np.array_equal(r4[0], res[0]) and np.array_equal(r4[1], res[1])),
                               "what": "what a batch costs when its distances do not fit 14 bits (VIADDMNMX.U32)"}
        ctx.set_option("sssp_narrow", 0)
        ctx.set_option("sssp_prune", 0)
        ctx.set_option("sssp_slots", 1)
        r3, secs3, st3 = timed(1)
        line["round1_shape"] = {"pairs_per_s": nw / secs3, "ms": secs3 * 1e3, "rounds": st3["levels"],
                                "same_answers": bool(np.array_equal(r3[0], res[0]) and np.array_equal(r3[1], res[1])),
                                "what": "8-byte rows, no destination bound, one lane batch at a time (round 1: 1830 "
                                        "pairs/s with a host sync per round)"}
    finally:
        ctx.set_option("sssp_narrow", 1)
        ctx.set_option("sssp_prune", 1)
        ctx.set_option("sssp_slots", 4)
    if with_parity:
        import orc
        k = min(8, nw)
        v_off, tgt, _ = ctx.csr_download(csr_id)
        wc = ctx.csr_download_weights(csr_id)
        t0 = time.perf_counter()
        eo, ev, _ = orc.cheapest_path_length(V, v_off, tgt, wc, src[:k], dst[:k])
        line["parity"] = {"pairs": k, "vs_oracle": "identical" if (np.array_equal(eo, res[0][:k]) and
                                                                 np.array_equal(ev, res[1][:k])) else "MISMATCH",
                          "oracle_s": time.perf_counter() - t0}
    ctx.csr_set_weights(csr_id, w.astype(np.float64))
    rd, secsd, std_ = timed(1)
    line["double_weights"] = {"pairs_per_s": nw / secsd, "ms": secsd * 1e3, "rounds": std_["levels"],
                              "same_values_as_int64": bool(np.array_equal(rd[0], res[0].astype(np.float64)) and
                                                           np.array_equal(rd[1], res[1])),
                              "what": "the same weights as DOUBLE: whole numbers, so the integer rows answer them with "
                                      "the same bits (sums of whole-numbered doubles below 2^53 are exact)"}
    wf = w.astype(np.float64) + 0.25
    ctx.csr_set_weights(csr_id, wf)
    rf, secsf, stf = timed(1)
    line["double_weights_fractional"] = {"pairs_per_s": nw / secsf, "ms": secsf * 1e3, "rounds": stf["levels"],
                                         "what": "weights + 0.25: 8-byte double rows (bound, chunks, sparse rounds, 4 in "
                                                 "flight as for the integer rows)"}
    if with_parity:
        import orc
        k = min(8, nw)
        v_off, tgt, _ = ctx.csr_download(csr_id)
        wcf = ctx.csr_download_weights(csr_id)
        eo, ev, _ = orc.cheapest_path_length(V, v_off, tgt, wcf, src[:k], dst[:k])
        line["double_weights_fractional"]["parity"] = {
            "pairs": k, "vs_oracle": "identical" if (eo.tobytes() == rf[0][:k].tobytes() and
                                                     np.array_equal(ev, rf[1][:k])) else "MISMATCH"}
    return line


def bench_paths(ctx, V, E, npairs, steps, warmup, peak_gbs, with_parity=True):
    """BASELINE config 4: shortestpath with path reconstruction on an LDBC-SF100-shaped knows graph
    (V persons, E directed knows rows = both directions), npairs (src,dst) pairs, host buffers in and
    the LIST<BIGINT> child buffer out (the scalar function's own data flow)."""
    import pgq_b200
    mult, add = knows_params(V)
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_knows_device(SEED_BASE + 100, V, E, mult, add, d_src, d_dst)
    ctx.csr_build_device(200, V, E, d_src, d_dst, rowids=True)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    ctx.csr_prepare(200)
    v_off = np.empty(V + 2, dtype=np.int64)
    ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 200, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
    src, dst = make_pairs(V, v_off, npairs, 4242)
    out = {}
    full = {}
    calls = (("shortestpath", lambda a, b: ctx.shortestpath(200, a, b, mode=0)),
             ("iterativelength", lambda a, b: ctx.iterativelength(200, a, b)))
    for name, fn in calls:
        for _ in range(min(warmup, 2)):
            res = fn(src, dst)
        ctx.reset_stats()
        steps = max(steps, 5)  # ~20 ms calls: time each one, report the median and every sample
        samples = []
        for _ in range(steps):
            t0 = time.perf_counter()
            res = fn(src, dst)
            samples.append(time.perf_counter() - t0)
        secs = float(np.median(samples))
        st = ctx.stats()
        full[name] = res
        out[name] = {"pairs_per_s": npairs / secs, "ms": secs * 1e3,
                     "ms_samples": [round(x * 1e3, 2) for x in samples], "batches": int(st["batches"] // steps),
                     "levels": int(st["levels"] // steps), "host_syncs": int(st["host_syncs"] // steps),
                     "h2d_bytes_per_step": int(st["h2d_bytes"] // steps),
                     "d2h_bytes_per_step": int(st["d2h_bytes"] // steps)}
        if name == "shortestpath":
            off, ln, valid, vals = res
            out[name]["found"] = int(np.sum(valid))
            out[name]["path_values"] = int(len(vals))
        # the plugin's granularity: the same pair set through <=1024-row calls
        # (/root/reference/src/include/duckdb/common/vector_size.hpp:16-18)
        fn(src[:1024], dst[:1024])
        t0 = time.perf_counter()
        for lo in range(0, npairs, 1024):
            fn(src[lo:lo + 1024], dst[lo:lo + 1024])
        secs = time.perf_counter() - t0
        out[name]["e2e_chunked"] = {"pairs_per_s": npairs / secs, "ms": secs * 1e3, "rows_per_call": 1024,
                                    "calls": (npairs + 1023) // 1024}
    # roofline of the lengths pass on this graph: bytes this implementation touched (instrumented, untimed)
    try:
        ctx.set_option("stats", 1)
        ctx.reset_stats()
        ctx.iterativelength(200, src, dst)
        st = ctx.stats()
        ms = out["iterativelength"]["ms"]
        out["iterativelength"]["roofline"] = {
            "bound": "hbm", "achieved": st["model_bytes"] / (ms * 1e-3) / 1e9, "peak": peak_gbs, "unit": "GB/s",
            "frac": st["model_bytes"] / (ms * 1e-3) / 1e9 / peak_gbs, "traffic": None,
            "model_bytes": int(st["model_bytes"]),
            "note": "86 MB of bitmaps per batch live in the 126 MB L2: the bytes are touched, mostly not in DRAM; "
                    "the step is the whole call incl. H2D/D2H and host lane filling"}
    except Exception as ex:
        out["iterativelength"]["roofline"] = {"error": str(ex)[:200]}
    finally:
        ctx.set_option("stats", 0)
    # ---- parity at config size (untimed): oracle CSR from host-generated rows; oracle paths, byte for byte ----
    if with_parity:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        hs, hd = orc.gen_knows(SEED_BASE + 100, V, E, nthreads=min(host_cores(), 32))
        ov, oe, oeid = orc.csr_build(V, hs, hd)
        del hs, hd
        par = {"csr": csr_parity(ctx, 200, V, E, ov, oe)}
        try:
            avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
        except Exception:
            avail = 32 << 30
        per = V * 512 * 8 * 2 + (256 << 20)  # the oracle's parent arrays
        T = max(1, min(host_cores(), 32, (npairs + 511) // 512, int(avail * 0.6 // per)))
        t0 = time.perf_counter()
        chunks = oracle_paths_mt(V, ov, oe, oeid, src, dst, T)
        cpu_secs = time.perf_counter() - t0
        off, ln, valid, vals = full["shortestpath"]
        glen, gval = full["iterativelength"]
        same = True
        rows = 0
        for lo, hi, (ooff, oln, ovalid, ovals) in chunks:
            rows += hi - lo
            base = int(off[lo])
            end = int(off[hi - 1] + ln[hi - 1])
            same &= bool(np.array_equal(valid[lo:hi], ovalid) and np.array_equal(ln[lo:hi], oln) and
                         np.array_equal(off[lo:hi] - np.uint64(base), ooff) and np.array_equal(vals[base:end], ovals))
            # the lengths call must agree with the paths' lengths too
            same &= bool(np.array_equal(gval[lo:hi], ovalid) and
                         np.array_equal(np.where(ovalid == 1, glen[lo:hi] + 1, 0), oln.astype(np.int64)))
        par["paths"] = "identical" if same else "MISMATCH"
        par["paths_what"] = ("%d of %d pairs (offsets, lengths, validity and values byte for byte against the oracle's "
                             "parent-array shortestpath on %d host threads, %.1f s)" % (rows, npairs, T, cpu_secs))
        out["parity"] = par
        out["cpu_oracle"] = {"shortestpath_pairs_per_s": rows / cpu_secs, "cores": T, "kind": "port"}
    # next row f4 on the same graph: cheapest_path_length, BIGINT weights 1..100 per knows row
    try:
        out["cheapest_path_length"] = bench_cheapest(ctx, 200, V, E, src, dst, min(npairs, 1024), peak_gbs, with_parity)
    except Exception as ex:  # reported, never fatal for the headline
        out["cheapest_path_length"] = {"error": str(ex)[:200]}
    ctx.csr_delete(200)
    return {"workload": "BASELINE configs[3]: LDBC-SF100-shaped knows graph V=%d E=%d, %d pairs, host buffers "
                        "(e2e through the C ABI)" % (V, E, npairs), **out}


def bench_sql():
    """The same functions through the reference engine: LOAD of the extension into the DuckDB built from the reference
    (baseline/_ref/duckdb travels with the snapshot), C1-shaped tables, hand-written statements and one SQL/PGQ
    GRAPH_TABLE statement through the rewriter (tools/sql_bench.py; wall time of Connection::Query per statement, median
    of repetitions 1-3 of 4; checksums against the C ABI).  Reported, never fatal."""
    import subprocess
    need = [os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_b200.duckdb_extension"),
            os.path.join(ROOT, "duckdb-pgq.old_b200", "ext", "pgq_sql"),
            os.path.join(ROOT, "baseline", "_ref", "duckdb", "libduckdb.so")]
    if not all(os.path.exists(f) for f in need):
        return {"unavailable": "reference DuckDB host / extension not built in this snapshot"}
    try:
        p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sql_bench.py"), "--reps", "4"],
                           capture_output=True, text=True, timeout=600)
        d = json.loads(p.stdout.strip().splitlines()[-1])
        samples = {}  # statement -> times of repetitions 1.. (repetition 0 pays module loading and first allocations)
        for k, v in d["statements_ms"].items():
            if "_" in k and k.rsplit("_", 1)[1].isdigit() and int(k.rsplit("_", 1)[1]) >= 1 and v is not None:
                samples.setdefault(k.rsplit("_", 1)[0], []).append(v)
        ms = {k: round(float(np.median(v)), 3) for k, v in samples.items()}
        n = d["pairs"]
        return {"workload": "reference DuckDB + pgq_b200.duckdb_extension, %d persons, %d knows rows, %d pairs, PRAGMA "
                            "threads=%d" % (d["persons"], d["knows"], n, d["threads"]),
                "statement_ms": ms,
                "iterativelength_pairs_per_s": 2 * n / (ms["iterativelength"] * 1e-3) if ms.get("iterativelength") else None,
                "graph_table": dict(d["graph_table"], ms=ms.get("graph_table"),
                                    what="CREATE PROPERTY GRAPH + SELECT ... FROM GRAPH_TABLE (... ANY SHORTEST ...) through "
                                         "ext/pgq_rewrite.hpp: CSR build CTE, 2 x pairs path lengths, delete_csr"),
                "parity_vs_c_abi": d["parity_vs_c_abi"], "errors": d["errors"]}
    except Exception as ex:
        return {"error": str(ex)[:300]}


def bench_c1(ctx, steps, warmup, with_cpu, peak_gbs):
    """BASELINE configs[0]: LDBC SNB SF1-shaped Person-knows-Person graph (10 k vertices, 180 k directed
    edges): CSR build + iterativelength for 2048 pairs, everything from HOST buffers through the same
    calls the DuckDB functions make.  This is the one config the CPU side can run in full, so the
    oracle is timed on exactly the same work (rank 0, N=1 only)."""
    import pgq_b200
    V, E, npairs = 10000, 180000, 2048
    mult, add = knows_params(V)
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_knows_device(SEED_BASE + 1, V, E, mult, add, d_src, d_dst)
    src = np.empty(E, dtype=np.int64)
    dst = np.empty(E, dtype=np.int64)
    ctx.d2h(src, d_src)
    ctx.d2h(dst, d_dst)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    deg = np.bincount(src, minlength=V).astype(np.int64)
    vid = np.arange(V, dtype=np.int64)
    v_off = np.concatenate([[0], np.cumsum(deg), [E]]).astype(np.int64)
    ps, pd = make_pairs(V, v_off, npairs, 77)

    def gpu_once():
        ctx.create_csr_vertex(300, V, vid, deg)
        for lo in range(0, E, 1024):  # the engine hands the function 1024-row chunks
            ctx.create_csr_edge(300, V, E, src[lo:lo + 1024], dst[lo:lo + 1024])
        ctx.csr_finalize(300)
        t1 = time.perf_counter()
        outs = []
        for lo in range(0, npairs, 1024):
            outs.append(ctx.iterativelength(300, ps[lo:lo + 1024], pd[lo:lo + 1024]))
        t2 = time.perf_counter()
        ctx.csr_delete(300)
        return t1, t2, outs

    for _ in range(max(1, warmup)):
        gpu_once()
    tb, tq = [], []
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        t1, t2, outs = gpu_once()
        tb.append(t1 - t0)
        tq.append(t2 - t1)
    res = {"workload": "BASELINE configs[0]: LDBC-SF1-shaped knows graph V=%d E=%d, CSR build from 1024-row "
                       "chunks + iterativelength for %d pairs in 1024-row calls, host buffers" % (V, E, npairs),
           "gpu": {"csr_build_ms": float(np.mean(tb)) * 1e3, "iterativelength_ms": float(np.mean(tq)) * 1e3,
                   "pairs_per_s": npairs / float(np.mean(tq))}}
    # roofline of the whole-batch kernel: every level sweeps every vertex (seen read, next written) and gathers
    # one frontier word per in-edge of the still-open vertices; upper bound = all in-edges
    try:
        ctx.create_csr_vertex(300, V, vid, deg)
        ctx.create_csr_edge(300, V, E, src, dst)
        ctx.csr_finalize(300)
        ctx.reset_stats()
        for lo in range(0, npairs, 1024):
            ctx.iterativelength(300, ps[lo:lo + 1024], pd[lo:lo + 1024])
        st = ctx.stats()
        ctx.csr_delete(300)
        lv = float(st["levels"])
        algo = lv * (2.0 * V * 64 + E * 68.0)
        ms = float(np.mean(tq)) * 1e3
        res["gpu"]["roofline"] = {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak_gbs, "unit": "GB/s",
                                  "frac": algo / (ms * 1e-3) / 1e9 / peak_gbs, "traffic": None,
                                  "algo_bytes": int(algo), "levels": int(lv),
                                  "note": "the graph's bitmaps (1.9 MB) and edges live in L2: the kernel is bound by "
                                          "its per-level grid barrier (4.4 us) and the call's launch + copies, not "
                                          "by HBM; algo bytes = levels * (2*V*64 + E*68), an upper bound"}
    except Exception as ex:
        res["gpu"]["roofline"] = {"error": str(ex)[:200]}
    if with_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        t0 = time.perf_counter()
        v, e, _ = orc.csr_build(V, src, dst, want_eids=False)
        t1 = time.perf_counter()
        couts = []
        for lo in range(0, npairs, 1024):
            couts.append(orc.iterativelength(V, v, e, ps[lo:lo + 1024], pd[lo:lo + 1024]))
        t2 = time.perf_counter()
        same = all(np.array_equal(a[1], b[1]) and np.array_equal(np.where(b[1] == 1, a[0], -1),
                                                                 np.where(b[1] == 1, b[0], -1))
                   for a, b in zip(outs, couts))
        res["cpu_oracle_1thread"] = {"csr_build_ms": (t1 - t0) * 1e3, "iterativelength_ms": (t2 - t1) * 1e3,
                                     "pairs_per_s": npairs / (t2 - t1), "kind": "port",
                                     "parity": "identical" if same else "MISMATCH"}
    return res


W_BITS = 512  # lanes per vertex word on the timed path


def time_steps(ctx, dist, local, fn, steps):
    """K calls of fn bracketed by a barrier; CUDA-event time on the library's stream, max over ranks (ms)."""
    ctx.sync()
    barrier_max(dist, local, 0.0)
    ctx.timer_start()
    for _ in range(steps):
        fn()
    return barrier_max(dist, local, ctx.timer_stop_ms())


def main_ours(a):
    import pgq_b200
    rank, world, local, dist = dist_setup(a.gpus)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    ctx = pgq_b200.Context(local)
    scale, ef = a.scale, a.edgefactor
    V, E = 1 << scale, ef << scale
    cpu_legs = rank == 0 and world == 1 and not a.no_cpu_baseline
    # ---- setup (untimed): synthetic graph in HBM, CSR + reverse CSR resident ----
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
    ctx.csr_build_device(0, V, E, d_src, d_dst)
    ctx.dev_free(d_src)
    ctx.dev_free(d_dst)
    build_ms = ctx.stats()["csr_build_ms"]
    ctx.csr_prepare(0)
    transpose_ms = ctx.stats()["transpose_ms"]
    v_off = np.empty(V + 2, dtype=np.int64)
    ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 0, v_off.ctypes.data_as(pgq_b200.I64P), None, None))
    # ---- config-size parity of THIS graph's CSR (untimed, rank 0 at N=1): the oracle builds it on the host from
    # host-generated rows; the CPU legs below run on the ORACLE's arrays, never on a download of the GPU's ----
    csr3_parity = None
    o_v = o_e = None
    if cpu_legs:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        hs, hd = orc.gen_rmat(SEED_BASE + scale, scale, E=E, nthreads=min(host_cores(), 64))
        o_v, o_e, _ = orc.csr_build(V, hs, hd, want_eids=False)
        del hs, hd
        csr3_parity = csr_parity(ctx, 0, V, E, o_v, o_e)
    n = a.pairs
    src, dst = make_pairs(V, v_off, n, 1000 + rank)
    d_ps, d_pd = ctx.dev_alloc(n * 8), ctx.dev_alloc(n * 8)
    d_out, d_ov = ctx.dev_alloc(n * 8), ctx.dev_alloc(n)
    ctx.h2d(d_ps, src)
    ctx.h2d(d_pd, dst)
    # ---- instrumented pass (untimed): TEPS numerator, the reference algorithm's bytes and the bytes
    # THIS implementation touched (DESIGN.md section 4 model) ----
    ctx.set_option("stats", 1)
    ctx.reset_stats()
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    st = ctx.stats()
    ctx.set_option("stats", 0)
    edges_lane, algo_bytes, levels, batches = st["edges_lane"], st["algo_bytes"], st["levels"], st["batches"]
    model_bytes = st["model_bytes"]
    model_parts = {k: int(st[k]) for k in ("k_listed", "k_open", "k_new", "k_scanned", "k_gathered", "kp_items",
                                           "kp_edges", "kp_seen", "kp_next")}
    ref_out = np.empty(n, dtype=np.int64)
    ctx.d2h(ref_out, d_out)
    # ---- level kernels timed alone (untimed pass): one batch at a time on one stream, host-driven loop, CUDA
    # events around each level's kernels; gives the kernels' share of a step ----
    ctx.set_option("timing", 1)
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    ctx.reset_stats()
    ctx.timer_start()
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    step_alone_ms = ctx.timer_stop_ms()
    st = ctx.stats()
    level_ms, level_launches = st["level_kernel_ms"], int(st["level_kernel_launches"])
    ctx.set_option("timing", 0)
    # (device-side level loop: one host thread enqueues every batch, so the host's core count no longer matters;
    # the host-driven loop pins a spin-waiting thread per batch in flight)
    cores_per_rank = host_cores() // max(1, world)
    if a.loop:
        streams = a.streams if a.streams else 8
    else:
        streams = a.streams if a.streams else (8 if cores_per_rank >= 16 else max(1, min(4, cores_per_rank)))
    ctx.set_option("loop", a.loop)
    ctx.set_option("streams", streams)
    # ---- device-resident timed region ----
    for _ in range(a.warmup):
        ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    ctx.sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ctx.reset_stats()
    local_t0 = time.perf_counter()
    ms_total = time_steps(ctx, dist, local, lambda: ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov), a.steps)
    st = ctx.stats()
    launches = int(st["kernel_launches"])
    host_syncs = int(st["host_syncs"])
    got = np.empty(n, dtype=np.int64)
    ctx.d2h(got, d_out)
    assert np.array_equal(got, ref_out), "timed run disagrees with the instrumented run"
    # ---- e2e: host buffers through the reference-facing C-ABI call ----
    out_h, ov_h = ctx.iterativelength(0, src, dst)  # warm
    assert np.array_equal(np.where(ov_h == 1, out_h, -1), ref_out), "host and device variants disagree"
    ctx.reset_stats()
    e2e_ms = time_steps(ctx, dist, local, lambda: ctx.iterativelength(0, src, dst), a.steps)
    st2 = ctx.stats()
    # ---- the plugin's granularity: the same pair set through <= 1024-row calls (vector_size.hpp:16-18) ----
    def chunked():
        for lo in range(0, n, 1024):
            ctx.iterativelength(0, src[lo:lo + 1024], dst[lo:lo + 1024])
    chunked_steps = max(1, min(a.steps, 2))
    chunked_ms = time_steps(ctx, dist, local, chunked, chunked_steps) / chunked_steps
    clocks = sampler.stop() if rank == 0 else None
    del local_t0
    total_edges = barrier_sum(dist, local, float(edges_lane))
    total_pairs = barrier_sum(dist, local, float(n))
    ms_per_step = ms_total / a.steps
    value = total_edges / (ms_per_step * 1e-3) / 1e9
    e2e_value = total_edges / (e2e_ms / a.steps * 1e-3) / 1e9
    # ---- roofline of the dominant kernels (the per-level pull/push kernels) ----
    # frac     = bytes THIS implementation must move for the levels it ran (what its kernels touched, counted by the
    #            instrumented pass; a gathered word counts 64 B wherever it was served from) / time / peak.
    # reference_model = the REFERENCE algorithm's bytes per level: what the CPU loop would move;
    #            the frontier filter, the early exit and the done bitmap skip most of it, so that ratio exceeds 1.
    model_achieved = model_bytes / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else 0.0
    ref_achieved = algo_bytes / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else 0.0
    prof, traffic_src = load_traffic("r02_traffic.json", {"scale": scale, "edgefactor": ef, "lanes": W_BITS})
    traffic = dram_frac = None
    if prof and ms_per_step > 0:
        traffic = float(prof["level_kernel_dram_bytes_per_batch"]) * float(batches)
        dram_frac = traffic / (ms_per_step * 1e-3) / 1e9 / peak_gbs
    line = {
        "metric": "msbfs_gteps", "value": value, "unit": "GTEPS", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 bitset (512 lanes/vertex), u32 ids", "data": "synthetic",
        "config": {"workload": "BASELINE configs[2]: MS-BFS iterativelength, R-MAT scale %d ef%d, %d src/dst "
                               "pairs per GPU in 512-lane batches, CSR replicated" % (scale, ef, n),
                   "V": V, "E": E, "pairs_per_gpu": n, "batches": int(batches), "levels": int(levels),
                   "l2": "inputs larger than L2 (3 bitmaps of V*64 B = %.1f GB)" % (3 * V * 64 / 1e9),
                   "parallelism": "source-batch sharding x%d" % world,
                   "batches_in_flight": int(streams),
                   "level_loop": "device (CUDA graph WHILE/SWITCH)" if a.loop else "host"},
        "pairs_per_s": total_pairs / (ms_per_step * 1e-3),
        "e2e": {"value": e2e_value, "unit": "GTEPS", "pairs_per_s": total_pairs / (e2e_ms / a.steps * 1e-3),
                "h2d_bytes_per_step": int(st2["h2d_bytes"] // a.steps),
                "d2h_bytes_per_step": int(st2["d2h_bytes"] // a.steps)},
        "e2e_chunked": {"value": total_edges / (chunked_ms * 1e-3) / 1e9, "unit": "GTEPS",
                        "pairs_per_s": total_pairs / (chunked_ms * 1e-3), "rows_per_call": 1024,
                        "calls_per_step": (n + 1023) // 1024,
                        "what": "the same pair set through <= 1024-row pgq_iterativelength calls, one after the other "
                                "(the plugin's call granularity with one worker thread); the DuckDB shim's batcher "
                                "merges the calls of concurrent workers, see sql"},
        "gpu_launches": launches,
        "host_syncs_in_level_loops": host_syncs,
        "roofline": {"bound": "hbm", "achieved": model_achieved, "peak": peak_gbs, "unit": "GB/s",
                     "frac": model_achieved / peak_gbs, "traffic": traffic, "dram_frac": dram_frac,
                     "traffic_source": traffic_src, "peak_source": peak_src,
                     "kernel": "MS-BFS level step = pull_tiny/pull_list/pull_seg/pull_warp/pull_hub (+push, bitmap) kernels",
                     "model": "bytes this implementation touches (DESIGN.md section 4): per listed vertex 12 B (list "
                              "entry, two in-offsets), per open vertex its 64-B seen word, per vertex that gained "
                              "bits 128 B (next + seen stores), 4 B per scanned in-edge, 64 B per gathered frontier "
                              "word; push: 88 B per work item, 4 B per out-edge, 8 B per 32-lane atomic; per level "
                              "3 bitmap passes; per batch the seen memset and the frontier scrub",
                     "model_bytes_per_step": int(model_bytes), "model_parts": model_parts,
                     "reference_model": {"achieved": ref_achieved, "frac": ref_achieved / peak_gbs,
                                         "algo_bytes_per_step": int(algo_bytes),
                                         "what": "SURVEY.md 8(d): 4*V*W/8 per level + E_act*(4+W/8) + 8*V_act, the bytes "
                                                 "of the reference's loop; > 1 means work the kernels skip"},
                     "launches": level_launches,
                     "level_kernel_ms_per_step_alone": level_ms,
                     "alone": "level kernels of one step timed with one batch in flight (host loop, CUDA events around "
                              "each level): %.0f GB/s of model bytes" % (model_bytes / (level_ms * 1e-3) / 1e9
                                                                        if level_ms > 0 else 0.0),
                     "step_ms_alone": step_alone_ms,
                     "share_of_step": level_ms / step_alone_ms if step_alone_ms else None},
        "setup": {"csr_build_ms": build_ms, "reverse_csr_ms": transpose_ms,
                  "csr_build_GEps": E / (build_ms * 1e-3) / 1e9 if build_ms else None},
        "parity": {"csr_config3": csr3_parity},
        "clocks": clocks,
    }
    if rank == 0 and W_BITS == 512:
        # what the memory system gives for the dense level's access pattern (random 64-byte word gathers, no
        # arithmetic): beyond L2 it is a fraction of the sequential-copy peak, whatever the kernel does
        try:
            g_l2, _ = ctx.probe_gather(int(0.245e9) // 64, 1 << 26, 4)
            g_far, _ = ctx.probe_gather((1 << 30) // 64, 1 << 26, 4)
            line["roofline"]["random_gather_gbs"] = {
                "table_245MB_(one_dense_frontier)": g_l2, "table_1GB_(whole_visit_array)": g_far,
                "what": "measured live: random 64-byte gathers, 4 threads x 16 B per word, 4 in flight per "
                        "thread, indices streamed; GB/s counts 68 B per gather"}
        except Exception as ex:  # a probe must not cost the bench line
            line["roofline"]["random_gather_gbs"] = {"error": str(ex)[:200]}
    # ---- multi-GPU sections (N > 1): strong scaling of the stated pair set, ONE process driving all GPUs through a
    # single C-ABI call, and BASELINE configs[4] (1-D vertex partition) ----
    if world > 1:
        multi_gpu_sections(a, ctx, dist, rank, world, local, line, V, E, v_off, ms_per_step, edges_lane, streams)
    if rank == 0 and world == 1 and not a.no_csr:
        line["csr_build"] = bench_csr(ctx, a.csr_scale, ef, a.steps, a.warmup, peak_gbs, not a.no_cpu_baseline)
        line["parity"]["csr_config2"] = line["csr_build"]["parity"]
    if rank == 0 and world == 1 and not a.no_paths:
        line["config1"] = bench_c1(ctx, a.steps, a.warmup, not a.no_cpu_baseline, peak_gbs)
        line["parity"]["config1"] = line["config1"].get("cpu_oracle_1thread", {}).get("parity")
    if rank == 0 and world == 1 and not a.no_paths:
        line["paths"] = bench_paths(ctx, 450000, 40000000, 16384, max(1, a.steps // 2), a.warmup, peak_gbs,
                                    not a.no_cpu_baseline)
        line["parity"]["paths_config4"] = line["paths"].get("parity", {}).get("paths")
        line["parity"]["csr_config4"] = line["paths"].get("parity", {}).get("csr")
        line["sql"] = bench_sql()
    if cpu_legs:
        T, cores = cpu_sample_threads(V, a.cpu_threads)
        rows_pt = min(512, n)
        T = max(1, min(T, n // rows_pt))
        # the ORACLE's own CSR (host-built from host-generated rows), not a download of the GPU's
        secs, rows, edges, res = run_cpu_sample(V, o_v, o_e, src, dst, T, rows_pt)
        same = True
        for t in range(T):
            lo = t * rows_pt
            eo, ev, _ = res[t]
            same &= bool(np.array_equal(np.where(ev == 1, eo, -1), ref_out[lo:lo + rows_pt]))
        line["parity"]["lengths_config3"] = "identical" if same else "MISMATCH"
        line["parity"]["lengths_config3_what"] = "%d of %d pairs against the oracle on the oracle's own CSR" % (rows, n)
        line["cpu_baseline"] = {"value": edges / secs / 1e9, "unit": "GTEPS", "cores": T, "host_cores": cores,
                                "kind": "port", "pairs_per_s": rows / secs,
                                "sample": "%d pairs (%d threads x one %d-row call) of the same pair set, %.1f s"
                                          % (rows, T, rows_pt, secs)}
    if rank == 0:
        bad = [k for k, v in line["parity"].items() if v == "MISMATCH"]
        if bad:
            line["parity"]["FAILED"] = bad
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def multi_gpu_sections(a, ctx, dist, rank, world, local, line, V, E, v_off, ms_weak, edges_lane_rank, streams):
    """N > 1 only.  (1) strong scaling: the stated 65 536-pair set split N ways (shard_rows), no collective;
    (2) ONE process, ONE pgq_iterativelength call, all N GPUs (pgq_init_devices; rank 0, the other ranks idle);
    (3) BASELINE configs[4]: 1-D vertex partition + per-level exchange, both exchange modes, with parity against the
    single-GPU path at the bench graph's scale."""
    import pgq_b200
    n = a.pairs
    steps = max(1, a.steps)
    src0, dst0 = make_pairs(V, v_off, n, 1000)  # rank 0's set, identical on every rank
    set_edges = barrier_sum(dist, local, float(edges_lane_rank) if rank == 0 else 0.0)
    # ---- (1) strong scaling ----
    lo, hi = pgq_b200.shard_rows(n, world, rank)
    m = hi - lo
    d = [ctx.dev_alloc(max(m, 1) * 8) for _ in range(3)] + [ctx.dev_alloc(max(m, 1))]
    ctx.h2d(d[0], src0[lo:hi])
    ctx.h2d(d[1], dst0[lo:hi])
    run = (lambda: ctx.iterativelength_device(0, m, d[0], d[1], d[2], d[3])) if m else (lambda: None)
    for _ in range(2):
        run()
    strong_ms = time_steps(ctx, dist, local, run, steps) / steps
    for p in d:
        ctx.dev_free(p)
    line["strong"] = {"pairs_total": n, "pairs_per_gpu": [pgq_b200.shard_rows(n, world, r)[1] -
                                                          pgq_b200.shard_rows(n, world, r)[0] for r in range(world)],
                      "ms": strong_ms, "value": set_edges / (strong_ms * 1e-3) / 1e9, "unit": "GTEPS",
                      "pairs_per_s": n / (strong_ms * 1e-3), "one_gpu_ms": ms_weak,
                      "speedup_vs_1": ms_weak / strong_ms if strong_ms > 0 else None,
                      "what": "BASELINE configs[2] as stated: ONE 65 536-pair set sharded by source batch over the "
                              "GPUs (device-resident, max over ranks); one_gpu_ms = the weak-scaling step, where "
                              "every GPU answers a whole 65 536-pair set"}
    # ---- (2) one process, all GPUs, one call ----
    # The other ranks wait on a HOST barrier (gloo): an NCCL barrier would park a spinning kernel of another process
    # on every GPU rank 0 is about to use, and the driver would time-slice rank 0's kernels against it (measured:
    # 1.44x on 2 GPUs with the NCCL wait, gpurun check 15).
    barrier_max(dist, local, 0.0)
    import torch
    torch.cuda.synchronize()
    host_group = dist.new_group(backend="gloo")
    dist.barrier(group=host_group)
    if rank == 0:
        try:
            mc = pgq_b200.Context(devices=list(range(world)))
            mc.set_option("streams", streams)
            scale = a.scale
            ds, dd = mc.dev_alloc(E * 8), mc.dev_alloc(E * 8)
            mc.gen_rmat_device(SEED_BASE + scale, scale, E, ds, dd)
            mc.csr_build_device(0, V, E, ds, dd)
            mc.dev_free(ds)
            mc.dev_free(dd)
            out1, ov1 = mc.iterativelength(0, src0, dst0)  # warm: replicates the CSR over NVLink
            ref, rv = ctx.iterativelength(0, src0, dst0)
            same = bool(np.array_equal(ov1, rv) and np.array_equal(out1, ref))
            mc.iterativelength(0, src0, dst0)
            t0 = time.perf_counter()
            for _ in range(steps):
                mc.iterativelength(0, src0, dst0)
            ms = (time.perf_counter() - t0) / steps * 1e3
            t0 = time.perf_counter()
            for _ in range(steps):
                ctx.iterativelength(0, src0, dst0)
            ms1 = (time.perf_counter() - t0) / steps * 1e3
            line["one_process"] = {"devices": world, "ms": ms, "value": set_edges / (ms * 1e-3) / 1e9, "unit": "GTEPS",
                                   "pairs_per_s": n / (ms * 1e-3), "one_device_ms": ms1,
                                   "speedup_vs_1_device": ms1 / ms if ms > 0 else None,
                                   "parity_vs_one_device": "identical" if same else "MISMATCH",
                                   "what": "pgq_init_devices: ONE process, ONE pgq_iterativelength call with host "
                                           "buffers, batches dealt over all GPUs (CSR replicated by peer copies at "
                                           "the first call); wall clock incl. H2D / D2H; the other ranks idle"}
            mc.close()
        except Exception as ex:
            line["one_process"] = {"error": str(ex)[:300]}
    dist.barrier(group=host_group)
    barrier_max(dist, local, 0.0)
    # ---- (3) configs[4]: 1-D vertex partition ----
    try:
        line_part = part_section(a, ctx, dist, rank, world, local, V, v_off, src0, dst0)
        if rank == 0:
            line["part"] = line_part
            line["parity"]["part_vs_single_gpu"] = line_part.get("parity_vs_single_gpu")
    except Exception as ex:
        if rank == 0:
            line["part"] = {"error": str(ex)[:300]}
    barrier_max(dist, local, 0.0)


def part_build(pc, scale, ef, csr_id, chunk=1 << 26):
    import pgq_b200
    C = pgq_b200.C
    V, E = 1 << scale, ef << scale
    pc.part_csr_create(csr_id, V)
    chunk = min(chunk, E)
    d_src, d_dst = pc.dev_alloc(chunk * 8), pc.dev_alloc(chunk * 8)
    t0 = time.perf_counter()
    for row0 in range(0, E, chunk):
        k = min(chunk, E - row0)
        pc._chk(pc.lib.pgq_gen_rmat_rows_device(pc._ctx, C.c_uint64(SEED_BASE + scale), C.c_int(scale),
                                                C.c_int64(row0), C.c_int64(k), C.c_void_p(d_src), C.c_void_p(d_dst)))
        pc.part_csr_add_edges_device(csr_id, d_src, d_dst, k)
    pc.part_csr_finalize(csr_id)
    pc.sync()
    secs = time.perf_counter() - t0
    pc.dev_free(d_src)
    pc.dev_free(d_dst)
    return V, E, secs


def part_section(a, ctx, dist, rank, world, local, V0, v_off, src0, dst0):
    """BASELINE configs[4]: R-MAT ef16 1-D vertex-partitioned over the N GPUs: scale 28 at N=8 (27 at 4, 26 at 2),
    every rank generates the stream and keeps the rows whose target it owns; 256 pairs in 64-lane words, all ranks
    answer collectively.  Exchange per level: ncclAllGather of the word slices (the stated baseline) and the fused
    peer-store variant (the product)."""
    import pgq_b200
    box = [pgq_b200.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    pc = pgq_b200.Context(local)
    pc.part_init(rank, world, box[0])
    out = {}
    # parity first, at the bench graph's own scale: the partitioned answers against the single-GPU path's
    pV, pE, _ = part_build(pc, a.scale, a.edgefactor, 1)
    k = min(1024, len(src0))
    pc.set_option("lanes", 256)
    po, pv = pc.part_iterativelength(1, src0[:k], dst0[:k])
    if rank == 0:
        so, sv = ctx.iterativelength(0, src0[:k], dst0[:k])
        same = bool(np.array_equal(sv, pv) and np.array_equal(np.where(sv == 1, so, -1), np.where(pv == 1, po, -1)))
        out["parity_vs_single_gpu"] = "identical" if same else "MISMATCH"
        out["parity_what"] = "scale %d, %d pairs, %d ranks, ncclAllGather exchange, 256 lanes" % (a.scale, k, world)
    pc.csr_delete(1)
    scale = {2: 26, 4: 27, 8: 28}.get(world, 25)
    steps = 2
    V, E, build_s = part_build(pc, scale, a.edgefactor, 0)
    out.update({"workload": "BASELINE configs[4]: R-MAT scale %d ef%d (V=%d E=%d) 1-D vertex-partitioned over %d B200"
                            % (scale, a.edgefactor, V, E, world),
                "slice_build_s": build_s,
                "timing": "wall clock of the collective call (host buffers in and out), max over ranks; pull_ms / "
                          "exchange_ms = CUDA-event sums over the levels on rank 0"})

    def measure(ps, pd):
        o, ov = pc.part_iterativelength(0, ps, pd)  # warm
        # instrumented pass: TEPS numerator of this pair set (same definition as the single-GPU path)
        pc.set_option("stats", 1)
        pc.reset_stats()
        pc.part_iterativelength(0, ps, pd)
        edges = barrier_sum(dist, local, float(pc.stats()["edges_lane"]))
        pc.set_option("stats", 0)
        pc.set_option("timing", 1)
        pc.reset_stats()
        pc.sync()
        barrier_max(dist, local, 0.0)
        t0 = time.perf_counter()
        for _ in range(steps):
            pc.part_iterativelength(0, ps, pd)
        ms = barrier_max(dist, local, (time.perf_counter() - t0) / steps * 1e3)
        st = pc.stats()
        pc.set_option("timing", 0)
        n = len(ps)
        return (o, ov), {"pairs": n, "ms_per_step": ms, "pairs_per_s": n / (ms * 1e-3),
                         "value": edges / (ms * 1e-3) / 1e9, "unit": "GTEPS", "edges_lane": edges,
                         "batches": int(st["batches"] // steps), "levels": int(st["levels"] // steps),
                         "pull_ms": st["level_kernel_ms"] / steps, "exchange_ms": st["coll_ms"] / steps,
                         "coll_bytes_per_step_per_rank": int(st["coll_bytes"] // steps)}

    rng = np.random.RandomState(1234)
    ps = rng.randint(0, V, size=2048).astype(np.int64)
    pd = rng.randint(0, V, size=2048).astype(np.int64)
    # round 1's configuration: 256 pairs in 64-lane words, both exchange modes
    pc.set_option("lanes", 64)
    r_ag, out["allgather"] = measure(ps[:256], pd[:256])
    out["allgather"]["lanes"] = 64
    pc.part_p2p_setup(dist, 0, 64, world)
    r_fu, out["fused"] = measure(ps[:256], pd[:256])
    out["fused"]["lanes"] = 64
    same = bool(np.array_equal(r_ag[0], r_fu[0]) and np.array_equal(r_ag[1], r_fu[1]))
    out["fused_equals_allgather"] = "identical" if same else "MISMATCH"
    # 512-lane words (2 x 17 GB of frontier replicas per rank at scale 28): every level's sweep serves 8x the lanes
    try:
        pc.set_option("lanes", 512)
        pc.part_p2p_setup(dist, 0, 512, world)
        r_w, out["fused_512"] = measure(ps, pd)
        out["fused_512"]["lanes"] = 512
        same = bool(np.array_equal(r_w[0][:256], r_fu[0]) and np.array_equal(r_w[1][:256], r_fu[1]))
        out["fused_512"]["first_256_equal_64_lane_run"] = "identical" if same else "MISMATCH"
    except Exception as ex:
        out["fused_512"] = {"error": str(ex)[:300]}
    pc.close()
    return out


def main_reference(a):
    """The CPU restatement timed on the host cores (no GPU work at all)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    scale, ef = a.scale, a.edgefactor
    V, E = 1 << scale, ef << scale
    cores = host_cores()
    src_e, dst_e = orc.gen_rmat(SEED_BASE + scale, scale, E=E, nthreads=min(cores, 64))
    t0 = time.perf_counter()
    v, e, _ = orc.csr_build(V, src_e, dst_e, want_eids=False)
    csr_secs = time.perf_counter() - t0
    del src_e, dst_e
    T, _ = cpu_sample_threads(V, a.cpu_threads)
    rows_pt = min(512, a.pairs)
    T = max(1, min(T, a.pairs // rows_pt))
    src, dst = make_pairs(V, v, a.pairs, 1000)
    # Time-boxed: one pass over the sample is tens of seconds, so the requested K steps would run for a quarter of an
    # hour (round 1: the scaling runs hit their limit four times).  The loop is deterministic CPU work: after the
    # time box the passes timed so far are reported, `steps` says how many, `config.steps_requested` what was asked.
    box_s = float(a.ref_seconds)
    t_start = time.perf_counter()
    times, edges_l, rows_l = [], [], []
    if a.warmup > 0:
        run_cpu_sample(V, v, e, src, dst, T, rows_pt)  # one warm-up pass faults the pages in
    for i in range(max(1, a.steps)):
        secs, rows, edges, _ = run_cpu_sample(V, v, e, src, dst, T, rows_pt)
        times.append(secs)
        edges_l.append(edges)
        rows_l.append(rows)
        if time.perf_counter() - t_start + secs > box_s:
            break
    steps_timed = len(times)
    secs = float(np.mean(times))
    value = float(np.mean(edges_l)) / secs / 1e9
    sample = "%d pairs per step (%d threads x one %d-row call) of the %d-pair set" % (rows_l[0], T, rows_pt, a.pairs)
    line = {"impl": "reference", "metric": "msbfs_gteps", "value": value, "unit": "GTEPS", "n_gpus": a.gpus,
            "steps": steps_timed, "warmup": min(a.warmup, 1), "ms_per_step": secs * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64 bitset (512 lanes/vertex), int64 ids",
            "data": "synthetic",
            "config": {"workload": "BASELINE configs[2]: MS-BFS iterativelength, R-MAT scale %d ef%d; CPU oracle "
                                   "(port of the reference algorithm), bounded sample" % (scale, ef),
                       "V": V, "E": E, "pairs_per_step": rows_l[0], "steps_requested": a.steps,
                       "steps_timed": steps_timed, "time_box_s": box_s,
                       "note": "time-boxed: deterministic CPU passes, %d of %d requested steps timed" % (steps_timed, a.steps)},
            "pairs_per_s": rows_l[0] / secs,
            "cpu_baseline": {"value": value, "unit": "GTEPS", "cores": T, "host_cores": cores, "kind": "port",
                             "sample": sample, "csr_build_GEps_1thread": E / csr_secs / 1e9},
            "e2e": {"value": value, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_ours(args)
