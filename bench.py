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


def bench_csr(ctx, scale, ef, steps, warmup, peak_gbs):
    """BASELINE config 2: CSR build from the edge table, R-MAT scale 22 ef16."""
    import pgq_b200
    V, E = 1 << scale, ef << scale
    d_src, d_dst = ctx.dev_alloc(E * 8), ctx.dev_alloc(E * 8)
    ctx.gen_rmat_device(SEED_BASE + scale, scale, E, d_src, d_dst)
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
    return {"workload": "csr_build rmat scale %d ef%d (V=%d E=%d)" % (scale, ef, V, E),
            "value": E / (ms * 1e-3) / 1e9, "unit": "GE/s", "ms": ms,
            "e2e": {"value": E / (float(np.mean(e2e_ms)) * 1e-3) / 1e9, "unit": "GE/s",
                    "ms": float(np.mean(e2e_ms)), "h2d_bytes_per_step": 16 * E, "d2h_bytes_per_step": 0},
            "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak_gbs, "unit": "GB/s",
                         "frac": algo / (ms * 1e-3) / 1e9 / peak_gbs, "traffic": None,
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


def bench_paths(ctx, V, E, npairs, steps, warmup):
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
    for name, fn in (("shortestpath", lambda: ctx.shortestpath(200, src, dst, mode=0)),
                     ("iterativelength", lambda: ctx.iterativelength(200, src, dst))):
        for _ in range(min(warmup, 2)):
            res = fn()
        ctx.reset_stats()
        t0 = time.perf_counter()
        for _ in range(steps):
            res = fn()
        secs = (time.perf_counter() - t0) / steps
        st = ctx.stats()
        out[name] = {"pairs_per_s": npairs / secs, "ms": secs * 1e3, "batches": int(st["batches"] // steps),
                     "levels": int(st["levels"] // steps),
                     "h2d_bytes_per_step": int(st["h2d_bytes"] // steps),
                     "d2h_bytes_per_step": int(st["d2h_bytes"] // steps)}
        if name == "shortestpath":
            off, ln, valid, vals = res
            out[name]["found"] = int(np.sum(valid))
            out[name]["path_values"] = int(len(vals))
    # next row f4 on the same graph: cheapest_path_length, BIGINT weights 1..100 per knows row
    try:
        w = np.random.default_rng(4243).integers(1, 101, size=E).astype(np.int64)
        ctx.csr_set_weights(200, w)
        nw = min(npairs, 1024)
        ctx.cheapest_path_length(200, src[:nw], dst[:nw])
        ctx.reset_stats()
        t0 = time.perf_counter()
        res = ctx.cheapest_path_length(200, src[:nw], dst[:nw])
        secs = time.perf_counter() - t0
        st = ctx.stats()
        out["cheapest_path_length"] = {"pairs": nw, "pairs_per_s": nw / secs, "ms": secs * 1e3,
                                       "batches": int(st["batches"]), "rounds": int(st["levels"]),
                                       "found": int(np.sum(res[1])), "weights": "int64 uniform 1..100"}
    except Exception as ex:  # reported, never fatal for the headline
        out["cheapest_path_length"] = {"error": str(ex)[:200]}
    ctx.csr_delete(200)
    return {"workload": "BASELINE configs[3]: LDBC-SF100-shaped knows graph V=%d E=%d, %d pairs, host buffers "
                        "(e2e through the C ABI)" % (V, E, npairs), **out}


def bench_c1(ctx, steps, warmup, with_cpu):
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
    n = a.pairs
    src, dst = make_pairs(V, v_off, n, 1000 + rank)
    d_ps, d_pd = ctx.dev_alloc(n * 8), ctx.dev_alloc(n * 8)
    d_out, d_ov = ctx.dev_alloc(n * 8), ctx.dev_alloc(n)
    ctx.h2d(d_ps, src)
    ctx.h2d(d_pd, dst)
    # ---- instrumented pass (untimed): TEPS numerator + algorithmic bytes ----
    ctx.set_option("stats", 1)
    ctx.reset_stats()
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    st = ctx.stats()
    ctx.set_option("stats", 0)
    edges_lane, algo_bytes, levels, batches = st["edges_lane"], st["algo_bytes"], st["levels"], st["batches"]
    ref_out = np.empty(n, dtype=np.int64)
    ctx.d2h(ref_out, d_out)
    # ---- level kernels timed alone (untimed pass): one batch at a time on one stream, CUDA events around
    # each level's kernels; used for the kernels' share of a step ----
    ctx.set_option("timing", 1)
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    ctx.reset_stats()
    ctx.timer_start()
    ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    step_alone_ms = ctx.timer_stop_ms()
    st = ctx.stats()
    level_ms, level_launches = st["level_kernel_ms"], int(st["level_kernel_launches"])
    ctx.set_option("timing", 0)
    # batches in flight per call: each has a host thread that spin-waits on its stream once per level, so
    # keep ranks x streams within the host's cores
    cores_per_rank = host_cores() // max(1, world)
    streams = a.streams if a.streams else (8 if cores_per_rank >= 16 else max(1, min(4, cores_per_rank)))
    ctx.set_option("streams", streams)
    # ---- device-resident timed region ----
    for _ in range(a.warmup):
        ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    ctx.sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier_max(dist, local, 0.0)
    ctx.reset_stats()
    ctx.timer_start()
    for _ in range(a.steps):
        ctx.iterativelength_device(0, n, d_ps, d_pd, d_out, d_ov)
    ms_total = ctx.timer_stop_ms()
    ms_total = barrier_max(dist, local, ms_total)
    st = ctx.stats()
    launches = int(st["kernel_launches"])
    # ---- e2e: host buffers through the reference-facing C-ABI call ----
    out_h, ov_h = ctx.iterativelength(0, src, dst)  # warm
    assert np.array_equal(np.where(ov_h == 1, out_h, -1), ref_out), "host and device variants disagree"
    barrier_max(dist, local, 0.0)
    ctx.reset_stats()
    ctx.timer_start()
    for _ in range(a.steps):
        ctx.iterativelength(0, src, dst)
    e2e_ms = barrier_max(dist, local, ctx.timer_stop_ms())
    st2 = ctx.stats()
    clocks = sampler.stop() if rank == 0 else None
    total_edges = barrier_sum(dist, local, float(edges_lane))
    total_pairs = barrier_sum(dist, local, float(n))
    ms_per_step = ms_total / a.steps
    value = total_edges / (ms_per_step * 1e-3) / 1e9
    e2e_value = total_edges / (e2e_ms / a.steps * 1e-3) / 1e9
    # roofline over the timed region: algorithmic bytes of a step / CUDA-event time of a step (batches overlap
    # on several streams, so the whole step is the honest denominator; it includes memsets, checks and host gaps)
    achieved = algo_bytes / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else 0.0
    # measured DRAM traffic of the level kernels (ncu pass committed under profiles/, per 512-lane batch)
    traffic = dram_frac = traffic_src = None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        traffic = float(prof["level_kernel_dram_bytes_per_batch"]) * float(batches)
        dram_frac = traffic / (ms_per_step * 1e-3) / 1e9 / peak_gbs if ms_per_step > 0 else None
        traffic_src = "profiles/r01_traffic.json (%s)" % prof.get("command", "")
    except Exception:
        pass
    line = {
        "metric": "msbfs_gteps", "value": value, "unit": "GTEPS", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 bitset (512 lanes/vertex), u32 ids", "data": "synthetic",
        "config": {"workload": "BASELINE configs[2]: MS-BFS iterativelength, R-MAT scale %d ef%d, %d src/dst "
                               "pairs per GPU in 512-lane batches, CSR replicated" % (scale, ef, n),
                   "V": V, "E": E, "pairs_per_gpu": n, "batches": int(batches), "levels": int(levels),
                   "l2": "inputs larger than L2 (3 bitmaps of V*64 B = %.1f GB)" % (3 * V * 64 / 1e9),
                   "parallelism": "source-batch sharding x%d" % world,
                   "batches_in_flight": int(streams)},
        "pairs_per_s": total_pairs / (ms_per_step * 1e-3),
        "e2e": {"value": e2e_value, "unit": "GTEPS", "pairs_per_s": total_pairs / (e2e_ms / a.steps * 1e-3),
                "h2d_bytes_per_step": int(st2["h2d_bytes"] // a.steps),
                "d2h_bytes_per_step": int(st2["d2h_bytes"] // a.steps)},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                     "frac": achieved / peak_gbs, "traffic": traffic, "dram_frac": dram_frac,
                     "traffic_source": traffic_src, "peak_source": peak_src,
                     "kernel": "MS-BFS level step = pull_list/pull_warp/pull_hub (+push, bitmap) kernels",
                     "note": "achieved = the REFERENCE algorithm's bytes per level (SURVEY.md 8(d)) / time; the "
                             "frontier filter, early exit and L2 hits move fewer DRAM bytes (traffic, dram_frac), "
                             "so frac can exceed 1",
                     "launches": level_launches,
                     "algo_bytes_per_step": int(algo_bytes),
                     "level_kernel_ms_per_step_alone": level_ms,
                     "alone": "level kernels of one step timed with one batch in flight (streams=1, CUDA events "
                              "around each level): achieved %.0f GB/s" % (algo_bytes / (level_ms * 1e-3) / 1e9
                                                                         if level_ms > 0 else 0.0),
                     "step_ms_alone": step_alone_ms,
                     "share_of_step": level_ms / step_alone_ms if step_alone_ms else None},
        "setup": {"csr_build_ms": build_ms, "reverse_csr_ms": transpose_ms},
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
            line["roofline"]["random_gather_gbs"] = {"error": str(ex)}
    if rank == 0 and world == 1 and not a.no_csr:
        line["csr_build"] = bench_csr(ctx, a.csr_scale, ef, a.steps, a.warmup, peak_gbs)
    if rank == 0 and world == 1 and not a.no_paths:
        line["config1"] = bench_c1(ctx, a.steps, a.warmup, not a.no_cpu_baseline)
    if rank == 0 and world == 1 and not a.no_paths:
        line["paths"] = bench_paths(ctx, 450000, 40000000, 16384, max(1, a.steps // 2), a.warmup)
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        e_h = np.empty(E, dtype=np.int64)
        ctx._chk(ctx.lib.pgq_csr_download(ctx._ctx, 0, None, e_h.ctypes.data_as(pgq_b200.I64P), None))
        T, cores = cpu_sample_threads(V, a.cpu_threads)
        rows_pt = min(512, n)
        T = max(1, min(T, n // rows_pt))
        secs, rows, edges, res = run_cpu_sample(V, v_off, e_h, src, dst, T, rows_pt)
        # the CPU sample doubles as a parity check at full size
        for t in range(T):
            lo = t * rows_pt
            eo, ev, _ = res[t]
            assert np.array_equal(np.where(ev == 1, eo, -1), ref_out[lo:lo + rows_pt]), "GPU != oracle at full size"
        line["cpu_baseline"] = {"value": edges / secs / 1e9, "unit": "GTEPS", "cores": T, "host_cores": cores,
                                "kind": "port", "pairs_per_s": rows / secs,
                                "sample": "%d pairs (%d threads x one %d-row call) of the same pair set, %.1f s"
                                          % (rows, T, rows_pt, secs),
                                "parity": "sample results identical to the GPU's"}
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


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
    times, edges_l, rows_l = [], [], []
    for i in range(a.warmup + a.steps):
        if i < a.warmup and i > 0:
            continue  # one warm-up pass is enough to fault the pages in; each pass is tens of seconds
        secs, rows, edges, _ = run_cpu_sample(V, v, e, src, dst, T, rows_pt)
        if i >= a.warmup:
            times.append(secs)
            edges_l.append(edges)
            rows_l.append(rows)
    secs = float(np.mean(times))
    value = float(np.mean(edges_l)) / secs / 1e9
    sample = "%d pairs per step (%d threads x one %d-row call) of the %d-pair set" % (rows_l[0], T, rows_pt, a.pairs)
    line = {"impl": "reference", "metric": "msbfs_gteps", "value": value, "unit": "GTEPS", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": secs * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64 bitset (512 lanes/vertex), int64 ids",
            "data": "synthetic",
            "config": {"workload": "BASELINE configs[2]: MS-BFS iterativelength, R-MAT scale %d ef%d; CPU oracle "
                                   "(port of the reference algorithm), bounded sample" % (scale, ef),
                       "V": V, "E": E, "pairs_per_step": rows_l[0]},
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
