"""CPU suite: the reference arm of bench.py (the CPU restatement timed on host cores) keeps the driver's
contract -- exactly one JSON line on stdout with the agreed keys; under a multi-rank launch only rank 0 prints."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
        "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def _run(env_extra, *flags):
    env = dict(os.environ)
    env.update(env_extra)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--scale", "11",
                          "--pairs", "1024", "--steps", "1", "--warmup", "0", *flags],
                         capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-1500:]
    return [l for l in out.stdout.splitlines() if l.strip()]


def test_reference_arm_prints_one_json_line():
    lines = _run({})
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert KEYS <= set(d), KEYS - set(d)
    assert d["impl"] == "reference" and d["metric"] == "msbfs_gteps" and d["unit"] == "GTEPS"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_stay_silent():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2") == []
    lines = _run({"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"}, "--gpus", "2")
    assert len(lines) == 1 and json.loads(lines[0])["n_gpus"] == 2


import pytest  # noqa: E402


@pytest.mark.gpu
def test_own_arm_contract_on_a_small_graph():
    """GPU: the product arm at a small scale carries every key the measurement contract names."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--scale", "15", "--csr-scale", "15",
                          "--pairs", "4096", "--steps", "1", "--warmup", "3", "--cpu-threads", "2", "--no-paths"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-1500:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    need = (KEYS - {"impl"}) | {"gpu_launches", "roofline", "clocks", "csr_build"}
    assert need <= set(d), need - set(d)
    assert d["n_gpus"] == 1 and d["gpu_launches"] > 0 and d["value"] > 0 and d["e2e"]["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["frac"] <= 1.2, "roofline.frac must be a fraction of the peak (the implementation's own byte model)"
    assert r["reference_model"]["frac"] > 0 and r["model_bytes_per_step"] > 0
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] >= 1 and c["value"] > 0
    # parity at the run's size: CSRs memcmp-equal the oracle's host-built arrays, lengths on the oracle's own CSR
    par = d["parity"]
    assert par["csr_config3"] == "identical" and par["csr_config2"] == "identical", par
    assert par["lengths_config3"] == "identical" and "FAILED" not in par, par
    assert d["csr_build"]["unit"] == "GE/s" and d["csr_build"]["value"] > 0 and d["csr_build"]["parity"] == "identical"
    assert d["host_syncs_in_level_loops"] == 0 and d["e2e_chunked"]["value"] > 0
