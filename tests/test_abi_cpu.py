"""CPU suite: the C-ABI library loads, exports every symbol include/pgq_b200.h declares, and
fails loudly (no CPU fallback) when there is no GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "pgq_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(pgq_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol():
    import pgq_b200
    lib = pgq_b200.load_library()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "missing export %s" % s


def test_no_cpu_fallback():
    import torch
    import pgq_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pgq_b200.PgqError) as ei:
        pgq_b200.Context(0)
    assert "no CPU fallback" in str(ei.value)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "duckdb-pgq.old_b200")
    for dp, _, fns in os.walk(pkg):
        if "build" in dp:
            continue
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "oracle/" not in txt and "import orc" not in txt and "liborc" not in txt, fn
