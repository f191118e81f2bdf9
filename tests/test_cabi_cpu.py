"""CPU: the C-ABI library loads, exports every symbol include/pn_b200.h declares, and the ctypes
signature table covers the header one to one.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pn_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pn_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    from pacednegatives_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib


def test_header_symbols_exported(built_lib):
    names = declared_symbols()
    assert len(names) >= 15
    raw = ctypes.CDLL(built_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in pn_b200.h but not exported"


def test_signature_table_matches_header(built_lib):
    assert sorted(built_lib.SIGNATURES) == declared_symbols()
    lib = built_lib.lib()
    assert lib.pn_version() == 100
    assert lib.pn_launch_count() >= 0


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        built_lib.require_gpu()
    import pacednegatives_b200.weights as w
    with pytest.raises(RuntimeError):
        w.LCEWeights(0.5)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pacednegatives_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
