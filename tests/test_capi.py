"""The drop-in boundary on a CPU-only box: the C-ABI library builds, loads, exports every symbol
include/readserver_b200.h declares, and -- with no CUDA device -- fails loudly instead of computing
anything on the host."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from readserver_b200 import build as rs_build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "readserver_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rs_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(rs_build.build())
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/readserver_b200.h but not exported"


def test_engine_binds_every_declared_symbol():
    from readserver_b200 import engine
    lib = engine.lib()
    for n in declared_symbols():
        f = getattr(lib, n)
        assert f.argtypes is not None, f"{n} has no ctypes signature in engine.py"


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from readserver_b200 import engine
    runs = np.array([(1 << 5) | 3, (0 << 5) | 1], dtype=np.uint8)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(runs=runs, num_strings=1, num_symbols=4)
    assert e.value.status == 6 and "no CPU path" in str(e.value)


def test_product_never_imports_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "readserver_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M), f
                assert "rs_oracle" not in txt and "liboracle" not in txt and "rs_emul" not in txt, f
