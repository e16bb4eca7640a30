"""CPU-side checks of the C-ABI libraries: they load and export every symbol the headers declare.
No compute call is made (no GPU here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:b200|ggml_backend)_[a-z0-9_]+)\s*\(", src)))


def test_ops_library_exports_every_declared_symbol():
    from llamacog_b200 import build, ops
    assert os.path.exists(build.OPS_LIB), "libb200ops.so not built"
    lib = C.CDLL(build.OPS_LIB)
    names = _declared("b200_ops.h")
    assert names, "no declarations parsed"
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    assert sorted(ops.EXPORTS) == names


def test_ops_fail_loudly_without_library(monkeypatch):
    from llamacog_b200 import build, ops
    monkeypatch.setattr(build, "OPS_LIB", "/nonexistent/libb200ops.so")
    monkeypatch.setattr(ops, "_lib", None)
    with pytest.raises(RuntimeError):
        ops.lib()
