"""CPU-side checks of the plugin library: it loads without a GPU, exports the two loader symbols that
include/ggml-b200.h declares, returns a registry with the right ABI version and zero devices here, and
refuses to score on a box without an sm_100 device (so the host never picks it up by accident)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "llamacog_b200", "libggml-b200.so")


def _load():
    if not os.path.exists(PLUGIN):
        pytest.skip("plugin not built (needs the reference headers)")
    C.CDLL(os.path.join(ROOT, "llamacog_b200", "libb200ops.so"), mode=C.RTLD_GLOBAL)
    C.CDLL(os.path.join(ROOT, "baseline", "_ref", "lib", "libggml-base.so"), mode=C.RTLD_GLOBAL)
    return C.CDLL(PLUGIN)


def test_plugin_exports_loader_symbols():
    lib = _load()
    src = open(os.path.join(ROOT, "include", "ggml-b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = sorted(set(re.findall(r"\b(ggml_backend_[a-z0-9_]+)\s*\(", src)))
    assert names == ["ggml_backend_b200_split_rows", "ggml_backend_init", "ggml_backend_score"]
    for n in names:
        assert hasattr(lib, n)


def test_plugin_registry_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _load()
    assert lib.ggml_backend_score() == 0

    class Reg(C.Structure):
        _fields_ = [("api_version", C.c_int), ("get_name", C.c_void_p), ("get_device_count", C.c_void_p),
                    ("get_device", C.c_void_p), ("get_proc_address", C.c_void_p), ("context", C.c_void_p)]
    lib.ggml_backend_init.restype = C.POINTER(Reg)
    reg = lib.ggml_backend_init()
    assert reg.contents.api_version == 1          # GGML_BACKEND_API_VERSION, ggml-backend-impl.h:11
    count = C.CFUNCTYPE(C.c_size_t, C.c_void_p)(reg.contents.get_device_count)
    assert count(C.cast(reg, C.c_void_p)) == 0
    name = C.CFUNCTYPE(C.c_char_p, C.c_void_p)(reg.contents.get_name)
    assert name(C.cast(reg, C.c_void_p)) == b"B200"
    assert lib.ggml_backend_init() and C.addressof(lib.ggml_backend_init().contents) == C.addressof(reg.contents)   # idempotent
    # the names the host asks the registry for (reference ggml-cuda.cu:3474-3496): answered; anything else: NULL
    proc = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_char_p)(reg.contents.get_proc_address)
    for n in (b"ggml_backend_split_buffer_type", b"ggml_backend_get_features", b"ggml_backend_register_host_buffer",
              b"ggml_backend_unregister_host_buffer"):
        assert proc(C.cast(reg, C.c_void_p), n), n
    assert not proc(C.cast(reg, C.c_void_p), b"ggml_backend_no_such_thing")
