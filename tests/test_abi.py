"""The C-ABI library loads on a CPU-only box and exports every symbol include/cellector_cuda.h declares
(no compute calls here: those need a GPU and live in the -m gpu tests)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "cellector_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(clx_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    from cellector_b200 import api
    lib = api.load_library()
    names = _declared_symbols()
    assert len(names) >= 20, names
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/cellector_cuda.h but not exported"
    # and the ctypes table binds exactly the declared surface
    assert sorted("clx_" + k for k in api.ABI) == names


def test_create_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    from cellector_b200 import api
    lib = api.load_library()
    h = ctypes.c_void_p()
    rc = lib.clx_create(ctypes.byref(h), 0, 0)
    assert rc != 0 and not h.value
    assert b"no CPU fallback" in lib.clx_last_error(None)
