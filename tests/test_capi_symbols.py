"""CPU: the C-ABI library loads and exports every symbol include/nimb200.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "nimb200.h")).read()
    return sorted(set(re.findall(r"\b(nimb200_[a-zA-Z_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib_path = os.path.join(ROOT, "nimble_b200", "libnimb200.so")
    assert os.path.exists(lib_path), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(lib_path)
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), s
    lib.nimb200_version.restype = ctypes.c_int
    assert lib.nimb200_version() == 1


def test_struct_layouts_match_header_sizes():
    from nimble_b200 import table as T
    assert ctypes.sizeof(T.Op) == 24
    assert ctypes.sizeof(T.Decl) == 4 * (5 + 3 + 3 + 1 + 6)
    assert ctypes.sizeof(T.Dep) == 32
    assert ctypes.sizeof(T.Sampler) == 14 * 4 + 4 * 8


def test_no_device_is_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        return
    from nimble_b200 import _capi
    from tests import small_models as sm
    import nimble_b200 as nb
    m, conf = sm.SMALL["pump"]()
    cm = nb.compileNimble(nb.buildMCMC(conf))
    try:
        nb.runMCMC(cm, niter=5)
    except _capi.BackendError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("running without a CUDA device must fail")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "nimble_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("the oracle", "").replace("CPU oracle", "").replace("oracle's", "").replace("oracle so", ""), (dirpath, f)
