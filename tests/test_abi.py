"""CPU: the C-ABI library loads and exports every symbol include/mhh_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from microhh_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "mhh_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mhh_[A-Za-z0-9_]+)\s*\(", src)))


def test_header_and_python_lists_agree():
    assert header_symbols() == sorted(L.EXPORTS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(L.LIB_PATH), "libmhh_b200.so missing: run __graft_entry__.build()"
    lib = ctypes.CDLL(L.LIB_PATH)
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_create_fails_loudly_without_gpu():
    import numpy as np
    import torch
    if torch.cuda.is_available():
        return
    lib = L.load()
    d = L.GridDesc(8, 8, 8, 1, 1, 1, 1., 1., 1., 2, 1, 0)
    z = (np.arange(8) + 0.5) / 8
    ctx = ctypes.c_void_p()
    err = lib.mhh_create(ctypes.byref(d), z.ctypes.data_as(ctypes.c_void_p), ctypes.byref(ctx))
    assert err != 0 and not ctx.value   # no CPU fallback


def test_rk_sub_dt_host_only():
    lib = L.load()
    assert lib.mhh_rk_sub_dt(3, 0, ctypes.c_double(3.0)) == 1.0
    assert abs(lib.mhh_rk_sub_dt(3, 2, ctypes.c_double(1.0)) - 8. / 15.) < 1e-16


def test_f32_symbol_set():
    """The single-precision symbol set: generated headers up to date, and every declared function exported with `_f32`."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_f32_header as G
    names, body = G.generate()
    assert open(os.path.join(ROOT, "include", "mhh_b200_f32_names.h")).read() == names, "run tools/gen_f32_header.py"
    assert open(os.path.join(ROOT, "include", "mhh_b200_f32.h")).read() == body, "run tools/gen_f32_header.py"
    lib = ctypes.CDLL(L.LIB_PATH)
    missing = [s + "_f32" for s in header_symbols() if not hasattr(lib, s + "_f32")]
    assert not missing, missing
    assert L.Api("f32").mhh_rk_sub_dt(3, 2, 1.0) == 8. / 15.        # time steps stay double in the f32 set
