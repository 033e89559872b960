"""TEST INFRASTRUCTURE ONLY — ctypes front-end of ``oracle/_ref/liboracle_ref.so``.

That library is the reference's OWN hot-path sources (``/root/reference/src/*.cxx``)
compiled in place by ``oracle/Makefile`` plus thin harness TUs (``oracle/ref_harness``).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module; the product (``microhh_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "liboracle_ref.so")
# the timing build (-O3 -march=x86-64-v3, FP contraction on: oracle/Makefile) — bench.py's CPU legs call use_perf_build()
# BEFORE the first lib(); parity tests always use the plain build above
PERF_LIB_PATH = os.path.join(_HERE, "_ref", "liboracle_ref_perf.so")

_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_lib = None


def use_perf_build() -> bool:
    """Select the timing build for this process (no effect once the library is loaded, or when it was not built)."""
    global LIB_PATH
    if _lib is None and os.path.exists(PERF_LIB_PATH):
        LIB_PATH = PERF_LIB_PATH
    return LIB_PATH == PERF_LIB_PATH


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB_PATH} missing: run `make -C oracle` where /root/reference exists")
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_world_create.restype = C.c_void_p
        for name in ("ref_mean_2d", "ref_mean", "ref_advec2_cfl", "ref_pres2_divergence", "ref_pres4_divergence",
                     "ref_smag2_dnmul", "ref_rk_subdt", "ref_advec2i5_cfl", "ref_advec4_cfl",
                     "ref_check_momentum", "ref_check_tke", "ref_check_mass"):
            if hasattr(_lib, name):
                getattr(_lib, name).restype = C.c_double
    return _lib


def _p(a):
    """Pointer to a float64 C-contiguous array (or NULL)."""
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], "oracle arrays must be contiguous float64"
    return a.ctypes.data_as(C.c_void_p)


class RefWorld:
    """The reference's Grid/FFT/Pres objects for one grid (serial build, npx=npy=1)."""

    def __init__(self, itot, jtot, ktot, xsize, ysize, zsize, z, order=2, gc=None, rhoref=None, rhorefh=None):
        L = lib()
        if gc is None:
            gc = (1, 1, 1) if order == 2 else (3, 3, 3)
        z = np.ascontiguousarray(z, dtype=np.float64)
        assert z.shape == (ktot,)
        self.h = C.c_void_p(L.ref_world_create(
            C.c_int(itot), C.c_int(jtot), C.c_int(ktot), C.c_double(xsize), C.c_double(ysize), C.c_double(zsize),
            C.c_int(order), C.c_int(gc[0]), C.c_int(gc[1]), C.c_int(gc[2]), _p(z)))
        if not self.h:
            raise RuntimeError("ref_world_create failed")
        self.order = order
        d = np.zeros(20, dtype=np.int32)
        L.ref_grid_dims(self.h, d.ctypes.data_as(C.c_void_p))
        (self.itot, self.jtot, self.ktot, self.imax, self.jmax, self.kmax, self.igc, self.jgc, self.kgc,
         self.icells, self.jcells, self.kcells, self.ijcells, self.ncells,
         self.istart, self.iend, self.jstart, self.jend, self.kstart, self.kend) = (int(x) for x in d)
        self.shape = (self.kcells, self.jcells, self.icells)
        sc = np.zeros(6)
        L.ref_grid_array(self.h, b"scalars", _p(sc))
        self.dx, self.dy, self.dxi, self.dyi, self.dzhi4bot, self.dzhi4top = (float(x) for x in sc)
        self.xsize, self.ysize, self.zsize = xsize, ysize, zsize
        for name in ("z", "zh", "dz", "dzh", "dzi", "dzhi", "dzi4", "dzhi4"):
            a = np.zeros(self.kcells)
            L.ref_grid_array(self.h, name.encode(), _p(a))
            setattr(self, name, a)
        self.rhoref = np.ones(self.kcells) if rhoref is None else np.ascontiguousarray(rhoref, dtype=np.float64)
        self.rhorefh = np.ones(self.kcells) if rhorefh is None else np.ascontiguousarray(rhorefh, dtype=np.float64)
        L.ref_world_set_rhoref(self.h, _p(self.rhoref), _p(self.rhorefh))
        L.ref_pres_set_values(self.h)

    def close(self):
        if getattr(self, "h", None):
            lib().ref_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- generic call helper: every array argument goes through _p -------------------------
    def call(self, fname, *args):
        f = getattr(lib(), fname)
        cargs = [self.h]
        for a in args:
            if a is None or isinstance(a, np.ndarray):
                cargs.append(_p(a))
            elif isinstance(a, (int, np.integer)):
                cargs.append(C.c_int(int(a)))
            else:
                cargs.append(C.c_double(float(a)))
        return f(*cargs)
