"""ctypes binding of ``libmhh_b200.so`` (the C ABI in ``include/mhh_b200.h``).

There is no CPU fallback: if the shared library is missing this module raises on first use,
and ``mhh_create`` itself fails when no CUDA device is visible.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MHH_B200_LIB: another build of the same library (tools/ experiments: e.g. a -fmad=true build); never a fallback
LIB_PATH = os.environ.get("MHH_B200_LIB") or os.path.join(_HERE, "libmhh_b200.so")

ADVEC_2, ADVEC_2I5, ADVEC_4 = 2, 25, 4
BC_DIRICHLET, BC_NEUMANN, BC_FLUX, BC_USTAR, BC_OFF = 0, 1, 2, 3, 4
BCW_NORMAL, BCW_CONSERVATION = 0, 1
EDGE_EAST_WEST, EDGE_NORTH_SOUTH, EDGE_BOTH = 0, 1, 2


class GridDesc(C.Structure):
    _fields_ = [("itot", C.c_int), ("jtot", C.c_int), ("ktot", C.c_int),
                ("igc", C.c_int), ("jgc", C.c_int), ("kgc", C.c_int),
                ("xsize", C.c_double), ("ysize", C.c_double), ("zsize", C.c_double),
                ("order", C.c_int), ("npy", C.c_int), ("rank_y", C.c_int)]


class GridDescF32(C.Structure):
    """mhh_grid_desc_f32 (include/mhh_b200_f32.h)"""
    _fields_ = [("itot", C.c_int), ("jtot", C.c_int), ("ktot", C.c_int),
                ("igc", C.c_int), ("jgc", C.c_int), ("kgc", C.c_int),
                ("xsize", C.c_float), ("ysize", C.c_float), ("zsize", C.c_float),
                ("order", C.c_int), ("npy", C.c_int), ("rank_y", C.c_int)]


class MhhError(RuntimeError):
    pass


_lib = None

# every symbol include/mhh_b200.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = [
    "mhh_create", "mhh_destroy", "mhh_last_error", "mhh_set_stream", "mhh_set_rhoref", "mhh_get_dims",
    "mhh_get_metric", "mhh_get_metric_device",
    "mhh_launch_count", "mhh_profile_enable", "mhh_profile_report",
    "mhh_advec_exec", "mhh_advec_cfl", "mhh_advec_time_limit",
    "mhh_diff2_exec", "mhh_diff4_exec", "mhh_advec4_diff4_exec", "mhh_diff_const_dnmul",
    "mhh_smag2_exec_viscosity", "mhh_smag2_exec_viscosity_dry", "mhh_smag2_exec_viscosity_dry_limits", "mhh_smag2_exec", "mhh_advec2_smag2_exec", "mhh_advec2_smag2_exec_rk", "mhh_advec2_smag2_exec_rk_parts", "mhh_smag2_strain2", "mhh_smag2_dnmul", "mhh_set_lazy_tendencies", "mhh_advec2_smag2_set_force_tail", "mhh_set_tma", "mhh_tma_active", "mhh_pres_set_fused_input",
    "mhh_thermo_set_ref", "mhh_thermo_dry_exec", "mhh_thermo_dry_N2",
    "mhh_surface_init_solver", "mhh_surface_get_lut", "mhh_boundary_surface_exec", "mhh_boundary_surface_exec_stages", "mhh_surface_scalar",
    "mhh_buffer_plan", "mhh_buffer_exec", "mhh_force_pressure", "mhh_force_fixed_flux", "mhh_force_coriolis",
    "mhh_force_large_scale_source", "mhh_force_nudge", "mhh_force_subsidence",
    "mhh_pres_set_values", "mhh_pres_exec", "mhh_pres_check_divergence", "mhh_pres_work", "mhh_pres_input",
    "mhh_fft_force_generic", "mhh_fft_forward", "mhh_fft_backward", "mhh_pres_solve_spectral", "mhh_pres_output", "mhh_pres_scatter",
    "mhh_pres_get_coefs", "mhh_pres_input_to", "mhh_fft_x", "mhh_fft_y", "mhh_transpose_copy", "mhh_pres_solve_spectral_ws",
    "mhh_pres_output_stage", "mhh_halo_ns", "mhh_fft_x_chunks", "mhh_fft_y_chunks", "mhh_fft_chunks_supported",
    "mhh_peer_alloc", "mhh_peer_free", "mhh_peer_export", "mhh_peer_import", "mhh_peer_unmap", "mhh_peer_register", "mhh_peer_halo_push",
    "mhh_peer_barrier", "mhh_peer_error", "mhh_fft_x_chunks_to_peers", "mhh_fft_y_chunks_to_peers",
    "mhh_cyclic", "mhh_cyclic_2d", "mhh_ghost_cells", "mhh_ghost_cells_w", "mhh_set_bc",
    "mhh_rk_substep", "mhh_rk_sub_dt", "mhh_pres_output_rk", "mhh_pres_exec_rk",
    "mhh_mean_profile", "mhh_level_sums", "mhh_stats_mean", "mhh_stats_moment", "mhh_stats_cov", "mhh_mean", "mhh_mean_2d", "mhh_max",
    "mhh_check_momentum", "mhh_check_tke", "mhh_check_mass", "mhh_mean_static",
    "mhh_pres_set_output_wait", "mhh_slab_fft_x_fwd_sub", "mhh_slab_fft_y_fwd_sub", "mhh_slab_solve_sub", "mhh_slab_fft_y_bwd_sub", "mhh_slab_fft_x_bwd_sub",
]


def load():
    """Load the CUDA library or raise — never falls back to anything else."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MhhError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(the B200 path has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    lib.mhh_last_error.restype = C.c_char_p
    lib.mhh_pres_work.restype = C.c_void_p
    lib.mhh_get_metric_device.restype = C.c_void_p
    lib.mhh_launch_count.restype = C.c_ulonglong
    lib.mhh_rk_sub_dt.restype = C.c_double
    lib.mhh_rk_sub_dt.argtypes = [C.c_int, C.c_int, C.c_double]
    _lib = lib
    return lib


def ptr(t):
    """Device pointer of a torch tensor (or NULL)."""
    if t is None:
        return C.c_void_p(0)
    return C.c_void_p(t.data_ptr())


def ptr_array(tensors):
    """C array of device pointers."""
    arr = (C.c_void_p * max(len(tensors), 1))()
    for n, t in enumerate(tensors):
        arr[n] = t.data_ptr() if t is not None else 0
    return arr


def check(ctx, err, what):
    if err != 0:
        msg = load().mhh_last_error(ctx).decode() if ctx else ""
        raise MhhError(f"{what} failed with code {err}: {msg}")


# ---- precision-aware view of the library ------------------------------------------------------------------------------
_PROTO = None


def _prototypes():
    """{function name: [parameter kind]} parsed from include/mhh_b200.h: 'real' (a `double` scalar: float in the _f32 set),
    'f64' (`mhh_double`: always double), 'ptr' (any pointer), 'int' (everything else)."""
    global _PROTO
    if _PROTO is None:
        import re
        src = open(os.path.join(os.path.dirname(_HERE), "include", "mhh_b200.h")).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        _PROTO = {}
        for m in re.finditer(r"\b(mhh_[A-Za-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
            kinds = []
            for prm in m.group(2).split(","):
                prm = prm.strip()
                if not prm or prm == "void":
                    continue
                if "*" in prm:
                    kinds.append("ptr")
                elif re.match(r"(const\s+)?mhh_double\b", prm):
                    kinds.append("f64")
                elif re.match(r"(const\s+)?double\b", prm):
                    kinds.append("real")
                else:
                    kinds.append("int")
            _PROTO[m.group(1)] = kinds
    return _PROTO


class Api:
    """The library as one precision sees it: ``Api("f64").mhh_xxx`` is the export itself; ``Api("f32").mhh_xxx`` calls
    ``mhh_xxx_f32`` and adapts the arguments the double-precision host code passes: `double` scalars become float (time
    steps, `mhh_double`, stay double) and by-reference double outputs go through a float temporary.  Arrays are the caller's
    business: every device tensor and host array must already have the working type (Grid.dtype / Grid.npdtype)."""

    def __init__(self, precision="f64"):
        if precision not in ("f64", "f32"):
            raise MhhError("precision must be f64 or f32")
        self.precision, self._lib, self._cache = precision, load(), {}
        self.real = C.c_double if precision == "f64" else C.c_float
        self.GridDesc = GridDesc if precision == "f64" else GridDescF32

    def __getattr__(self, name):
        if not name.startswith("mhh_"):
            raise AttributeError(name)
        fn = self._cache.get(name)
        if fn is None:
            fn = self._cache[name] = self._make(name)
        return fn

    def _make(self, name):
        if self.precision == "f64":
            return getattr(self._lib, name)
        raw = getattr(self._lib, name + "_f32")
        raw.restype = getattr(self._lib, name).restype
        kinds = _prototypes().get(name)
        byref_t = type(C.byref(C.c_int()))

        def call(*args):
            conv, fix = [], []
            for n, a in enumerate(args):
                if isinstance(a, C.c_double):
                    kind = kinds[n] if kinds and n < len(kinds) else "real"
                    conv.append(a if kind == "f64" else C.c_float(a.value))
                elif isinstance(a, float):
                    kind = kinds[n] if kinds and n < len(kinds) else "real"
                    conv.append(C.c_double(a) if kind == "f64" else C.c_float(a))
                elif isinstance(a, byref_t) and isinstance(a._obj, C.c_double):
                    tmp = C.c_float()
                    conv.append(C.byref(tmp))
                    fix.append((a._obj, tmp))
                else:
                    conv.append(a)
            r = raw(*conv)
            for dst, tmp in fix:
                dst.value = tmp.value
            return r
        call.__name__ = name
        return call
