"""mhh_mean_static: the no-allocation variant of mhh_mean (per-level partial sums in a file-scope static __device__ array of
4096 slots instead of the context's cudaMalloc'ed scratch; microhh_b200/csrc/static_scratch.cu).  No counterpart in the
reference.  Called through the C ABI twice in a row (the array is shared and keeps the previous call's contents) and checked
against the existing entry point mhh_mean at the tolerance of test_reductions_and_diagnostics (1e-13 relative to the field's
scale), against the reference's calc_mean, and for the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from tests.helpers import make_case

pytestmark = pytest.mark.gpu


def _grid(case, precision="f64"):
    import torch
    from microhh_b200 import host as H
    c = case
    return H.Grid(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], swspatialorder=str(c["order"]),
                  gc=c["gc"], device=torch.device("cuda:0"), precision=precision)


def _mean(g, d):
    from microhh_b200 import lib as L
    out = C.c_double()
    L.check(g.ctx, g.lib.mhh_mean(g.ctx, L.ptr(d), C.byref(out)), "mhh_mean")
    return out.value


def _mean_static(g, d, out_dev):
    from microhh_b200 import lib as L
    L.check(g.ctx, g.lib.mhh_mean_static(g.ctx, L.ptr(d), L.ptr(out_dev)), "mhh_mean_static")


@pytest.mark.parametrize("kw", [dict(itot=32, jtot=24, ktot=16, alpha=1.2), dict(itot=96, jtot=40, ktot=130),
                                dict(itot=24, jtot=20, ktot=10, order=4)], ids=["stretched", "96x40x130", "o4"])
def test_mean_static_matches_mean_on_two_calls_in_a_row(cuda_lib, kw):
    import torch
    from oracle.ref import RefWorld
    case = make_case(**kw)
    g = _grid(case)
    th = case["s"][0]
    rng = np.random.default_rng(11)
    other = th[::-1].copy() * 0.5 + rng.standard_normal(th.shape)        # a second, different field
    fields = [torch.from_numpy(np.ascontiguousarray(a)).to(g.device) for a in (th, other)]
    want = [_mean(g, d) for d in fields]
    out = torch.full((2,), float("nan"), dtype=torch.float64, device=g.device)
    # two calls back to back on the context's stream, no synchronisation in between: the second reuses the static array
    _mean_static(g, fields[0], out[0:1])
    _mean_static(g, fields[1], out[1:2])
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    print(got, want)
    for a, b in zip(got, want):
        assert abs(a - b) <= 1e-13 * 300., (a, b)
    # and again in the other order (the array now holds the second field's partial sums)
    _mean_static(g, fields[0], out[1:2])
    torch.cuda.synchronize()
    assert abs(out[1].item() - want[0]) <= 1e-13 * 300.
    # against the reference's own Field3d_operators::calc_mean
    c = case
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"],
                 rhoref=c["rhoref"], rhorefh=c["rhorefh"])
    assert abs(got[0] - W.call("ref_mean", th)) <= 1e-13 * 300.
    W.close(); g.close()


def test_mean_static_single_precision(cuda_lib):
    """The _f32 symbol (mhh_mean_static_f32) against mhh_mean_f32: same kernels, float fields, FP64 accumulation."""
    import torch
    case = make_case(itot=32, jtot=24, ktot=16, alpha=1.2)
    g = _grid(case, precision="f32")
    rng = np.random.default_rng(5)
    flds = [torch.from_numpy((300. + rng.standard_normal(case["s"][0].shape)).astype(np.float32)).to(g.device) for _ in range(2)]
    want = [_mean(g, d) for d in flds]
    out = torch.zeros(2, dtype=torch.float32, device=g.device)
    for n, d in enumerate(flds):
        _mean_static(g, d, out[n:n + 1])
    torch.cuda.synchronize()
    for a, b in zip(out.cpu().numpy(), want):
        assert abs(a - b) <= 1e-6 * 300., (a, b)
    g.close()


def test_mean_static_rejects_more_levels_than_the_scratch_holds(cuda_lib):
    import torch
    from microhh_b200 import lib as L
    ktot = 4100                                     # > MHH_STATIC_SCRATCH_LEVELS = 4096
    z = (np.arange(ktot) + 0.5) * (1. / ktot)
    from microhh_b200 import host as H
    g = H.Grid(8, 4, ktot, 1., 1., 1., z, swspatialorder="2", device=torch.device("cuda:0"))
    d = torch.zeros((g.kcells, g.jcells, g.icells), dtype=torch.float64, device=g.device)
    out = torch.zeros(1, dtype=torch.float64, device=g.device)
    err = g.lib.mhh_mean_static(g.ctx, L.ptr(d), L.ptr(out))
    assert err == 1                                  # MHH_ERR_ARG
    msg = g.lib.mhh_last_error(g.ctx).decode()
    assert "4096" in msg and "static scratch" in msg, msg
    assert _mean(g, d) == 0.                         # the allocating entry point still takes the grid
    g.close()
