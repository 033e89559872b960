"""GPU parity of Buffer::exec and Force::exec through the C ABI / the host mirror classes against the
reference's own kernels (src/buffer.cxx:37-58, src/force.cxx:46-292 compiled in place, oracle/_ref).  Every term is one or
two IEEE operations per point on identical inputs (FMA contraction off on both sides), so the bar is bit-exact except for
fixed-flux, whose body force contains two xy-z mean reductions (summation order differs: 1e-13)."""
import numpy as np
import pytest

from tests.helpers import make_case, rel_err
from tests.test_gpu_ops import _up, _worlds

pytestmark = pytest.mark.gpu

CASES = [dict(itot=16, jtot=12, ktot=16), dict(itot=20, jtot=9, ktot=12, order=4, alpha=1.5), dict(itot=16, jtot=1, ktot=8)]
IDS = ["o2", "o4_stretched", "o2_2d"]


def _fields(g, case, slist=("th",)):
    from microhh_b200 import host as H
    f = H.Fields(g, slist=list(slist))
    for n in ("u", "v", "w"):
        f.mp[n].fld.copy_(_up(g, case[n]))
    for n, a in zip(slist, case["s"]):
        f.sp[n].fld.copy_(_up(g, a))
    return f


def _tend(g, f, rng, shape):
    t0 = {}
    for n in f.at:
        t0[n] = rng.uniform(-1, 1, shape)
        f.at[n].fld.copy_(_up(g, t0[n]))
    return t0


@pytest.mark.parametrize("kw", CASES, ids=IDS)
@pytest.mark.parametrize("swupdate", [False, True])
def test_buffer(cuda_lib, kw, swupdate):
    from microhh_b200 import host as H
    case = make_case(zsize=3200., **kw)
    g, W = _worlds(case)
    f = _fields(g, case)
    rng = np.random.default_rng(5)
    t0 = _tend(g, f, rng, W.shape)
    zstart, sigma, beta = 2300., 0.00223, 2.
    profs = {"u": rng.uniform(-1, 1, W.ktot), "v": rng.uniform(-1, 1, W.ktot), "th": 300. + rng.uniform(0, 10, W.ktot)}
    b = H.Buffer(g, f, swbuffer=True, swupdate=swupdate, zstart=zstart, sigma=sigma, beta=beta, profs=profs)
    assert b.bufferkstart == W.call("ref_buffer_kstart", zstart, 0)
    assert b.bufferkstarth == W.call("ref_buffer_kstart", zstart, 1)
    assert W.kstart < b.bufferkstart < W.kend
    f.exec()
    b.exec()
    for n in f.at:
        fld = case[n] if n in ("u", "v", "w") else case["s"][0]
        if swupdate:
            abuf = np.zeros(W.kcells)
            W.call("ref_mean_profile", abuf, fld)
            assert rel_err(f.ap[n].fld_mean.cpu().numpy(), abuf) <= 1e-14
            abuf = f.ap[n].fld_mean.cpu().numpy()      # the same profile on both sides: the kernel itself is then bit-exact
        else:
            abuf = np.zeros(W.kcells)
            if n in profs:
                abuf[W.kstart:W.kend] = profs[n]
        ref = t0[n].copy()
        W.call("ref_buffer", ref, fld, abuf, zstart, sigma, beta, 1 if n == "w" else 0)
        got = f.at[n].fld.cpu().numpy()
        assert np.array_equal(got, ref), n
        assert not np.array_equal(got, t0[n])
    g.close(); W.close()


def test_buffer_too_close_to_top(cuda_lib):
    from microhh_b200 import host as H, lib as L
    case = make_case(itot=8, jtot=8, ktot=8)
    g, W = _worlds(case)
    f = H.Fields(g)
    with pytest.raises(L.MhhError, match="Buffer is too close to the model top"):
        H.Buffer(g, f, swbuffer=True, zstart=case["zsize"] * 2., profs={})
    g.close(); W.close()


@pytest.mark.parametrize("kw", CASES, ids=IDS)
def test_force_coriolis_and_sources(cuda_lib, kw):
    from microhh_b200 import host as H
    case = make_case(**kw)
    g, W = _worlds(case)
    f = _fields(g, case)
    for fld in f.ap.values():       # the Coriolis stencil reads the cyclic ghost cells
        H.Boundary_cyclic(g).exec(fld.fld)
    u, v = f.mp["u"].fld.cpu().numpy(), f.mp["v"].fld.cpu().numpy()
    rng = np.random.default_rng(6)
    t0 = _tend(g, f, rng, W.shape)
    profs = {"u_geo": rng.uniform(5, 10, W.ktot), "v_geo": rng.uniform(-1, 1, W.ktot), "th_ls": rng.uniform(-1e-3, 1e-3, W.ktot),
             "u_ls": rng.uniform(-1e-3, 1e-3, W.ktot)}
    fo = H.Force(g, f, swlspres="geo", fc=1.39e-4, swls="1", lslist=["th", "u"], profs=profs, utrans=0.5, vtrans=-0.25)
    fo.exec(1.0)
    ug, vg = np.zeros(W.kcells), np.zeros(W.kcells)
    ug[W.kstart:W.kend], vg[W.kstart:W.kend] = profs["u_geo"], profs["v_geo"]
    rut, rvt = t0["u"].copy(), t0["v"].copy()
    W.call("ref_force_coriolis", rut, rvt, u, v, ug, vg, 1.39e-4, 0.5, -0.25)
    ls = np.zeros(W.kcells); ls[W.kstart:W.kend] = profs["u_ls"]
    W.call("ref_force_ls_source", rut, ls)
    rtht = t0["th"].copy()
    ls[W.kstart:W.kend] = profs["th_ls"]
    W.call("ref_force_ls_source", rtht, ls)
    assert np.array_equal(f.mt["u"].fld.cpu().numpy(), rut)
    assert np.array_equal(f.mt["v"].fld.cpu().numpy(), rvt)
    assert np.array_equal(f.st["th"].fld.cpu().numpy(), rtht)
    assert np.array_equal(f.mt["w"].fld.cpu().numpy(), t0["w"])
    g.close(); W.close()


@pytest.mark.parametrize("kw", CASES, ids=IDS)
@pytest.mark.parametrize("sw", ["uflux", "dpdx"])
def test_force_pressure(cuda_lib, kw, sw):
    from microhh_b200 import host as H
    case = make_case(**kw)
    g, W = _worlds(case)
    f = _fields(g, case)
    rng = np.random.default_rng(7)
    t0 = _tend(g, f, rng, W.shape)
    fo = H.Force(g, f, swlspres=sw, uflux=0.11, dpdx=-3e-3, utrans=0.02)
    fo.exec(0.37)
    ref = t0["u"].copy()
    if sw == "uflux":
        W.call("ref_force_fixed_flux", ref, case["u"], 0.11, 0.02, 0.37)
        assert rel_err(f.mt["u"].fld.cpu().numpy(), ref) <= 1e-13
    else:
        W.call("ref_force_pressure", ref, 3e-3)
        assert np.array_equal(f.mt["u"].fld.cpu().numpy(), ref)
    assert np.array_equal(f.mt["v"].fld.cpu().numpy(), t0["v"])
    g.close(); W.close()


@pytest.mark.parametrize("kw", CASES[:1] + CASES[2:], ids=IDS[:1] + IDS[2:])
@pytest.mark.parametrize("swwls", ["mean", "local"])
def test_force_subsidence_and_nudging(cuda_lib, kw, swwls):
    from microhh_b200 import host as H
    case = make_case(**kw)
    g, W = _worlds(case)
    f = _fields(g, case)
    rng = np.random.default_rng(8)
    t0 = _tend(g, f, rng, W.shape)
    wls = rng.uniform(-0.01, 0.01, W.ktot)          # both signs: both upwind branches
    profs = {"w_ls": wls, "nudgefac": rng.uniform(0, 1e-3, W.ktot), "th_nudge": 300. + rng.uniform(0, 10, W.ktot),
             "u_nudge": rng.uniform(-1, 1, W.ktot)}
    fo = H.Force(g, f, swwls=swwls, swwls_mom=True, swnudge="1", nudgelist=["th", "u"], profs=profs, utrans=0.3)
    f.exec()
    fo.exec(1.0)
    wl = np.zeros(W.kcells); wl[W.kstart:W.kend] = wls
    fac = np.zeros(W.kcells); fac[W.kstart:W.kend] = profs["nudgefac"]
    flds = {"u": case["u"], "v": case["v"], "w": case["w"], "th": case["s"][0]}
    means = {n: f.ap[n].fld_mean.cpu().numpy() for n in flds}
    for n in ("u", "v", "th"):
        ref = t0[n].copy()
        W.call("ref_force_wls", 0 if swwls == "mean" else 1, ref, means[n] if swwls == "mean" else flds[n], wl)
        if n in ("th", "u"):
            nud = np.zeros(W.kcells)
            nud[W.kstart:W.kend] = profs[n + "_nudge"] - (0.3 if n == "u" else 0.)
            W.call("ref_force_nudge", ref, means[n], nud, fac)
        assert np.array_equal(f.at[n].fld.cpu().numpy(), ref), n
        assert not np.array_equal(ref, t0[n])
    ref = t0["w"].copy()
    if swwls == "local":
        W.call("ref_force_wls", 2, ref, flds["w"], wl)
    assert np.array_equal(f.mt["w"].fld.cpu().numpy(), ref)
    g.close(); W.close()


def test_force_illegal_switch(cuda_lib):
    from microhh_b200 import host as H, lib as L
    case = make_case(itot=8, jtot=8, ktot=8)
    g, W = _worlds(case)
    f = H.Fields(g)
    with pytest.raises(L.MhhError, match="swlspres"):
        H.Force(g, f, swlspres="bogus")
    with pytest.raises(L.MhhError, match="lslist"):
        H.Force(g, f, swls="1", lslist=["qt"], profs={"qt_ls": np.zeros(8)})
    g.close(); W.close()
