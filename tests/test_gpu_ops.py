"""GPU parity, per call, of the small operators around the stencils: Boundary_cyclic, vertical ghost rows (2nd/4th order,
every Boundary_type), set_bc, RK3/RK4 substeps, Field3d_operators reductions, Fields diagnostics, thermo_dry pieces —
each through the C ABI against the reference's own code (oracle/_ref), src/fields.cxx included (compiled in place through
Boost / NetCDF shims); the numpy restatement in oracle/restate.py is checked beside it.  Copies and ghost rows are bit-exact; arithmetic is within 1e-13."""
import ctypes as C

import numpy as np
import pytest

from tests.helpers import make_case, rel_err

pytestmark = pytest.mark.gpu


def _worlds(case):
    import torch
    from microhh_b200 import host as H
    from oracle.ref import RefWorld
    c = case
    g = H.Grid(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], swspatialorder=str(c["order"]),
               gc=c["gc"], device=torch.device("cuda:0"))
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"],
                 rhoref=c["rhoref"], rhorefh=c["rhorefh"])
    return g, W


def _up(g, a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(g.device)


CASES = [dict(itot=16, jtot=12, ktot=8), dict(itot=16, jtot=12, ktot=8, order=4, alpha=1.5), dict(itot=16, jtot=1, ktot=8),
         dict(itot=16, jtot=1, ktot=8, order=4), dict(itot=20, jtot=8, ktot=10, gc=(3, 3, 2))]
IDS = ["o2", "o4_stretched", "o2_2d", "o4_2d", "gc332"]


@pytest.mark.parametrize("kw", CASES, ids=IDS)
def test_cyclic_bit_exact(cuda_lib, kw):
    from microhh_b200 import host as H
    case = make_case(**kw)
    g, W = _worlds(case)
    bc = H.Boundary_cyclic(g)
    for edge in (0, 1, 2):
        a = case["u"].copy()
        d = _up(g, a)
        bc.exec(d, edge)
        W.call("ref_cyclic", a, edge)
        assert np.array_equal(d.cpu().numpy(), a), edge
    a2 = case["v"][0].copy()
    d2 = _up(g, a2)
    bc.exec_2d(d2)
    W.call("ref_cyclic_2d", a2)
    assert np.array_equal(d2.cpu().numpy(), a2)
    g.close(); W.close()


@pytest.mark.parametrize("kw", CASES[:4], ids=IDS[:4])
@pytest.mark.parametrize("bcbot,bctop", [(0, 0), (1, 1), (2, 1), (0, 2)])
def test_ghost_cells(cuda_lib, kw, bcbot, bctop):
    case = make_case(**kw)
    g, W = _worlds(case)
    rng = np.random.default_rng(3)
    pl = [rng.uniform(-1, 1, (W.jcells, W.icells)) for _ in range(4)]
    a = case["s"][0].copy()
    d = _up(g, a)
    dp = [_up(g, p) for p in pl]
    from microhh_b200 import lib as L
    L.check(g.ctx, g.lib.mhh_ghost_cells(g.ctx, L.ptr(d), bcbot, bctop, L.ptr(dp[0]), L.ptr(dp[1]), L.ptr(dp[2]), L.ptr(dp[3])), "ghost")
    W.call("ref_ghost_cells", a, bcbot, bctop, pl[0], pl[1], pl[2], pl[3])
    got = d.cpu().numpy()
    assert np.array_equal(got[W.kstart:W.kend], a[W.kstart:W.kend])
    assert rel_err(got, a) <= 1e-15
    if case["order"] == 4:
        for t in (0, 1):
            w = case["w"].copy()
            w[W.kstart + 1:W.kend] = rng.uniform(-1, 1, w[W.kstart + 1:W.kend].shape)
            dw = _up(g, w)
            L.check(g.ctx, g.lib.mhh_ghost_cells_w(g.ctx, L.ptr(dw), t), "ghost_w")
            W.call("ref_ghost_cells_w", w, t)
            assert np.array_equal(dw.cpu().numpy(), w), t
    g.close(); W.close()


def test_set_bc(cuda_lib):
    from microhh_b200 import lib as L
    case = make_case(itot=16, jtot=12, ktot=8)
    g, W = _worlds(case)
    for sw, aval in ((0, 1.25), (1, 0.003), (2, -0.1)):
        ref = [np.full((W.jcells, W.icells), 7.) for _ in range(3)]
        dev = [_up(g, r) for r in ref]
        W.call("ref_set_bc", ref[0], ref[1], ref[2], sw, aval, 1e-5, 0.5)
        L.check(g.ctx, g.lib.mhh_set_bc(g.ctx, L.ptr(dev[0]), L.ptr(dev[1]), L.ptr(dev[2]), sw, C.c_double(aval), C.c_double(1e-5),
                                        C.c_double(0.5)), "set_bc")
        for r, d in zip(ref, dev):
            assert np.array_equal(d.cpu().numpy(), r), sw
    g.close(); W.close()


@pytest.mark.parametrize("rkorder", [3, 4])
def test_rk_substeps(cuda_lib, rkorder):
    """rk3/rk4, src/timeloop.cxx:250-334: a += cB*dt*at on the interior; at *= cA_next (0 after the last substep).
    The CPU reference zeroes/rescales ALL cells of the tendency (ghosts too): whole arrays are compared."""
    from microhh_b200 import lib as L
    case = make_case(itot=16, jtot=12, ktot=8, nscalars=2)
    g, W = _worlds(case)
    rng = np.random.default_rng(1)
    names = ["u", "v", "w", "s0", "s1"]
    a = {n: rng.uniform(-1, 1, W.shape) for n in names}
    at = {n: rng.uniform(-1, 1, W.shape) for n in names}
    da = {n: _up(g, a[n]) for n in names}
    dat = {n: _up(g, at[n]) for n in names}
    dt = 0.37
    for sub in range(3 if rkorder == 3 else 5):
        pa, pat = L.ptr_array([da[n] for n in names]), L.ptr_array([dat[n] for n in names])
        L.check(g.ctx, g.lib.mhh_rk_substep(g.ctx, rkorder, sub, C.c_double(dt), len(names), pa, pat), "rk")
        for n in names:
            W.call("ref_rk", rkorder, a[n], at[n], sub, dt)
            assert rel_err(da[n].cpu().numpy(), a[n]) <= 1e-15, (n, sub)
            assert rel_err(dat[n].cpu().numpy(), at[n]) <= 1e-15, (n, sub)
        import oracle.ref as R
        sd = R.lib().ref_rk_subdt(C.c_int(rkorder), C.c_double(dt), C.c_int(sub))
        assert g.lib.mhh_rk_sub_dt(rkorder, sub, C.c_double(dt)) == sd
    g.close(); W.close()


@pytest.mark.parametrize("kw", [dict(itot=32, jtot=24, ktot=16, alpha=1.2), dict(itot=64, jtot=1, ktot=12), dict(itot=24, jtot=20, ktot=10, order=4)],
                         ids=["3d_stretched", "2d", "o4"])
def test_reductions_and_diagnostics(cuda_lib, kw):
    from microhh_b200 import host as H
    from oracle import restate as RS
    case = make_case(**kw)
    g, W = _worlds(case)
    ops = H.Field3d_operators(g)
    th = case["s"][0]
    d = _up(g, th)
    prof = ops.calc_mean_profile(d)
    ref = np.zeros(W.kcells)
    W.call("ref_mean_profile", ref, th)
    assert np.abs(prof - ref).max() <= 1e-13 * np.abs(ref).max()
    assert abs(ops.calc_mean(d) - W.call("ref_mean", th)) <= 1e-13 * 300.
    assert abs(ops.calc_mean_2d(_up(g, th[3])) - W.call("ref_mean_2d", th[3].copy())) <= 1e-13 * 300.
    assert ops.calc_max(_up(g, case["u"])) == RS.calc_max(W, case["u"])
    if case["order"] == 2:
        f = H.Fields(g, slist=["s0"])
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
        f.sp["s0"].fld.copy_(d)
        u, v, w = case["u"], case["v"], case["w"]
        # Fields::check_momentum/tke/mass: the reference's own calc_momentum_2nd / calc_tke_2nd / calc_mass compiled in place
        # (oracle/ref_harness/k_fields.cxx), and the numpy restatement beside it
        for got, want, rs in ((f.check_momentum(), W.call("ref_check_momentum", u, v, w), RS.check_momentum(W, u, v, w)),
                              (f.check_tke(), W.call("ref_check_tke", u, v, w), RS.check_tke(W, u, v, w)),
                              (f.check_mass(), W.call("ref_check_mass", th), RS.check_mass(W, th))):
            assert abs(got - want) <= 1e-12 * max(abs(want), 1e-3), (got, want)
            assert abs(rs - want) <= 1e-12 * max(abs(want), 1e-3), (rs, want)
    g.close(); W.close()


def test_reduction_propagates_nan(cuda_lib):
    """The CFL guard of Model::print_status (src/model.cxx:832-836) needs a NaN/inf in the fields to reach the host."""
    from microhh_b200 import host as H
    case = make_case(itot=16, jtot=12, ktot=8)
    g, W = _worlds(case)
    f = H.Fields(g)
    for n in "uvw":
        f.mp[n].fld.copy_(_up(g, case[n]))
    f.mp["u"].fld[5, 6, 7] = float("nan")
    cfl = H.Advec.factory(g, f, "2").get_cfl(1.0)
    assert not (0. <= cfl < 10.)
    g.close(); W.close()


@pytest.mark.parametrize("order", [2, 4])
def test_thermo_dry(cuda_lib, order):
    from microhh_b200.model import Thermo
    from microhh_b200 import host as H
    case = make_case(itot=16, jtot=12, ktot=10, order=order, alpha=1.3)
    g, W = _worlds(case)
    f = H.Fields(g, slist=["th"])
    th = case["s"][0]
    f.sp["th"].fld.copy_(_up(g, th))
    rng = np.random.default_rng(2)
    thref = 300. + rng.uniform(0, 1, W.kcells)
    threfh = 300. + rng.uniform(0, 1, W.kcells)
    t = Thermo(g, f, "dry", thref=thref, threfh=threfh)
    wt0 = rng.uniform(-1, 1, W.shape)
    f.mt["w"].fld.copy_(_up(g, wt0))
    t.exec()
    ref = wt0.copy()
    W.call("ref_thermo_buoyancy_tend", ref, th, threfh)
    assert rel_err(f.mt["w"].fld.cpu().numpy(), ref) <= 1e-14
    if order == 2:
        n2 = t.get_thermo_field("N2").cpu().numpy()
        rn2 = np.zeros(W.shape)
        W.call("ref_thermo_N2", rn2, th, thref)
        assert rel_err(n2[W.kstart:W.kend], rn2[W.kstart:W.kend]) <= 1e-14
    g.close(); W.close()


@pytest.mark.parametrize("sm", [0, 1])
def test_fused_dry_viscosity_equals_two_step_path(cuda_lib, sm):
    """mhh_smag2_exec_viscosity_dry (N2 evaluated in-kernel from th) is bit-identical to thermo N2 + mhh_smag2_exec_viscosity."""
    from microhh_b200 import host as H
    from microhh_b200.model import Thermo
    case = make_case(itot=24, jtot=16, ktot=12, alpha=1.3, surface_model=sm, anelastic=True)
    g, W = _worlds(case)
    f = H.Fields(g, slist=["th"])
    for n in "uvw":
        f.mp[n].fld.copy_(_up(g, case[n]))
    f.sp["th"].fld.copy_(_up(g, case["s"][0]))
    rng = np.random.default_rng(2)
    t = Thermo(g, f, "dry", thref=300. + rng.uniform(0, 1, W.kcells), threfh=300. + rng.uniform(0, 1, W.kcells))
    b = H.Boundary(g, f, swboundary="surface" if sm else "default")
    d = H.Diff.factory(g, f, b, "smag2", cs=0.17, tPr=1. / 3.)
    P = {k: _up(g, v) for k, v in case["planes"].items()}
    d.exec_viscosity(t, surface=P)                  # fused
    fused = f.sd["evisc"].fld.cpu().numpy().copy()
    t.fuse_N2 = False
    f.sd["evisc"].fld.zero_()
    d.exec_viscosity(t, surface=P)                  # N2 field, then evisc
    assert np.array_equal(f.sd["evisc"].fld.cpu().numpy(), fused)
    g.close(); W.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sm", [0, 1])
@pytest.mark.parametrize("dims", [(24, 16, 12), (40, 10, 37), (33, 9, 70)], ids=["even", "ragged_tile", "two_chunks"])
def test_fused_viscosity_limits_equal_separate_reductions(cuda_lib, sm, dims):
    """mhh_smag2_exec_viscosity_dry_limits: the evisc field is bit-identical to mhh_smag2_exec_viscosity_dry and the two
    maxima equal mhh_advec_cfl / dt and mhh_smag2_dnmul run afterwards (same expressions, exact maxima); a NaN in the
    velocity still propagates into the CFL maximum (the reference's guard, src/model.cxx:832-836)."""
    from microhh_b200 import host as H
    from microhh_b200.model import Thermo
    itot, jtot, ktot = dims
    case = make_case(itot=itot, jtot=jtot, ktot=ktot, alpha=1.3, surface_model=sm, anelastic=True)
    g, W = _worlds(case)
    f = H.Fields(g, slist=["th"])
    b = H.Boundary(g, f, swboundary="surface" if sm else "default")
    for n in "uvw":
        f.mp[n].fld.copy_(_up(g, case[n]))
        b.boundary_cyclic.exec(f.mp[n].fld)
    f.sp["th"].fld.copy_(_up(g, case["s"][0]))
    rng = np.random.default_rng(2)
    t = Thermo(g, f, "dry", thref=300. + rng.uniform(0, 1, W.kcells), threfh=300. + rng.uniform(0, 1, W.kcells))
    a = H.Advec.factory(g, f, "2")
    d = H.Diff.factory(g, f, b, "smag2", cs=0.17, tPr=1. / 3.)
    P = {k: _up(g, v) for k, v in case["planes"].items()}
    lim = d.exec_viscosity(t, surface=P, limits=True)
    assert lim is not None
    fused = f.sd["evisc"].fld.cpu().numpy().copy()
    f.sd["evisc"].fld.zero_()
    assert d.exec_viscosity(t, surface=P) is None
    assert np.array_equal(f.sd["evisc"].fld.cpu().numpy(), fused)
    dt = 0.37
    assert lim[0] * dt == a.get_cfl(dt)
    assert lim[1] == d._dnmul()
    assert a.time_limit_from(lim[0] * dt, 10**9) == a.get_time_limit(10**9, dt)
    assert d.time_limit_from(lim[1], 10**9, dt) == d.get_time_limit(10**9, dt)
    f.mp["u"].fld[g.kstart + 1, g.jstart + 2, g.istart + 3] = float("nan")
    lim = d.exec_viscosity(t, surface=P, limits=True)
    assert np.isnan(lim[0]) and np.isnan(a.get_cfl(dt))
    g.close(); W.close()


@pytest.mark.parametrize("rkorder,substep", [(3, 0), (3, 1), (3, 2), (4, 3), (4, 4)])
def test_fused_pressure_tail_equals_output_then_rk(cuda_lib, rkorder, substep):
    """mhh_pres_exec_rk (Pres_2::output fused with Timeloop::exec for u,v,w) is bit-identical to mhh_pres_exec followed
    by mhh_rk_substep, including the cleared tendency ghost cells after the last substep."""
    import torch
    from microhh_b200 import host as H
    case = make_case(itot=24, jtot=16, ktot=12, alpha=1.3)
    g, W = _worlds(case)
    rng = np.random.default_rng(11)
    res = []
    for fused in (False, True):
        f = H.Fields(g, slist=["th"])
        b = H.Boundary(g, f)
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
            b.boundary_cyclic.exec(f.mp[n].fld)
        f.sp["th"].fld.copy_(_up(g, case["s"][0]))
        r = np.random.default_rng(11)
        for n in ("u", "v", "w", "th"):
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        pres = H.Pres.factory(g, f, "2")
        pres.set_values()
        tl = H.Timeloop(g, f, rkorder=rkorder, dt=0.73)
        tl.substep = substep
        done = pres.exec(tl.get_sub_time_step(), timeloop=tl if fused else None)
        assert done == fused
        tl.exec(skip=("u", "v", "w") if done else ())
        res.append({n: f.ap[n].fld.cpu().numpy().copy() for n in f.ap} | {n + "t": f.at[n].fld.cpu().numpy().copy() for n in f.at}
                   | {"p": f.sd["p"].fld.cpu().numpy().copy()})
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    g.close(); W.close()


@pytest.mark.parametrize("sm", [0, 1])
@pytest.mark.parametrize("thermo", ["0", "dry"])
@pytest.mark.parametrize("dims", [(24, 16, 12), (40, 10, 37), (33, 9, 5)], ids=["even", "ragged_tile", "odd"])
def test_fused_tendency_pass_equals_thermo_advec_diff(cuda_lib, sm, thermo, dims):
    """mhh_advec2_smag2_exec (buoyancy + advec_2 + smag2 diffusion in one pass per field) is bit-identical to
    Thermo::exec, Advec::exec, Diff::exec called one after the other on non-zero starting tendencies; the tile edges
    (imax, jmax not multiples of the tile; ktot not a multiple of the k chunk) are covered by the ragged / odd sizes."""
    from microhh_b200 import host as H
    from microhh_b200.model import Thermo
    itot, jtot, ktot = dims
    case = make_case(itot=itot, jtot=jtot, ktot=ktot, alpha=1.3, surface_model=sm, anelastic=True, nscalars=2)
    g, W = _worlds(case)
    res = []
    for fused in (False, True):
        f = H.Fields(g, slist=["th", "q"], visc=1e-3, svisc=2e-3)
        f.set_rhoref(case["rhoref"], case["rhorefh"])
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
        f.sp["th"].fld.copy_(_up(g, case["s"][0]))
        f.sp["q"].fld.copy_(_up(g, case["s"][1]))
        r = np.random.default_rng(3)
        for n in f.at:
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        for fld in list(f.mp.values()) + list(f.sp.values()):
            fld.flux_bot.copy_(_up(g, r.uniform(-1e-2, 1e-2, (W.jcells, W.icells))))
            fld.flux_top.copy_(_up(g, r.uniform(-1e-2, 1e-2, (W.jcells, W.icells))))
        rr = np.random.default_rng(2)
        t = Thermo(g, f, thermo, thref=300. + rr.uniform(0, 1, W.kcells), threfh=300. + rr.uniform(0, 1, W.kcells))
        b = H.Boundary(g, f, swboundary="surface" if sm else "default")
        a = H.Advec.factory(g, f, "2")
        d = H.Diff.factory(g, f, b, "smag2", cs=0.17, tPr=1. / 3.)
        f.sd["evisc"].fld.copy_(_up(g, r.uniform(0.1, 1., W.shape)))
        if fused:
            d.exec_with_advec2(f.sp["th"].fld if thermo == "dry" else None)
        else:
            t.exec(); a.exec(); d.exec()
        res.append({n: f.at[n].fld.cpu().numpy().copy() for n in f.at})
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    g.close(); W.close()


@pytest.mark.parametrize("dims", [(64, 16, 12), (512, 8, 3), (16, 6, 5)], ids=["n64", "n512", "n16"])
def test_fused_pressure_input_equals_input_then_fft(cuda_lib, dims):
    """mhh_pres_set_fused_input: Pres_2::input evaluated inside the forward x transform gives the same p and tendencies,
    bit for bit, as the separate input and transform stages."""
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    itot, jtot, ktot = dims
    case = make_case(itot=itot, jtot=jtot, ktot=ktot, alpha=1.3, anelastic=True)
    g, W = _worlds(case)
    res = []
    for fused in (0, 1):
        f = H.Fields(g, slist=[])
        f.set_rhoref(case["rhoref"], case["rhorefh"])
        b = H.Boundary(g, f)
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
            b.boundary_cyclic.exec(f.mp[n].fld)
        r = np.random.default_rng(11)
        for n in "uvw":
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        f.mt["w"].fld[:g.kstart + 1] = 0.
        f.mt["w"].fld[g.kend:] = 0.
        pres = H.Pres.factory(g, f, "2")
        pres.set_values()
        L.check(g.ctx, g.lib.mhh_pres_set_fused_input(g.ctx, fused), "mhh_pres_set_fused_input")
        pres.exec(0.37)
        res.append({n + "t": f.mt[n].fld.cpu().numpy().copy() for n in "uvw"} | {"p": f.sd["p"].fld.cpu().numpy().copy()})
    L.check(g.ctx, g.lib.mhh_pres_set_fused_input(g.ctx, 0), "mhh_pres_set_fused_input")
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    assert np.abs(res[0]["p"]).max() > 0.
    g.close(); W.close()


@pytest.mark.parametrize("rkorder,substep", [(3, 0), (3, 2), (4, 4)])
def test_fused_pressure_tail_order4(cuda_lib, rkorder, substep):
    """4th order: mhh_pres_exec_rk = Pres_4::exec, Boundary::set_ghost_cells_w(Normal), Timeloop::exec(u,v,w) — bit-identical
    to the three calls, including w's ghost rows (set from the not yet updated w, src/model.cxx:412)."""
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    case = make_case(itot=24, jtot=16, ktot=12, alpha=1.3, order=4)
    g, W = _worlds(case)
    res = []
    for fused in (False, True):
        f = H.Fields(g, slist=[])
        b = H.Boundary(g, f)
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
            b.boundary_cyclic.exec(f.mp[n].fld)
        r = np.random.default_rng(11)
        for n in "uvw":
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        f.mt["w"].fld[:g.kstart + 1] = 0.
        f.mt["w"].fld[g.kend:] = 0.
        pres = H.Pres.factory(g, f, "4")
        pres.set_values()
        tl = H.Timeloop(g, f, rkorder=rkorder, dt=0.73)
        tl.substep = substep
        b.set_ghost_cells_w(L.BCW_CONSERVATION)
        done = pres.exec(tl.get_sub_time_step(), timeloop=tl if fused else None)
        assert done == fused
        if not done:
            b.set_ghost_cells_w(L.BCW_NORMAL)
        tl.exec(skip=("u", "v", "w") if done else ())
        res.append({n: f.ap[n].fld.cpu().numpy().copy() for n in f.ap} | {n + "t": f.at[n].fld.cpu().numpy().copy() for n in f.at}
                   | {"p": f.sd["p"].fld.cpu().numpy().copy()})
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    g.close(); W.close()


@pytest.mark.parametrize("dims", [dict(itot=24, jtot=16, ktot=12), dict(itot=20, jtot=9, ktot=37), dict(itot=16, jtot=1, ktot=8)],
                         ids=["even", "ragged_2chunks", "2d"])
@pytest.mark.parametrize("nscal", [0, 1])
def test_fused_advec4_diff4_equals_separate_calls(cuda_lib, dims, nscal):
    """mhh_advec4_diff4_exec = Advec_4::exec, set_ghost_cells_w(Normal), Diff_4::exec (src/model.cxx:387-392): bit-identical
    tendencies and w ghost rows, on a stretched grid, with and without a scalar, over more than one k-chunk."""
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    case = make_case(alpha=1.3, order=4, nscalars=nscal, **dims)
    g, W = _worlds(case)
    res = []
    for fused in (False, True):
        f = H.Fields(g, slist=["s"] if nscal else [], visc=1e-2, svisc=2e-2)
        b = H.Boundary(g, f)
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
        if nscal:
            f.sp["s"].fld.copy_(_up(g, case["s"][0]))
        b.set_prognostic_cyclic_bcs()
        b.set_ghost_cells()
        r = np.random.default_rng(12)
        for n in f.at:
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        advec = H.Advec.factory(g, f, "4")
        diff = H.Diff.factory(g, f, b, "4")
        b.set_ghost_cells_w(L.BCW_CONSERVATION)
        if fused:
            diff.exec_with_advec4()
        else:
            advec.exec()
            b.set_ghost_cells_w(L.BCW_NORMAL)
            diff.exec()
        res.append({n + "t": f.at[n].fld.cpu().numpy().copy() for n in f.at} | {"w": f.mp["w"].fld.cpu().numpy().copy()})
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    assert np.abs(res[0]["ut"]).max() > 1e-2
    g.close(); W.close()


@pytest.mark.parametrize("rkorder,substep", [(3, 0), (3, 2), (4, 2), (4, 4)])
@pytest.mark.parametrize("swbuffer", [False, True])
def test_fused_scalar_tail_equals_buffer_then_rk(cuda_lib, rkorder, substep, swbuffer):
    """mhh_advec2_smag2_exec_rk: the scalars' Buffer term and RK update inside their tendency pass are bit-identical to
    Advec/Diff::exec, Buffer::exec, Timeloop::exec — interior of the scalar (returned in the swapped second buffer), the whole
    tendency array (zeroed with its ghost cells after the last substep), u, v, w tendencies untouched by the tail."""
    from microhh_b200 import host as H
    from microhh_b200.model import Thermo
    case = make_case(itot=40, jtot=10, ktot=37, alpha=1.3, surface_model=1, anelastic=True, nscalars=2, zsize=3200.)
    g, W = _worlds(case)
    res = []
    for fused in (False, True):
        f = H.Fields(g, slist=["th", "q"], visc=1e-3, svisc=2e-3)
        f.set_rhoref(case["rhoref"], case["rhorefh"])
        for n in "uvw":
            f.mp[n].fld.copy_(_up(g, case[n]))
        f.sp["th"].fld.copy_(_up(g, case["s"][0]))
        f.sp["q"].fld.copy_(_up(g, case["s"][1]))
        r = np.random.default_rng(3)
        for n in f.at:
            f.at[n].fld.copy_(_up(g, r.uniform(-1e-2, 1e-2, W.shape)))
        for fld in list(f.mp.values()) + list(f.sp.values()):
            fld.flux_bot.copy_(_up(g, r.uniform(-1e-2, 1e-2, (W.jcells, W.icells))))
            fld.flux_top.copy_(_up(g, r.uniform(-1e-2, 1e-2, (W.jcells, W.icells))))
        t = Thermo(g, f, "dry", thref=np.full(W.kcells, 300.), threfh=np.full(W.kcells, 300.))
        b = H.Boundary(g, f, swboundary="surface")
        a = H.Advec.factory(g, f, "2")
        d = H.Diff.factory(g, f, b, "smag2", cs=0.17, tPr=1. / 3.)
        f.sd["evisc"].fld.copy_(_up(g, r.uniform(0.1, 1., W.shape)))
        profs = {"th": 300. + r.uniform(0, 10, W.ktot), "q": r.uniform(0, 1, W.ktot), "u": r.uniform(-1, 1, W.ktot), "v": r.uniform(-1, 1, W.ktot)}
        buf = H.Buffer(g, f, swbuffer=swbuffer, zstart=2000., sigma=0.00223, beta=2., profs=profs)
        tl = H.Timeloop(g, f, rkorder=rkorder, dt=0.73)
        tl.substep = substep
        if fused:
            d.exec_with_advec2(f.sp["th"].fld, buffer=buf, timeloop=tl)
            buf.exec(skip=("th", "q"))
            tl.exec(skip=("th", "q"))
        else:
            t.exec(); a.exec(); d.exec()
            buf.exec()
            tl.exec()
        inner = (slice(W.kstart, W.kend), slice(W.jstart, W.jend), slice(W.istart, W.iend))
        res.append({n: f.ap[n].fld.cpu().numpy()[inner].copy() for n in f.ap} | {n + "t": f.at[n].fld.cpu().numpy().copy() for n in f.at})
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k]), k
    assert not np.array_equal(res[0]["th"], case["s"][0][W.kstart:W.kend, W.jstart:W.jend, W.istart:W.iend])
    g.close(); W.close()


@pytest.mark.parametrize("dims", [(64, 20, 70), (100, 12, 40)], ids=["2chunks", "ragged"])
def test_tma_and_cp_async_plane_staging_are_bit_identical(cuda_lib, dims):
    """The tile kernels stage their planes with TMA bulk tensor copies (default) or per-thread cp.async (mhh_set_tma(ctx, 0)):
    same values in shared memory, so evisc, all tendencies, the updated scalar and the dt maxima must agree bit for bit."""
    from tests.helpers import run_substep_gpu_product
    import microhh_b200.host as H
    case = make_case(itot=dims[0], jtot=dims[1], ktot=dims[2], surface_model=1, anelastic=True, alpha=1.3)
    out = []
    for tma in (True, False):
        os_env = __import__("os").environ
        if not tma:
            os_env["MHH_TMA"] = "0"
        try:
            out.append(run_substep_gpu_product(case, scalar_rk=False, limits=True))
            out.append(run_substep_gpu_product(case, scalar_rk=True, limits=False))
        finally:
            os_env.pop("MHH_TMA", None)
    for a, b in ((out[0], out[2]), (out[1], out[3])):
        for n in a:
            if n == "scalars":
                assert a[n] == b[n]
            else:
                assert np.array_equal(a[n], b[n]), n
    # and TMA is really the active path on this (even icells) grid
    g = H.Grid(dims[0], dims[1], dims[2], 100., 100., 100., case["z"] * 100. / 3200., swspatialorder="2", gc=(1, 1, 1))
    assert g.lib.mhh_tma_active(g.ctx) == 1
    g.close()


def test_fused_force_tail_equals_separate_buffer_and_coriolis(cuda_lib):
    """mhh_advec2_smag2_set_force_tail: Buffer::exec and calc_coriolis_2nd of u,v,w inside the fused momentum pass against the
    same terms as separate passes (mhh_buffer_exec, mhh_force_coriolis) — the same operations in the same order: bit-identical
    fields, dt history and status over 6 adaptive steps of a gabls1-like case (MO surface layer, sponge, geostrophic wind)."""
    import torch
    from microhh_b200.model import Model
    from tests.test_gpu_multistep import les_case
    ini, z, init, _ = les_case("gabls1_stable", n=(40, 24, 72))
    n = ini["grid"]["ktot"]
    ini["buffer"] = dict(swbuffer=True, zstart=250., sigma=0.00223, beta=2.)
    ini["force"] = dict(swlspres="geo", fc=1.39e-4)
    profs = dict(u=np.full(n, 8.), v=np.zeros(n), th=265. + 0.01 * np.clip(z - 100., 0., None), u_geo=np.full(n, 7.5), v_geo=np.full(n, 0.3))
    dev = torch.device("cuda:0")
    out = []
    for fuse in (True, False):
        m = Model(dict(ini, b200=dict(fuse_force=fuse)), z, device=dev, profs=profs, mo=True)
        assert m.fuse_force == fuse
        f = m.fields
        for nm in f.ap:
            f.ap[nm].fld.copy_(torch.from_numpy(np.ascontiguousarray(init[nm])).to(dev))
        dts = []
        for _ in range(6):
            m.step()
            dts.append(m.timeloop.idt)
        out.append(({nm: f.ap[nm].fld.cpu().numpy().copy() for nm in f.ap}, dts, m.print_status()))
        m.grid.close()
    assert out[0][1] == out[1][1]
    for nm in out[0][0]:
        assert np.array_equal(out[0][0][nm], out[1][0][nm]), nm
    assert out[0][2] == out[1][2]
    assert np.abs(out[0][0]["v"]).max() > 1e-4


@pytest.mark.parametrize("kw", [dict(itot=32, jtot=24, ktot=16, alpha=1.2), dict(itot=20, jtot=12, ktot=10, order=4)], ids=["o2", "o4"])
def test_stats_profiles(cuda_lib, kw):
    """mhh_stats_mean / _moment / _cov against the reference's calc_mean, calc_moment (powers 2, 3, 4) and calc_cov
    (src/stats.cxx:264-398 compiled in place, oracle/ref_harness/k_stats.cxx) over the default mask."""
    from microhh_b200 import host as H
    case = make_case(**kw)
    g, W = _worlds(case)
    st = H.Stats(g)
    w, th = case["w"], case["s"][0]
    dw, dth = _up(g, w), _up(g, th)
    ks, ke = W.kstart, W.kend + 1
    for a, d in ((w, dw), (th, dth)):
        ref = np.zeros(W.kcells)
        W.call("ref_stats_mean", ref, a)
        got = st.calc_mean(d)
        assert np.abs(got[ks:ke] - ref[ks:ke]).max() <= 1e-13 * max(np.abs(ref).max(), 1e-300)
        for power in (2, 3, 4):
            r2 = np.zeros(W.kcells)
            W.call("ref_stats_moment", r2, a, ref, 0.25, power)
            g2 = st.calc_moment(d, ref, power, offset=0.25)
            assert np.abs(g2[ks:ke] - r2[ks:ke]).max() <= 1e-12 * np.abs(r2).max(), power
    mw, mth = np.zeros(W.kcells), np.zeros(W.kcells)
    W.call("ref_stats_mean", mw, w); W.call("ref_stats_mean", mth, th)
    rc = np.zeros(W.kcells)
    W.call("ref_stats_cov", rc, w, mw, 0., 1, th, mth, 0., 1)
    gc = st.calc_cov(dw, mw, dth, mth)
    assert np.abs(gc[ks:ke] - rc[ks:ke]).max() <= 1e-12 * np.abs(rc).max()
    rc2 = np.zeros(W.kcells)
    W.call("ref_stats_cov", rc2, w, mw, 0.1, 2, th, mth, -0.2, 1)
    gc2 = st.calc_cov(dw, mw, dth, mth, pow1=2, pow2=1, offset1=0.1, offset2=-0.2)
    assert np.abs(gc2[ks:ke] - rc2[ks:ke]).max() <= 1e-12 * np.abs(rc2).max()
    g.close(); W.close()
