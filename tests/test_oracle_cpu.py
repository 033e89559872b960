"""CPU (no GPU): pins the oracle — the reference's own sources compiled in place (oracle/_ref) plus oracle/restate.py —
against (1) the committed golden vectors, (2) the reference's own known-answer tests restated: Taylor-Green convergence
(cases/taylorgreen/taylorgreen_test.py) and conservation (cases/conservation/conservation_test.py), (3) closed-form answers
(projection leaves a divergence-free field, FFT round trip, uniform-field diagnostics).  Skipped where oracle/_ref is not
built (it is built by __graft_entry__.build() wherever /root/reference exists and travels to the GPU box)."""
import numpy as np
import pytest

from oracle import ref as R
from tests import golden_lib as G
from tests.helpers import make_case

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built")


def _close(a, b, tol=1e-13):
    return np.abs(a - b).max() <= tol * max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("tag", list(G.KERNEL_CASES))
def test_oracle_reproduces_golden_kernels(tag):
    want = G.load("kernels_" + tag)
    got = G.oracle_kernels(tag)
    assert sorted(got) == sorted(want)
    for k in want:
        assert _close(got[k], want[k]), k


def test_oracle_reproduces_golden_steps():
    _, got = G.oracle_les_substep()
    want = G.load("les_substep")
    for k in want:
        assert _close(got[k], want[k], 1e-12), k
    got = G.oracle_dns4_steps()
    want = G.load("dns4_steps")
    for k in "uvwp":
        assert _close(got[k], want[k], 1e-11), k
    assert int(got["idt"]) == int(want["idt"])


def test_oracle_reproduces_golden_round2_features():
    """calc_evisc_neutral (Mason / van Driest), Stats profiles, dry buoyancy, sponge layer, Coriolis: the reference's outputs
    on seeded cases against the committed vectors (tests/golden/kernels_r02.npz)."""
    want = G.load("kernels_r02")
    got = G.oracle_r02()
    assert sorted(got) == sorted(want)
    for k in want:
        assert _close(got[k], want[k]), k
        assert np.abs(want[k]).max() > 0., k          # no vector is trivially zero


# ---------------------------------------------------------------------------------------------------------------------
def oracle_taylorgreen(itot, ktot, order):
    from oracle.les import OracleLES
    visc = (8. * np.pi ** 2 * 100.) ** (-1.)
    gc = 1 if order == 2 else 3
    dz, dx = 0.5 / ktot, 1. / itot
    z = np.linspace(0.5 * dz, 0.5 - 0.5 * dz, ktot)
    shape = (ktot + 2 * gc, 1 + 2 * gc, itot + 2 * gc)
    x = (np.arange(shape[2]) - gc + 0.5) * dx
    xh = (np.arange(shape[2]) - gc) * dx
    zc = (np.arange(shape[0]) - gc + 0.5) * dz
    zh = (np.arange(shape[0]) - gc) * dz
    u0, w0 = np.zeros(shape), np.zeros(shape)
    ks, ke, js, is_, ie = gc, gc + ktot, gc, gc, gc + itot
    u0[ks:ke, js, is_:ie] = np.sin(2. * np.pi * xh[None, is_:ie]) * np.cos(np.pi * zc[ks:ke, None] / 0.5)
    w0[ks:ke, js, is_:ie] = -np.cos(2. * np.pi * x[None, is_:ie]) * np.sin(np.pi * zh[ks:ke, None] / 0.5)
    o = OracleLES(itot, 1, ktot, 1., 1., 0.5, z, dict(u=u0, v=np.zeros(shape), w=w0), swadvec=str(order), swdiff=str(order),
                  order=order, visc=visc, svisc=0., swthermo="0", mbcbot="freeslip", mbctop="freeslip", rkorder=3, dt=0.0025,
                  dtmax=0.01, cflmax=0.5, dnmax=0.2, endtime=1.)
    n = 0
    while o.itime < o.iendtime:
        o.step(); n += 1
        assert n < 5000
    dec = np.exp(-8. * np.pi ** 2 * visc)
    ru = np.sin(2. * np.pi * xh[None, is_:ie]) * np.cos(2. * np.pi * zc[ks:ke, None]) * dec
    rw = -np.cos(2. * np.pi * x[None, is_:ie]) * np.sin(2. * np.pi * zh[ks:ke, None]) * dec
    rp = (0.25 * (np.cos(4. * np.pi * x[None, is_:ie]) + np.cos(4. * np.pi * zc[ks:ke, None])) - 0.25) * dec ** 2
    err = {v: float(dx * dz * np.abs(a[ks:ke, js, is_:ie] - r).sum()) for v, a, r in (("u", o.a["u"], ru), ("w", o.a["w"], rw), ("p", o.p, rp))}
    div = o.divergence()
    o.close()
    return err, div


@pytest.mark.parametrize("order", [2, 4])
def test_oracle_taylorgreen_known_answer(order):
    """The reference's own known-answer test (cases/taylorgreen/taylorgreen_test.py:101-131): L1 error vs the analytic
    Taylor-Green solution at t = 1 converges with the scheme's order (pins advec+diff+pres+RK3+BCs end to end)."""
    e16, d16 = oracle_taylorgreen(16, 8, order)
    e32, d32 = oracle_taylorgreen(32, 16, order)
    print(e16, e32)
    assert d16 < 1e-12 and d32 < 1e-12
    for v in ("u", "w", "p"):
        conv = np.log(e16[v] / e32[v]) / np.log(2.)
        assert conv > order - 0.3, (v, conv)
    assert e32["u"] < (1e-5 if order == 2 else 2e-7)


def test_oracle_advection_conserves_mass_and_momentum():
    """cases/conservation: flux-form advection in a periodic box with w = 0 at the walls neither creates nor destroys
    volume-integrated scalar (advec_2, advec_2i5 with and without the limiter) or momentum (advec_2).  advec_4 is not
    discretely conservative at its biased wall rows (that is what the out-of-scope advec_4m is for), so it is not listed."""
    from oracle.ref import RefWorld
    for tag, kw in (("2", dict()), ("2i5", dict(gc=(3, 3, 2)))):
        c = make_case(itot=12, jtot=10, ktot=8, seed=5, **kw)
        W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"])
        u, v, w, s = c["u"], c["v"], c["w"], c["s"][0] - 300.
        zero2 = np.zeros((W.jcells, W.icells))
        for a in (u, v, w, s):
            W.call("ref_cyclic", a, 2)
        for a in (u, v, s):
            W.call("ref_ghost_cells", a, 1, 1, zero2.copy(), zero2.copy(), zero2.copy(), zero2.copy())
        if c["order"] == 4:
            W.call("ref_ghost_cells_w", w, 1)
        dz = W.dz[W.kstart:W.kend, None, None]
        I = (slice(W.kstart, W.kend), slice(W.jstart, W.jend), slice(W.istart, W.iend))
        variants = [("s", 0)] + ([("s_lim", 1)] if tag == "2i5" else [])
        for name, lim in variants:
            st = np.zeros(W.shape)
            if tag == "2":
                W.call("ref_advec2_s", st, s, u, v, w)
            elif tag == "4":
                W.call("ref_advec4_s", st, s, u, v, w)
            else:
                W.call("ref_advec2i5_s", st, s, u, v, w, lim)
            assert abs((st[I] * dz).sum()) <= 1e-12 * (np.abs(st[I]) * dz).sum(), (tag, name)
        if tag == "2":
            ut, vt, wt = np.zeros(W.shape), np.zeros(W.shape), np.zeros(W.shape)
            W.call("ref_advec2_uvw", ut, vt, wt, u, v, w)
            for t in (ut, vt):
                assert abs((t[I] * dz).sum()) <= 1e-12 * (np.abs(t[I]) * dz).sum()
        W.close()


def test_oracle_projection_and_fft_roundtrip():
    from oracle.ref import RefWorld
    for order in (2, 4):
        c = make_case(itot=12, jtot=10, ktot=8, order=order, alpha=1.4, seed=9)
        W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=order, gc=c["gc"])
        x = np.random.default_rng(0).uniform(-1, 1, W.shape)
        n = W.itot * W.jtot * W.ktot
        y = x.copy()
        out = np.zeros(W.shape)
        W.call("ref_fft_forward", y)
        W.call("ref_fft_backward", y, out)
        assert _close(out.ravel()[:n], x.ravel()[:n], 1e-13)
        u, v, w = c["u"], c["v"], c["w"]
        for a in (u, v, w):
            W.call("ref_cyclic", a, 2)
        ut, vt, wt, p = (np.zeros(W.shape) for _ in range(4))
        dt = 0.7
        if order == 4:
            W.call("ref_ghost_cells_w", w, 1)
        W.call("ref_pres%d_exec" % order, p, u, v, w, ut, vt, wt, dt)
        un, vn, wn = u + dt * ut, v + dt * vt, w + dt * wt
        for a in (un, vn, wn):
            W.call("ref_cyclic", a, 2)
        if order == 4:
            W.call("ref_ghost_cells_w", wn, 1)
        assert W.call("ref_pres%d_divergence" % order, un, vn, wn) < 1e-12
        W.close()


def test_restatement_closed_forms():
    from oracle import restate as RS
    from oracle.ref import RefWorld
    c = make_case(itot=8, jtot=6, ktot=10, alpha=1.7)
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"])
    one = np.ones(W.shape)
    assert abs(RS.check_momentum(W, 1. * one, 2. * one, 0. * one) - 3.) < 1e-13
    assert abs(RS.check_tke(W, 1. * one, 2. * one, 0. * one) - 2.5) < 1e-13
    assert abs(RS.check_mass(W, 7. * one) - 7.) < 1e-13
    assert abs(W.call("ref_mean", 7. * one) - 7.) < 1e-13
    W.close()


def test_fields_diagnostics_reference_vs_restatement():
    """Fields::check_momentum/tke/mass (src/fields.cxx:144-234) compiled in place (k_fields.cxx: Boost and NetCDF shimmed) pin
    the numpy restatement on random fields, 2nd- and 4th-order ghost depths, stretched z."""
    from oracle import restate as RS
    from oracle.ref import RefWorld
    for kw in (dict(itot=12, jtot=10, ktot=14, alpha=1.4), dict(itot=16, jtot=1, ktot=8), dict(itot=10, jtot=8, ktot=12, order=4, alpha=1.2)):
        c = make_case(**kw)
        W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"])
        u, v, w, th = c["u"], c["v"], c["w"], c["s"][0]
        for ref, rs in ((W.call("ref_check_momentum", u, v, w), RS.check_momentum(W, u, v, w)),
                        (W.call("ref_check_tke", u, v, w), RS.check_tke(W, u, v, w)), (W.call("ref_check_mass", th), RS.check_mass(W, th))):
            assert abs(ref - rs) <= 1e-13 * max(abs(ref), 1e-3), (kw, ref, rs)
        W.close()
