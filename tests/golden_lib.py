"""Golden-vector plumbing shared by tests/golden/make_golden.py (generator, runs where the oracle is built) and the tests.

Every vector is an OUTPUT of the reference's own code (oracle/_ref = /root/reference sources compiled in place) on inputs
regenerated from a seed by tests.helpers.make_case; the .npz files under tests/golden/ hold outputs only."""
from __future__ import annotations

import os

import numpy as np

from tests.helpers import make_case, run_substep_oracle, stretched_z

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

KERNEL_CASES = {
    "o2": dict(itot=8, jtot=6, ktot=6, anelastic=True, alpha=1.2, surface_model=1, seed=101),
    "o4": dict(itot=8, jtot=6, ktot=6, order=4, alpha=1.5, seed=102),
    "2i5": dict(itot=8, jtot=6, ktot=8, gc=(3, 3, 2), anelastic=True, seed=103),
}


def oracle_kernels(tag):
    """Per-call outputs of the reference kernels for KERNEL_CASES[tag]: {name: array}."""
    from oracle.ref import RefWorld
    c = make_case(**KERNEL_CASES[tag])
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"],
                 rhoref=c["rhoref"], rhorefh=c["rhorefh"])
    u, v, w, s = c["u"], c["v"], c["w"], c["s"][0]
    z = lambda: np.zeros(W.shape)
    out = {}
    for name in ("dz", "dzh", "dzi", "dzhi", "dzi4", "dzhi4", "z", "zh"):
        out["grid_" + name] = getattr(W, name).copy()
    ut, vt, wt, st = z(), z(), z(), z()
    if tag == "o2":
        P = c["planes"]
        W.call("ref_advec2_uvw", ut, vt, wt, u, v, w)
        W.call("ref_advec2_s", st, s, u, v, w)
        out.update(advec_ut=ut.copy(), advec_vt=vt.copy(), advec_wt=wt.copy(), advec_st=st.copy())
        out["cfl"] = np.array(W.call("ref_advec2_cfl", u, v, w, 2.0))
        d = [z() for _ in range(4)]
        W.call("ref_diff2_c", d[0], u, 1e-2); W.call("ref_diff2_c", d[1], v, 1e-2); W.call("ref_diff2_w", d[2], w, 1e-2)
        out.update(diff2_ut=d[0], diff2_vt=d[1], diff2_wt=d[2])
        ev = z()
        N2 = z()
        W.call("ref_thermo_N2", N2, s, np.full(W.kcells, 300.))
        W.call("ref_smag2_strain2", 1, ev, u, v, w, P["dudz"], P["dvdz"])
        out["strain2"] = ev.copy()
        W.call("ref_smag2_evisc", 1, ev, u, v, w, N2, P["dbdz"], P["z0m"], c["cs"], c["tPr"])
        out["evisc"] = ev.copy()
        out["N2"] = N2
        out["dnmul"] = np.array(W.call("ref_smag2_dnmul", ev, c["tPr"]))
        e = [z() for _ in range(4)]
        W.call("ref_smag2_diff_uvw", 1, e[0], e[1], e[2], u, v, w, ev, P["ufluxbot"], P["ufluxtop"], P["vfluxbot"], P["vfluxtop"], c["visc"])
        W.call("ref_smag2_diff_c", 1, e[3], s, ev, P["sfluxbot"], P["sfluxtop"], c["tPr"], c["svisc"])
        out.update(smag2_ut=e[0], smag2_vt=e[1], smag2_wt=e[2], smag2_st=e[3])
        pexec, pdiv = "ref_pres2_exec", "ref_pres2_divergence"
    elif tag == "o4":
        W.call("ref_advec4_uvw", ut, vt, wt, u, v, w)
        W.call("ref_advec4_s", st, s, u, v, w)
        out.update(advec_ut=ut.copy(), advec_vt=vt.copy(), advec_wt=wt.copy(), advec_st=st.copy())
        out["cfl"] = np.array(W.call("ref_advec4_cfl", u, v, w, 2.0))
        d = [z() for _ in range(3)]
        W.call("ref_diff4_c", d[0], u, 1e-2); W.call("ref_diff4_c", d[1], v, 1e-2); W.call("ref_diff4_w", d[2], w, 1e-2)
        out.update(diff4_ut=d[0], diff4_vt=d[1], diff4_wt=d[2])
        pexec, pdiv = "ref_pres4_exec", "ref_pres4_divergence"
    else:
        W.call("ref_advec2i5_uvw", ut, vt, wt, u, v, w)
        W.call("ref_advec2i5_s", st, s, u, v, w, 0)
        sl = z()
        W.call("ref_advec2i5_s", sl, s, u, v, w, 1)
        out.update(advec_ut=ut.copy(), advec_vt=vt.copy(), advec_wt=wt.copy(), advec_st=st.copy(), advec_st_lim=sl)
        out["cfl"] = np.array(W.call("ref_advec2i5_cfl", u, v, w, 2.0))
        pexec, pdiv = "ref_pres2_exec", "ref_pres2_divergence"
    # pressure solve on top of the advection tendencies, then one RK3 substep
    p = z()
    uu, vv, ww = u.copy(), v.copy(), w.copy()
    W.call(pexec, p, uu, vv, ww, ut, vt, wt, 0.5)
    out.update(pres_p=p, pres_ut=ut.copy(), pres_vt=vt.copy(), pres_wt=wt.copy())
    W.call("ref_rk", 3, uu, ut, 0, 1.5)
    out.update(rk_u=uu, rk_ut=ut)
    g = s.copy()
    zero2 = np.zeros((W.jcells, W.icells))
    W.call("ref_cyclic", g, 2)
    W.call("ref_ghost_cells", g, 0, 1, np.full_like(zero2, 299.), zero2.copy(), zero2.copy(), np.full_like(zero2, 0.003))
    out["bc_s"] = g
    W.close()
    return out


def oracle_les_substep():
    c = make_case(itot=16, jtot=12, ktot=10, surface_model=1, anelastic=True, alpha=1.3, seed=104)
    o = run_substep_oracle(c)
    sc = o.pop("scalars")
    o.update({"scalar_" + k: np.array(v) for k, v in sc.items()})
    return c, o


DNS4_INI = {"grid": dict(itot=16, jtot=12, ktot=12, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="4"),
            "advec": dict(swadvec="4", cflmax=1.0), "diff": dict(swdiff="4", dnmax=0.4), "pres": dict(swpres="4"),
            "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-3),
            "time": dict(rkorder=3, dt=0.01, dtmax=0.1, adaptivestep=True)}


def dns4_init():
    g = DNS4_INI["grid"]
    z = stretched_z(g["ktot"], 2., alpha=2.0)
    rng = np.random.default_rng(105)
    shape = (g["ktot"] + 6, g["jtot"] + 6, g["itot"] + 6)
    zf = np.zeros(g["ktot"] + 6); zf[3:-3] = z
    init = dict(u=(zf * (2. - zf))[:, None, None] * 0.11 + 0.01 * rng.uniform(-1, 1, shape), v=0.01 * rng.uniform(-1, 1, shape),
                w=0.01 * rng.uniform(-1, 1, shape))
    init["w"][:4] = 0.; init["w"][-3:] = 0.
    return z, init


def oracle_dns4_steps(nsteps=3):
    from oracle.les import OracleLES
    z, init = dns4_init()
    g = DNS4_INI["grid"]
    o = OracleLES(g["itot"], g["jtot"], g["ktot"], g["xsize"], g["ysize"], g["zsize"], z, init, swadvec="4", swdiff="4", order=4,
                  visc=1e-3, svisc=0., swthermo="0", mbcbot="noslip", mbctop="noslip", rkorder=3, dt=0.01, dtmax=0.1, dnmax=0.4)
    for _ in range(nsteps):
        o.step()
    out = {n: o.a[n].copy() for n in "uvw"}
    out["p"] = o.p.copy()
    out["idt"] = np.array(o.idt, dtype=np.int64)
    o.close()
    return out


def oracle_r02():
    """Outputs of the reference for what round 2 added to the path: Diff_smag2 without a thermo scalar (calc_evisc_neutral, both
    wall treatments), Stats profiles, dry buoyancy, sponge layer, Coriolis force — small seeded cases."""
    from oracle.ref import RefWorld
    out = {}
    for tag, kw in (("mason", dict(surface_model=1)), ("vandriest", dict(surface_model=0, mbc="noslip", xsize=100., ysize=100., zsize=100.))):
        c = make_case(itot=12, jtot=8, ktot=10, nscalars=0, seed=201, **kw)
        if tag == "vandriest":
            c["visc"] = 1e-3
            c["u"] *= 0.05; c["v"] *= 0.05
        o = run_substep_oracle(c)
        for k in ("evisc", "ut", "vt", "wt", "p"):
            out[f"neutral_{tag}_{k}"] = o[k]
    c = make_case(itot=12, jtot=8, ktot=10, zsize=3200., seed=202)
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"],
                 rhoref=c["rhoref"], rhorefh=c["rhorefh"])
    u, v, w, th = c["u"], c["v"], c["w"], c["s"][0]
    prof = lambda: np.zeros(W.kcells)
    mw, mth = prof(), prof()
    W.call("ref_stats_mean", mw, w); W.call("ref_stats_mean", mth, th)
    m2, m3, cov = prof(), prof(), prof()
    W.call("ref_stats_moment", m2, th, mth, 0., 2)
    W.call("ref_stats_moment", m3, w, mw, 0.25, 3)
    W.call("ref_stats_cov", cov, w, mw, 0., 1, th, mth, 0., 1)
    out.update(stats_mean_w=mw, stats_mean_th=mth, stats_th2=m2, stats_w3=m3, stats_wth=cov)
    rng = np.random.default_rng(203)
    threfh = 300. + 0.003 * W.zh
    bt = np.zeros(W.shape)
    W.call("ref_thermo_buoyancy_tend", bt, th, threfh)
    out["buoyancy_wt"] = bt
    abuf = prof(); abuf[W.kstart:W.kend] = rng.uniform(-1, 1, W.ktot)
    for name, fld, is_w in (("u", u, 0), ("w", w, 1)):
        t = np.zeros(W.shape)
        W.call("ref_buffer", t, fld, abuf, 2300., 0.00223, 2., is_w)
        out["buffer_" + name + "t"] = t
    g2 = [a.copy() for a in (u, v)]
    for a in g2:
        W.call("ref_cyclic", a, 2)
    ug, vg = prof(), prof()
    ug[W.kstart:W.kend] = rng.uniform(5, 10, W.ktot); vg[W.kstart:W.kend] = rng.uniform(-1, 1, W.ktot)
    cut, cvt = np.zeros(W.shape), np.zeros(W.shape)
    W.call("ref_force_coriolis", cut, cvt, g2[0], g2[1], ug, vg, 1.39e-4, 0.5, -0.25)
    out.update(coriolis_ut=cut, coriolis_vt=cvt)
    W.close()
    return out


def load(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz")))
