"""Shared test scaffolding: seeded synthetic cases in the reference layout, one RK substep of the
LES configuration through (a) the CUDA path via the C ABI and (b) the oracle (reference sources
compiled in place, oracle/_ref)."""
from __future__ import annotations

import numpy as np

GRAV = 9.81


def stretched_z(ktot, zsize, alpha=0.0):
    """Uniform (alpha=0) or tanh-stretched (cases/moser180/moser180_input.py:19-24 style) cell centres."""
    if alpha == 0.0:
        return (np.arange(ktot) + 0.5) * (zsize / ktot)
    zh = np.zeros(ktot + 1)
    for k in range(ktot + 1):
        eta = -1.0 + 2.0 * k / ktot
        zh[k] = 0.5 * zsize * (1.0 + np.tanh(alpha * eta) / np.tanh(alpha))
    return 0.5 * (zh[1:] + zh[:-1])


def make_case(itot=32, jtot=32, ktot=32, xsize=3200., ysize=3200., zsize=3200., seed=666, order=2, gc=None,
              alpha=0.0, anelastic=False, surface_model=0, nscalars=1, mbc="freeslip"):
    """Synthetic inputs in the style of kernel_tuner/helpers.py:10-48, physically scaled."""
    rng = np.random.default_rng(seed)
    if gc is None:
        gc = (1, 1, 1) if order == 2 else (3, 3, 3)
    z = stretched_z(ktot, zsize, alpha)
    kc, jc, ic = ktot + 2 * gc[2], jtot + 2 * gc[1], itot + 2 * gc[0]
    shape = (kc, jc, ic)
    ks, ke = gc[2], gc[2] + ktot
    u = rng.uniform(-1, 1, shape)
    v = rng.uniform(-1, 1, shape)
    w = rng.uniform(-1, 1, shape)
    w[:ks + 1] = 0.
    w[ke:] = 0.
    zfull = np.zeros(kc)
    zfull[ks:ke] = z
    scal = []
    for n in range(nscalars):
        th = 300. + 0.003 * zfull[:, None, None] + rng.uniform(-0.05, 0.05, shape)
        scal.append(th)
    if anelastic:
        zz = np.concatenate([[-z[0]] * ks, z, [2 * zsize - z[-1]] * (kc - ke)])
        zzh = np.zeros(kc)
        zzh[ks + 1:ke] = 0.5 * (z[1:] + z[:-1])
        zzh[ke:] = zsize
        rhoref = np.exp(-zz / 8000.)
        rhorefh = np.exp(-zzh / 8000.)
    else:
        rhoref = np.ones(kc)
        rhorefh = np.ones(kc)
    planes = {}
    if surface_model:
        planes["z0m"] = np.full((jc, ic), 0.1) + rng.uniform(0, 0.01, (jc, ic))
        planes["dudz"] = rng.uniform(0.01, 0.1, (jc, ic))
        planes["dvdz"] = rng.uniform(-0.05, 0.05, (jc, ic))
        planes["dbdz"] = rng.uniform(-1e-4, 1e-4, (jc, ic))
        for nm in ("ufluxbot", "vfluxbot", "sfluxbot"):
            planes[nm] = rng.uniform(-0.01, 0.01, (jc, ic))
        for nm in ("ufluxtop", "vfluxtop", "sfluxtop"):
            planes[nm] = np.zeros((jc, ic))
    return dict(itot=itot, jtot=jtot, ktot=ktot, xsize=xsize, ysize=ysize, zsize=zsize, z=z, order=order, gc=gc,
                u=u, v=v, w=w, s=scal, rhoref=rhoref, rhorefh=rhorefh, thref=300., planes=planes,
                surface_model=surface_model, mbc=mbc, visc=1e-5, svisc=1e-5, cs=0.23, tPr=1. / 3., dt=2.0, substep=0,
                ks=ks, ke=ke)


def calc_N2(case, th, dzi):
    """thermo_dry calc_N2 (src/thermo_dry.cxx:66-79) with a constant thref — input to smag2, same array for both paths."""
    N2 = np.zeros_like(th)
    ks, ke = case["ks"], case["ke"]
    N2[ks:ke] = GRAV / case["thref"] * 0.5 * (th[ks + 1:ke + 1] - th[ks - 1:ke - 1]) * dzi[ks:ke, None, None]
    return N2


def rel_err(got, ref):
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-300))


# ------------------------------------------------------------------------------------------------
def run_substep_oracle(case, buoyancy=False):
    """One LES RK3 substep with the reference's own code (oracle/_ref).  buoyancy: Thermo_dry::exec on wt before the
    advection (src/model.cxx:366), as the product path's fused tendency pass includes it."""
    from oracle.ref import RefWorld
    c = case
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], order=c["order"], gc=c["gc"],
                 rhoref=c["rhoref"], rhorefh=c["rhorefh"])
    u, v, w = c["u"].copy(), c["v"].copy(), c["w"].copy()
    s = [x.copy() for x in c["s"]]
    sm = c["surface_model"]
    P = c["planes"]
    zero2 = np.zeros((W.jcells, W.icells))
    ut, vt, wt = np.zeros(W.shape), np.zeros(W.shape), np.zeros(W.shape)
    st = [np.zeros(W.shape) for _ in s]
    evisc, p = np.zeros(W.shape), np.zeros(W.shape)
    # boundary->set_prognostic_cyclic_bcs, set_ghost_cells (src/model.cxx:346-348)
    for f in (u, v, w, *s):
        W.call("ref_cyclic", f, 2)
    mbc = 0 if c.get("mbc", "freeslip") == "noslip" else 1      # Boundary_type: Dirichlet (value 0) / Neumann (gradient 0)
    for f in (u, v):
        W.call("ref_ghost_cells", f, mbc, mbc, zero2.copy(), zero2.copy(), zero2.copy(), zero2.copy())
    for f in s:
        W.call("ref_ghost_cells", f, 1, 1, zero2.copy(), zero2.copy(), np.full_like(zero2, 0.003), np.full_like(zero2, 0.003))
    # diff->exec_viscosity
    N2 = calc_N2(c, s[0], W.dzi) if s else None
    W.call("ref_smag2_strain2", sm, evisc, u, v, w, P.get("dudz"), P.get("dvdz"))
    if s:
        W.call("ref_smag2_evisc", sm, evisc, u, v, w, N2, P.get("dbdz"), P.get("z0m"), c["cs"], c["tPr"])
    else:
        W.call("ref_smag2_evisc_neutral", sm, evisc, u, v, w, P.get("z0m"), c["cs"], c["visc"])
    cfl = W.call("ref_advec2_cfl", u, v, w, c["dt"])
    dnmul = W.call("ref_smag2_dnmul", evisc, c["tPr"])
    if buoyancy and s:
        W.call("ref_thermo_buoyancy_tend", wt, s[0], np.full(W.kcells, c["thref"]))
    # advec, diff
    W.call("ref_advec2_uvw", ut, vt, wt, u, v, w)
    for a, at in zip(s, st):
        W.call("ref_advec2_s", at, a, u, v, w)
    W.call("ref_smag2_diff_uvw", sm, ut, vt, wt, u, v, w, evisc, P.get("ufluxbot", zero2), P.get("ufluxtop", zero2),
           P.get("vfluxbot", zero2), P.get("vfluxtop", zero2), c["visc"])
    for a, at in zip(s, st):
        W.call("ref_smag2_diff_c", sm, at, a, evisc, P.get("sfluxbot", zero2), P.get("sfluxtop", zero2), c["tPr"], c["svisc"])
    tend = dict(ut=ut.copy(), vt=vt.copy(), wt=wt.copy(), **{f"st{n}": x.copy() for n, x in enumerate(st)})
    # pres, rk
    import ctypes as C
    from oracle.ref import lib as _reflib
    sub_dt = _reflib().ref_rk_subdt(C.c_int(3), C.c_double(c["dt"]), C.c_int(c["substep"]))
    W.call("ref_pres2_exec", p, u, v, w, ut, vt, wt, sub_dt)
    for a, at in zip((u, v, w, *s), (ut, vt, wt, *st)):
        W.call("ref_rk", 3, a, at, c["substep"], c["dt"])
    div = None
    uu, vv, ww = u.copy(), v.copy(), w.copy()
    for f in (uu, vv, ww):
        W.call("ref_cyclic", f, 2)
    div = W.call("ref_pres2_divergence", uu, vv, ww)
    out = dict(u=u, v=v, w=w, p=p, evisc=evisc, ut_after=ut, **{f"s{n}": x for n, x in enumerate(s)}, **tend)
    out["scalars"] = dict(cfl=cfl, dnmul=dnmul, div=div, sub_dt=sub_dt)
    W.close()
    return out


def run_substep_gpu(case):
    """The same substep through the CUDA path (host mirror classes over the C ABI)."""
    import torch
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    c = case
    dev = torch.device("cuda:0")
    g = H.Grid(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], swspatialorder=str(c["order"]),
               gc=c["gc"], device=dev)
    names = [f"s{n}" for n in range(len(c["s"]))]
    f = H.Fields(g, slist=names, visc=c["visc"], svisc=c["svisc"])
    f.set_rhoref(c["rhoref"], c["rhorefh"])
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    f.mp["u"].fld.copy_(up(c["u"])); f.mp["v"].fld.copy_(up(c["v"])); f.mp["w"].fld.copy_(up(c["w"]))
    for n, a in zip(names, c["s"]):
        f.sp[n].fld.copy_(up(a))
    sm = c["surface_model"]
    b = H.Boundary(g, f, mbcbot=c.get("mbc", "freeslip"), mbctop=c.get("mbc", "freeslip"),
                   sbc={n: dict(sbot=0.003, stop=0.003, sbcbot="neumann", sbctop="neumann") for n in names},
                   swboundary="surface" if sm else "default")
    b.set_values()
    P = {k: up(v) for k, v in c["planes"].items()}
    if sm:
        f.mp["u"].flux_bot.copy_(P["ufluxbot"]); f.mp["v"].flux_bot.copy_(P["vfluxbot"])
        f.mp["u"].flux_top.copy_(P["ufluxtop"]); f.mp["v"].flux_top.copy_(P["vfluxtop"])
        for n in names:
            f.sp[n].flux_bot.copy_(P["sfluxbot"]); f.sp[n].flux_top.copy_(P["sfluxtop"])
    advec = H.Advec.factory(g, f, "2")
    diff = H.Diff.factory(g, f, b, "smag2", cs=c["cs"], tPr=c["tPr"])
    pres = H.Pres.factory(g, f, "2")
    pres.set_values()
    tl = H.Timeloop(g, f, rkorder=3, dt=c["dt"], adaptivestep=False)
    tl.substep = c["substep"]

    b.set_prognostic_cyclic_bcs()
    b.set_ghost_cells()

    class _Thermo:
        def get_switch(self_inner):
            return "dry" if names else "0"

        def get_thermo_field(self_inner, name):
            th = f.sp[names[0]].fld.cpu().numpy()
            return up(calc_N2(c, th, g.dzi))

    diff.exec_viscosity(_Thermo(), surface=P)
    cfl = advec.get_cfl(c["dt"])
    dnmul = diff.get_dn(1.0)
    advec.exec()
    diff.exec()
    tend = dict(ut=f.mt["u"].fld.cpu().numpy().copy(), vt=f.mt["v"].fld.cpu().numpy().copy(), wt=f.mt["w"].fld.cpu().numpy().copy(),
                **{f"st{n}": f.st[nm].fld.cpu().numpy().copy() for n, nm in enumerate(names)})
    sub_dt = tl.get_sub_time_step()
    pres.exec(sub_dt)
    tl.exec()
    # divergence of the updated velocities (after a halo fill on copies)
    keep = {n: f.mp[n].fld.clone() for n in "uvw"}
    for n in "uvw":
        b.boundary_cyclic.exec(f.mp[n].fld)
    div = pres.check_divergence()
    for n in "uvw":
        f.mp[n].fld.copy_(keep[n])
    out = dict(u=f.mp["u"].fld.cpu().numpy(), v=f.mp["v"].fld.cpu().numpy(), w=f.mp["w"].fld.cpu().numpy(),
               p=f.sd["p"].fld.cpu().numpy(), evisc=f.sd["evisc"].fld.cpu().numpy(), ut_after=f.mt["u"].fld.cpu().numpy(),
               **{f"s{n}": f.sp[nm].fld.cpu().numpy() for n, nm in enumerate(names)}, **tend)
    out["scalars"] = dict(cfl=cfl, dnmul=dnmul, div=div, sub_dt=sub_dt)
    g.close()
    return out


def run_substep_gpu_product(case, scalar_rk=False, limits=True, precision="f64"):
    """The same substep through the entry points the product path (microhh_b200.model.Model) uses: the fused viscosity pass
    with N2 from th and both dt maxima (mhh_smag2_exec_viscosity_dry[_limits]), the one-pass tendencies
    (mhh_advec2_smag2_exec[_rk]: buoyancy + advec_2 + diff_smag2, optionally the scalars' RK update) and the pressure tail
    fused with the RK update of u,v,w (mhh_pres_exec_rk).  The first scalar is th (thermo dry, thref = case["thref"])."""
    import torch
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    from microhh_b200.model import Thermo
    c = case
    dev = torch.device("cuda:0")
    g = H.Grid(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"], swspatialorder=str(c["order"]),
               gc=c["gc"], device=dev, precision=precision)
    names = ["th"] + [f"s{n}" for n in range(1, len(c["s"]))]
    f = H.Fields(g, slist=names, visc=c["visc"], svisc=c["svisc"])
    f.set_rhoref(c["rhoref"], c["rhorefh"])
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=g.npdtype)).to(dev)
    f.mp["u"].fld.copy_(up(c["u"])); f.mp["v"].fld.copy_(up(c["v"])); f.mp["w"].fld.copy_(up(c["w"]))
    for n, a in zip(names, c["s"]):
        f.sp[n].fld.copy_(up(a))
    sm = c["surface_model"]
    b = H.Boundary(g, f, mbcbot=c.get("mbc", "freeslip"), mbctop=c.get("mbc", "freeslip"),
                   sbc={n: dict(sbot=0.003, stop=0.003, sbcbot="neumann", sbctop="neumann") for n in names},
                   swboundary="surface" if sm else "default")
    b.set_values()
    P = {k: up(v) for k, v in c["planes"].items()}
    if sm:
        f.mp["u"].flux_bot.copy_(P["ufluxbot"]); f.mp["v"].flux_bot.copy_(P["vfluxbot"])
        f.mp["u"].flux_top.copy_(P["ufluxtop"]); f.mp["v"].flux_top.copy_(P["vfluxtop"])
        for n in names:
            f.sp[n].flux_bot.copy_(P["sfluxbot"]); f.sp[n].flux_top.copy_(P["sfluxtop"])
    advec = H.Advec.factory(g, f, "2")
    diff = H.Diff.factory(g, f, b, "smag2", cs=c["cs"], tPr=c["tPr"])
    pres = H.Pres.factory(g, f, "2")
    pres.set_values()
    thermo = Thermo(g, f, "dry", thref0=c["thref"])
    tl = H.Timeloop(g, f, rkorder=3, dt=c["dt"], adaptivestep=False)
    tl.substep = c["substep"]
    b.set_prognostic_cyclic_bcs()
    b.set_ghost_cells()
    lim = diff.exec_viscosity(thermo, surface=P, limits=limits)
    if lim is not None:
        cfl, dnmul = lim[0] * c["dt"], lim[1]
    else:
        cfl, dnmul = advec.get_cfl(c["dt"]), diff.get_dn(1.0)
    evisc = f.sd["evisc"].fld.cpu().numpy().copy()
    diff.exec_with_advec2(f.sp["th"].fld, timeloop=tl if scalar_rk else None)
    tend = {}
    if not scalar_rk:
        tend = dict(ut=f.mt["u"].fld.cpu().numpy().copy(), vt=f.mt["v"].fld.cpu().numpy().copy(), wt=f.mt["w"].fld.cpu().numpy().copy(),
                    **{f"st{n}": f.st[nm].fld.cpu().numpy().copy() for n, nm in enumerate(names)})
    sub_dt = tl.get_sub_time_step()
    fused = pres.exec(sub_dt, timeloop=tl)
    assert fused
    tl.exec(skip=("u", "v", "w") + (tuple(names) if scalar_rk else ()))
    keep = {n: f.mp[n].fld.clone() for n in "uvw"}
    for n in "uvw":
        b.boundary_cyclic.exec(f.mp[n].fld)
    div = pres.check_divergence()
    for n in "uvw":
        f.mp[n].fld.copy_(keep[n])
    out = dict(u=f.mp["u"].fld.cpu().numpy(), v=f.mp["v"].fld.cpu().numpy(), w=f.mp["w"].fld.cpu().numpy(),
               p=f.sd["p"].fld.cpu().numpy(), evisc=evisc, ut_after=f.mt["u"].fld.cpu().numpy(),
               **{f"s{n}": f.sp[nm].fld.cpu().numpy() for n, nm in enumerate(names)}, **tend)
    out["scalars"] = dict(cfl=cfl, dnmul=dnmul, div=div, sub_dt=sub_dt)
    g.close()
    return out
