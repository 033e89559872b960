"""GPU parity at the sizes the headline number runs at: the tile kernels march 64-level chunks (MHH_TILE_KCHUNK) over
32-wide x tiles and re-derive every carried face flux / shear / interpolation sum at each chunk start, so these cases cross
several chunk seams and x tiles, with ragged edges, a stretched anelastic column, surface model on and off — through the
fused entry points of the product path (mhh_smag2_exec_viscosity_dry[_limits], mhh_advec2_smag2_exec[_rk],
mhh_pres_exec_rk) as well as the operator-by-operator calls, against the reference's own sources (oracle/_ref).
Also: the BASELINE grids (drycblles 64^3, moser180 256x192x128), Diff_smag2's neutral branch (calc_evisc_neutral,
src/diff_smag2.cxx:158-252; never taken when a thermo scalar exists), and a neutral LES.
Tolerances: north_star's 1e-12 relative for per-call tendencies and updated fields, 1e-9 for profiles."""
import numpy as np
import pytest

from tests.helpers import make_case, rel_err, run_substep_gpu, run_substep_gpu_product, run_substep_oracle

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _compare(case, got, want, names, tol=TOL):
    ks, ke = case["ks"], case["ke"]
    gi, gj = case["gc"][0], case["gc"][1]
    errs = {n: rel_err(got[n][ks:ke, gj:-gj, gi:-gi], want[n][ks:ke, gj:-gj, gi:-gi]) for n in names if n != "p"}
    if "p" in names:
        errs["p"] = rel_err(got["p"][ks:ke, gj:-gj, gi:-gi], want["p"][ks:ke, gj:-gj, gi:-gi])
    print(errs, got["scalars"], want["scalars"])
    for n, e in errs.items():
        assert e <= (1e-10 if n == "p" else tol), (n, e)      # p: different (exact-arithmetic-equivalent) FFT algorithm
    for k in ("cfl", "dnmul", "sub_dt"):
        assert abs(got["scalars"][k] - want["scalars"][k]) <= 1e-13 * abs(want["scalars"][k]), (k, got["scalars"][k], want["scalars"][k])
    assert got["scalars"]["div"] <= max(10 * want["scalars"]["div"], 1e-13)


SEAM_CASES = [
    dict(itot=100, jtot=44, ktot=200, anelastic=True, alpha=1.2, surface_model=1),      # 4 ragged x tiles, 4 k chunks (64,64,64,8)
    dict(itot=100, jtot=44, ktot=200, anelastic=True, alpha=1.2, surface_model=0),
    dict(itot=96, jtot=39, ktot=129, surface_model=1),                                   # a 1-level last chunk, odd jtot
    dict(itot=72, jtot=20, ktot=70, surface_model=1, gc=(2, 2, 1)),                      # two ghost cells (cp.async staging: box starts are not 16-byte aligned)
]
SEAM_IDS = ["100x44x200_surf_anel_stretched", "100x44x200_nosurf_anel_stretched", "96x39x129_surf", "72x20x70_gc221"]


@pytest.mark.parametrize("kw", SEAM_CASES, ids=SEAM_IDS)
def test_product_entries_across_chunk_seams(cuda_lib, kw):
    """Fused viscosity (+ dt maxima), fused tendency pass (buoyancy + advec_2 + smag2), fused pressure tail + RK: per-call
    tendencies, evisc, p and the updated fields against the reference on a grid with >= 3 k chunks and >= 3 x tiles."""
    case = make_case(**kw)
    want = run_substep_oracle(case, buoyancy=True)
    got = run_substep_gpu_product(case, scalar_rk=False, limits=True)
    _compare(case, got, want, ("evisc", "ut", "vt", "wt", "st0", "u", "v", "w", "s0", "p"))
    # the variant of the scalar pass that carries the RK update (mhh_advec2_smag2_exec_rk), and the non-LIM viscosity entry
    got2 = run_substep_gpu_product(case, scalar_rk=True, limits=False)
    _compare(case, got2, want, ("evisc", "u", "v", "w", "s0", "p"))
    gi, gj = case["gc"][0], case["gc"][1]
    for n in ("u", "v", "w", "s0", "p", "evisc"):
        assert np.array_equal(got[n][case["ks"]:case["ke"], gj:-gj, gi:-gi], got2[n][case["ks"]:case["ke"], gj:-gj, gi:-gi]), n


@pytest.mark.parametrize("kw", SEAM_CASES[:2], ids=SEAM_IDS[:2])
def test_separate_entries_across_chunk_seams(cuda_lib, kw):
    """The operator-by-operator entries (mhh_smag2_exec_viscosity, mhh_advec_exec, mhh_smag2_exec, mhh_pres_exec,
    mhh_rk_substep) on the same seam-crossing grids."""
    case = make_case(**kw)
    _compare(case, run_substep_gpu(case), run_substep_oracle(case), ("evisc", "ut", "vt", "wt", "st0", "u", "v", "w", "s0", "p"))


@pytest.mark.parametrize("n", [64, 128], ids=["drycblles_64", "cube_128"])
def test_full_substep_at_baseline_grids(cuda_lib, n):
    """One full substep at drycblles' 64^3 (BASELINE configs[0]) and at 128^3, product entries, surface model on."""
    case = make_case(itot=n, jtot=n, ktot=n, surface_model=1)
    want = run_substep_oracle(case, buoyancy=True)
    _compare(case, run_substep_gpu_product(case), want, ("evisc", "ut", "vt", "wt", "st0", "u", "v", "w", "s0", "p"))


def test_three_les_steps_128x128x160(cuda_lib):
    """Three adaptive RK3 steps of the product path (Model) against the reference stepper at 128 x 128 x 160 (3 k chunks,
    4 x tiles): fields to 1e-10 (rounding differences grow over 9 substeps), profiles to 1e-11, dt history identical."""
    from tests.test_gpu_multistep import check_profiles, les_case, run_both
    ini, z, init, surface = les_case("gabls1_stable", n=(128, 128, 160))
    res = run_both(ini, z, init, 3, surface=surface)
    check_profiles(res, 1e-11, 2)
    ks, ke, js, je, is_, ie = res["dims"]
    for n in res["want"]:
        e = rel_err(res["got"][n][ks:ke, js:je, is_:ie], res["want"][n][ks:ke, js:je, is_:ie])
        print(n, e)
        assert e <= 1e-10, (n, e)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)


def test_moser180_grid_order4_step(cuda_lib):
    """BASELINE configs[3] at its real size 256 x 192 x 128 (tanh-stretched z, advec_4, diff_4, pres_4, RK3, uflux forcing):
    one full step (3 substeps) of the product path (fused advec4+diff4, fused pres4 tail + RK) against the reference."""
    from tests.helpers import stretched_z
    from tests.test_gpu_multistep import check_profiles, run_both
    itot, jtot, ktot = 256, 192, 128
    z = stretched_z(ktot, 2., alpha=2.0)
    rng = np.random.default_rng(4)
    shape = (ktot + 6, jtot + 6, itot + 6)
    zf = np.zeros(ktot + 6); zf[3:-3] = z
    init = dict(u=(zf * (2. - zf))[:, None, None] * 0.11 + 0.01 * rng.uniform(-1, 1, shape), v=0.01 * rng.uniform(-1, 1, shape),
                w=0.01 * rng.uniform(-1, 1, shape))
    init["w"][:4] = 0.; init["w"][-3:] = 0.
    ini = {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="4"),
           "advec": dict(swadvec="4", cflmax=1.0), "diff": dict(swdiff="4", dnmax=0.4), "pres": dict(swpres="4"),
           "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-5),
           "force": dict(swlspres="uflux", uflux=0.0282),
           "time": dict(rkorder=3, dt=0.001, dtmax=0.1, adaptivestep=True)}
    res = run_both(ini, z, init, 1)
    check_profiles(res, 1e-11, 0)
    ks, ke, js, je, is_, ie = res["dims"]
    for n in res["want"]:
        e = rel_err(res["got"][n][ks:ke, js:je, is_:ie], res["want"][n][ks:ke, js:je, is_:ie])
        print(n, e)
        assert e <= 1e-11, (n, e)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)


# ---- Diff_smag2 without a thermo scalar: calc_evisc_neutral (src/diff_smag2.cxx:158-252) ---------------------------------
@pytest.mark.parametrize("kw", [
    dict(itot=40, jtot=24, ktot=70, surface_model=1),                                   # Mason n = 1 (:243-247)
    dict(itot=40, jtot=24, ktot=70, surface_model=0),                                   # van Driest (:199-211) + mirrored rows (:214-226)
    dict(itot=36, jtot=20, ktot=24, surface_model=1, anelastic=True, alpha=1.5),
    dict(itot=36, jtot=20, ktot=24, surface_model=0, anelastic=True, alpha=1.5),
], ids=["mason_2chunks", "vandriest_2chunks", "mason_stretched", "vandriest_stretched"])
def test_evisc_neutral_matches_reference(cuda_lib, kw):
    if kw["surface_model"]:
        case = make_case(nscalars=0, **kw)
    else:
        # resolved walls: no-slip (u_tau comes from the wall shear), and a domain / viscosity at which the van Driest factor
        # is neither 0 nor saturated near the walls
        case = make_case(nscalars=0, mbc="noslip", xsize=100., ysize=100., zsize=100., **kw)
        case["visc"] = 1e-3
        case["u"] *= 0.05; case["v"] *= 0.05
    got = run_substep_gpu(case)
    want = run_substep_oracle(case)
    ks, ke = case["ks"], case["ke"]
    # evisc including its ghost cells: cyclic halo everywhere, mirrored wall rows when the walls are resolved
    lo, hi = (ks, ke) if kw["surface_model"] else (ks - 1, ke + 1)
    e_ev = rel_err(got["evisc"][lo:hi], want["evisc"][lo:hi])
    print("evisc", e_ev, got["scalars"], want["scalars"])
    assert e_ev <= TOL
    if not kw["surface_model"]:     # the damping really acts: evisc at the wall-adjacent level is far below its mid-channel value
        assert want["evisc"][ks, 1:-1, 1:-1].mean() < 0.5 * want["evisc"][(ks + ke) // 2, 1:-1, 1:-1].mean()
    _compare(case, got, want, ("ut", "vt", "wt", "u", "v", "w", "p"))


@pytest.mark.parametrize("surface_model", [1, 0], ids=["mason", "vandriest"])
def test_neutral_les_20_steps(cuda_lib, surface_model):
    """20 adaptive RK3 steps of a neutral LES (swthermo=0: exec_viscosity takes calc_evisc_neutral every substep) through
    Model against the reference stepper: profiles at 1e-9, same dt history."""
    from tests.test_gpu_multistep import check_profiles, run_both
    n = 32
    rng = np.random.default_rng(11)
    shape = (n + 2, n + 2, n + 2)
    s2 = shape[1:]
    z = (np.arange(n) + 0.5) * (1000. / n)
    zf = np.zeros(n + 2); zf[1:-1] = z
    init = dict(u=(3. + 2. * zf / 1000.)[:, None, None] + 0.3 * rng.uniform(-1, 1, shape), v=0.3 * rng.uniform(-1, 1, shape),
                w=0.1 * rng.uniform(-1, 1, shape))
    init["w"][:2] = 0.; init["w"][-1:] = 0.
    surface = None
    if surface_model:
        surface = dict(z0m=np.full(s2, 0.1), dudz=rng.uniform(0.02, 0.05, s2), dvdz=rng.uniform(-0.01, 0.01, s2),
                       ufluxbot=rng.uniform(-0.02, 0., s2), vfluxbot=rng.uniform(-0.002, 0.002, s2),
                       ufluxtop=np.zeros(s2), vfluxtop=np.zeros(s2))
    ini = {"grid": dict(itot=n, jtot=n, ktot=n, xsize=1000., ysize=1000., zsize=1000., swspatialorder="2"),
           "advec": dict(swadvec="2", cflmax=1.0), "diff": dict(swdiff="smag2", cs=0.17, dnmax=0.3), "pres": dict(swpres="2"),
           "boundary": dict(mbcbot="noslip", mbctop="freeslip" if surface_model else "noslip",
                            swboundary="surface" if surface_model else "default"),
           "fields": dict(visc=1e-5 if surface_model else 1e-2), "time": dict(rkorder=3, dt=1.0, dtmax=20., adaptivestep=True)}
    res = run_both(ini, z, init, 20, surface=surface)
    check_profiles(res, 1e-9, 100)
    assert res["status"]["div"] <= max(10 * res["odiv"], 1e-12)
