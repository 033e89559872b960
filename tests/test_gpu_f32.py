"""The `_f32` symbol set (include/mhh_b200_f32.h; the reference's -DUSESP build, CMakeLists.txt:29-34, main/microhh.cxx:40-46):
the same kernels compiled with float as the working type (csrc/sys.h), driven through the same host classes with
precision="f32", against the DOUBLE-precision reference sources (oracle/_ref) — i.e. against the values a single-precision
run approximates.  Tolerances are those of float arithmetic (eps = 6e-8): per-call tendencies 1e-4 of the field's largest
value (differences of nearly equal products), updated fields 1e-5, and north_star's 1e-5 for the mean profiles of a short run."""
import numpy as np
import pytest

from tests.helpers import make_case, rel_err, run_substep_gpu_product, run_substep_oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kw", [dict(itot=64, jtot=40, ktot=70, surface_model=1), dict(itot=48, jtot=32, ktot=24, anelastic=True, alpha=1.3)],
                         ids=["surface_2chunks", "anelastic_stretched"])
def test_f32_product_substep(cuda_lib, kw):
    case = make_case(**kw)
    want = run_substep_oracle(case, buoyancy=True)
    got = run_substep_gpu_product(case, scalar_rk=False, limits=True, precision="f32")
    ks, ke = case["ks"], case["ke"]
    errs = {n: rel_err(got[n][ks:ke, 1:-1, 1:-1].astype(np.float64), want[n][ks:ke, 1:-1, 1:-1])
            for n in ("evisc", "ut", "vt", "wt", "st0", "u", "v", "w", "s0", "p")}
    print(errs, got["scalars"], want["scalars"])
    for n in ("ut", "vt", "wt", "st0"):
        assert errs[n] <= 1e-4, (n, errs[n])
    # evisc goes through N2 = g/th0 * dth/dz: th ~ 300 K carries 3e-5 K of float rounding against differences of ~0.1 K
    assert errs["evisc"] <= 5e-4, errs["evisc"]
    for n in ("u", "v", "w", "s0"):
        assert errs[n] <= 1e-5, (n, errs[n])
    assert errs["p"] <= 2e-3, errs["p"]            # the Poisson solve amplifies the float rounding of its right-hand side
    for k in ("cfl", "dnmul"):
        assert abs(got["scalars"][k] - want["scalars"][k]) <= 1e-5 * abs(want["scalars"][k]), k
    assert got["scalars"]["div"] <= 1e-4             # float projection: divergence at the float rounding level of u/dx
    assert got["u"].dtype == np.float32


def test_f32_les_steps_profiles(cuda_lib):
    """20 adaptive RK3 steps of the gabls1-like LES in single precision (MO surface layer, sponge, Coriolis on the device)
    against the double-precision reference stepper: mean profiles within north_star's 1e-5, second moments within 5e-3."""
    from tests.test_gpu_multistep import les_case, profiles, run_both
    ini, z, init, _ = les_case("gabls1_stable", n=32)
    n = 32
    ini["buffer"] = dict(swbuffer=True, zstart=300., sigma=0.00223, beta=2.)
    ini["force"] = dict(swlspres="geo", fc=1.39e-4)
    profs = dict(u=np.full(n, 8.), v=np.zeros(n), th=265. + 0.01 * np.clip(z - 100., 0., None), u_geo=np.full(n, 8.), v_geo=np.zeros(n))
    res = run_both(ini, z, init, 20, profs=profs, mo=True, precision="f32")
    pg, pw = profiles(res["got"], *res["dims"]), profiles(res["want"], *res["dims"])
    scale = {k: max(np.abs(pw[k]).max(), np.sqrt(pw[k + "_2"].max()) if k + "_2" in pw else 0., 1e-300) for k in pw}
    errs = {k: float(np.abs(pg[k] - pw[k]).max() / scale[k]) for k in pw}
    print("f32 profile errors", errs, "dt history", res["dts"][-3:], "status", res["status"])
    for k in ("u", "v", "th"):
        assert errs[k] <= 1e-5, (k, errs[k])
    assert errs["w"] <= 1e-4, errs["w"]
    # second moments: th fluctuates by ~0.05 K on 265 K, i.e. each value carries 3e-4 of its fluctuation as float rounding
    for k in ("u_2", "v_2", "w_2", "th_2", "wth"):
        assert errs[k] <= 5e-3, (k, errs[k])
    ddt = max(abs(a - b) / b for a, b in res["dts"])
    assert ddt <= 1e-4, ddt
    assert res["status"]["div"] <= 1e-4
