"""The reference's own CUDA kernels (src/*.cu compiled in place for sm_100a: oracle/_ref/liboracle_refcu.so, measurement
infrastructure only) timed on the same B200, same fields and same launch shapes the reference uses, beside this repository's
kernels on one RK substep of the bench case:  python tools/bench_refcu.py [n]  (default 512^3).  CUDA events, 5 repeats after
2 warm-ups, inputs larger than L2."""
import ctypes as C, json, os, sys
import numpy as np, torch
sys.path.insert(0, ".")
import bench as B
from microhh_b200 import lib as L
from microhh_b200.model import Model

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
ini = B.les_ini(n, n, n)
z = (np.arange(n) + 0.5) * (ini["grid"]["zsize"] / n)
m = Model(ini, z, device=dev, profs=B.les_profs(z), mo=True)
g, f = m.grid, m.fields
host = B.synthetic_fields(g.shape, g.kstart, g.kend, z, seed=666)
for k, v in host.items():
    f.ap[k].fld.copy_(torch.from_numpy(v).to(dev))
for _ in range(2):
    m.step()
torch.cuda.synchronize()
g.profile_enable(True)
for _ in range(3):
    m.step()
torch.cuda.synchronize()
mine = {k: v[0] / v[1] for k, v in g.profile_report().items()}
g.profile_enable(False)

R = C.CDLL(os.path.join("oracle", "_ref", "liboracle_refcu.so"))
class RG(C.Structure):
    _fields_ = [(k, C.c_int) for k in "itot jtot ktot icells jcells kcells ijcells istart jstart kstart iend jend kend igc jgc kgc".split()] + \
               [(k, C.c_double) for k in ("dx", "dy", "zsize")]
rg = RG(g.itot, g.jtot, g.ktot, g.icells, g.jcells, g.kcells, g.ijcells, g.istart, g.jstart, g.kstart, g.iend, g.jend, g.kend,
        g.igc, g.jgc, g.kgc, g.dx, g.dy, g.zsize)
P = lambda t: C.c_void_p(t.data_ptr())
met = lambda name: torch.from_numpy(g.metric(name)).to(dev)
dzi, dzhi, dz, zc = met("dzi"), met("dzhi"), met("dz"), met("z")
one = torch.ones(g.kcells, dtype=torch.float64, device=dev)
mlen = torch.from_numpy(0.1 * (g.dx * g.dy * g.dz) ** (1. / 3.)).to(dev)
u, v, w, th = (f.ap[k].fld for k in ("u", "v", "w", "th"))
ut, vt, wt, tht = (f.at[k].fld for k in ("u", "v", "w", "th"))
for t in (ut, vt, wt, tht):
    t.zero_()
evisc, p = f.sd["evisc"].fld, f.sd["p"].fld
tmp1, tmp2 = torch.zeros_like(u), torch.zeros_like(u)
b = m.boundary
bm = torch.zeros(g.itot + g.jtot + 4 * g.kcells, dtype=torch.float64, device=dev) + 1.
stream = C.c_void_p(0)
def timeit(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ref = {}
ref["calc_strain2_g"] = timeit(lambda: R.refcu_smag2_strain2(C.byref(rg), P(evisc), P(u), P(v), P(w), P(b.dudz_mo), P(b.dvdz_mo), P(dzi), P(dzhi), stream))
ref["evisc_g"] = timeit(lambda: R.refcu_smag2_evisc(C.byref(rg), P(evisc), P(tmp1), P(b.dbdz_mo), P(mlen), P(b.z0m), P(zc), C.c_double(3.), stream))
ref["advec_uvw_g"] = timeit(lambda: R.refcu_advec2_uvw(C.byref(rg), P(ut), P(vt), P(wt), P(u), P(v), P(w), P(one), P(one), P(dzi), P(dzhi), stream))
ref["advec_s_g"] = timeit(lambda: R.refcu_advec2_s(C.byref(rg), P(tht), P(th), P(u), P(v), P(w), P(one), P(one), P(dzi), stream))
ref["diff_uvw_g"] = timeit(lambda: R.refcu_smag2_diff_uvw(C.byref(rg), P(ut), P(vt), P(wt), P(evisc), P(u), P(v), P(w), P(f.mp["u"].flux_bot), P(f.mp["u"].flux_top),
                                                          P(f.mp["v"].flux_bot), P(f.mp["v"].flux_top), P(dzi), P(dzhi), P(one), P(one), C.c_double(1e-5), stream))
ref["diff_c_g"] = timeit(lambda: R.refcu_smag2_diff_c(C.byref(rg), P(tht), P(th), P(evisc), P(f.sp["th"].flux_bot), P(f.sp["th"].flux_top), P(dzi), P(dzhi),
                                                      P(one), P(one), C.c_double(3.), C.c_double(1e-5), stream))
ref["pres_in_g"] = timeit(lambda: R.refcu_pres2_in(C.byref(rg), P(p), P(u), P(v), P(w), P(ut), P(vt), P(wt), P(dzi), P(one), P(one), C.c_double(1.), stream))
ref["solve_in_g"] = timeit(lambda: R.refcu_pres2_solve_in(C.byref(rg), P(p), P(tmp1), P(tmp2), P(bm), P(bm), P(dz), P(one), P(bm), P(bm), stream))
tmp2.fill_(4.)     # a diagonally dominant system so that the sweep stays finite
ref["tdma_g"] = timeit(lambda: R.refcu_pres2_tdma(C.byref(rg), P(bm), P(tmp2), P(bm), P(p), P(tmp1), stream))
ref["memcpy p -> tmp (cudaMemcpy, src/pres_2.cu:312)"] = timeit(lambda: tmp1.copy_(p))
ref["solve_out_g"] = timeit(lambda: R.refcu_pres2_solve_out(C.byref(rg), P(p), P(tmp1), stream))
ref["pres_out_g"] = timeit(lambda: R.refcu_pres2_out(C.byref(rg), P(ut), P(vt), P(wt), P(p), P(dzhi), stream))
ref["rk3_g (one field)"] = timeit(lambda: R.refcu_rk3(C.byref(rg), 1, P(tht), P(th), C.c_double(1e-6), stream))
assert torch.cuda.current_stream().cuda_stream == 0
pairs = [("strain2 + evisc (reference also runs a separate N2 kernel)", ["calc_strain2_g", "evisc_g"], ["evisc_kernel"]),
         ("momentum: advection + diffusion (this repo's pass also holds buoyancy, sponge, Coriolis)", ["advec_uvw_g", "diff_uvw_g"], ["advec2_smag2_uvw_kernel"]),
         ("scalar: advection + diffusion + RK update (this repo's pass also holds the sponge term)", ["advec_s_g", "diff_c_g", "rk3_g (one field)"], ["advec2_smag2_c_kernel"]),
         ("Pres_2::input", ["pres_in_g"], ["pres2_in_kernel"]),
         ("tridiagonal solve", ["solve_in_g", "tdma_g"], ["pres2_tdma_kernel"]),
         ("solve tail + output + RK of u,v,w", ["memcpy p -> tmp (cudaMemcpy, src/pres_2.cu:312)", "solve_out_g", "pres_out_g", "rk3_g (one field)", "rk3_g (one field)", "rk3_g (one field)"],
          ["pres2_out_rk_kernel", "pres2_bottom_ghost_kernel"])]
print(f"# reference .cu kernels vs this repository, {n}^3 FP64, one B200, ms per launch (CUDA events)")
for k, vv in ref.items():
    print(f"  reference {k:52s} {vv:8.3f} ms")
print("  this repo " + ", ".join(f"{k} {vv:.3f}" for k, vv in sorted(mine.items(), key=lambda kv: -kv[1])[:14]))
rows = []
for what, rk, mk in pairs:
    r, mm = sum(ref[k] for k in rk), sum(mine.get(k, 0.) for k in mk)
    rows.append(dict(what=what, reference_ms=round(r, 3), this_repo_ms=round(mm, 3), ratio=round(r / mm, 2) if mm else None))
    print(f"{what:95s} reference {r:7.3f} ms | this repo {mm:7.3f} ms | x{r / mm if mm else float('nan'):.2f}")
json.dump(dict(n=n, reference=ref, this_repo=mine, rows=rows), open("gpurun_out/refcu_%d.json" % n, "w"), indent=1)
