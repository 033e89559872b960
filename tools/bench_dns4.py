"""Per-kernel timing of the 4th-order DNS path (advec_4, diff_4, pres_4, RK3; moser180-like channel) and of the
advec_2i5 LES variant on one GPU:  python tools/bench_dns4.py [itot jtot ktot]"""
import sys, json
import numpy as np, torch
sys.path.insert(0, ".")
from microhh_b200.model import Model

def stretched(k, size, alpha=2.0):
    zh = 0.5 * size * (1. + np.tanh(alpha * (2. * np.arange(k + 1) / k - 1.)) / np.tanh(alpha))
    return 0.5 * (zh[1:] + zh[:-1])

itot, jtot, ktot = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (512, 512, 256)
for tag, ini in (
    ("dns4", {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="4"),
              "advec": dict(swadvec="4", cflmax=1.0), "diff": dict(swdiff="4", dnmax=0.4), "pres": dict(swpres="4"),
              "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-4),
              "time": dict(rkorder=3, dt=1e-4, dtmax=1e-3, adaptivestep=True)}),
    ("les2i5", {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="2", gc=(3, 3, 2)),
                "advec": dict(swadvec="2i5", cflmax=1.0), "diff": dict(swdiff="2", dnmax=0.4), "pres": dict(swpres="2"),
                "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-4),
                "time": dict(rkorder=3, dt=1e-4, dtmax=1e-3, adaptivestep=True)})):
    z = stretched(ktot, 2.)
    try:
        m = Model(ini, z, device=torch.device("cuda:0"))
    except Exception as e:
        print(tag, "skipped:", e); continue
    g, f = m.grid, m.fields
    rng = np.random.default_rng(1)
    zf = np.zeros(g.kcells); zf[g.kstart:g.kend] = z
    for n, amp, mean in (("u", 0.01, 0.11), ("v", 0.01, 0.), ("w", 0.01, 0.)):
        a = (zf * (2. - zf))[:, None, None] * mean + amp * rng.uniform(-1, 1, g.shape)
        if n == "w": a[:g.kstart + 1] = 0.; a[g.kend:] = 0.
        f.mp[n].fld.copy_(torch.from_numpy(a).cuda())
    for _ in range(2): m.step()
    torch.cuda.synchronize()
    g.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): m.step()
    e1.record(); torch.cuda.synchronize()
    prof = g.profile_report()
    pts = itot * jtot * ktot
    print(f"== {tag} {itot}x{jtot}x{ktot}: {e0.elapsed_time(e1) / 3:.2f} ms/step, {pts * 3 / (e0.elapsed_time(e1) * 1e-3):.3e} grid-point-steps/s, status {m.print_status()}")
    for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:12]:
        print(f"   {k:28s} {v[0] / v[1]:8.3f} ms x {v[1]:3d}   {pts * 8 / (v[0] / v[1] * 1e-3) / 1e9:8.1f} GB/s per 8 B/pt")
    g.close()
