"""Times the chunk-layout FFT stages of one rank of an npy-way y-slab decomposition on ONE GPU (no collective):
python tools/bench_fft_chunks.py [n]   -> ms per pass for npy = 1, 2, 4, 8 (512^3 points per rank)."""
import sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, ".")
from microhh_b200 import host as H, lib as L

SHAPES = [(512, 512, 512, 1), (512, 1024, 512, 2), (1024, 1024, 512, 4), (1024, 2048, 512, 8), (1024, 1024, 1024, 8)]
for itot, jtot, ktot, npy in SHAPES:
    n = 512
    z = (np.arange(ktot) + 0.5) / ktot
    g = H.Grid(itot, jtot, ktot, 1., 1., 1., z, npy=npy, rank_y=0, device=torch.device("cuda:0"))
    if not g.lib.mhh_fft_chunks_supported(g.ctx):
        print(itot, jtot, ktot, npy, "chunk FFT stages unsupported"); g.close(); continue
    L.check(g.ctx, g.lib.mhh_pres_set_values(g.ctx), "set_values")
    A = torch.rand(n * n * n, dtype=torch.float64, device="cuda")
    B = torch.empty_like(A)
    pa, pb = L.ptr(A), L.ptr(B)
    def timeit(fn, reps=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    res = {}
    for nk, tag in ((0, "full"), (ktot // 4, "k/4")):
        rng = [(0, 0)] if nk == 0 else [(c * nk, nk) for c in range(4)]
        res[tag] = (
            timeit(lambda: [g.lib.mhh_fft_x_chunks(g.ctx, -1, pa, pb, None, k0, k) for k0, k in rng]),
            timeit(lambda: [g.lib.mhh_fft_y_chunks(g.ctx, -1, pa, pb, k0, k) for k0, k in rng]),
            timeit(lambda: [g.lib.mhh_fft_y_chunks(g.ctx, +1, pa, pb, k0, k) for k0, k in rng]),
            timeit(lambda: [g.lib.mhh_fft_x_chunks(g.ctx, +1, pa, pb, None, k0, k) for k0, k in rng]))
    print(f"{itot}x{jtot}x{ktot} npy={npy} iblock={itot // npy} jmax={jtot // npy}", {k: [round(x, 3) for x in v] for k, v in res.items()}, "(fwd_x, fwd_y, bwd_y, bwd_x)")
    g.close()
