"""Times the vertical (Thomas) solve of one rank of an npy-way y-slab decomposition on ONE GPU: the whole block in one launch
(mhh_pres_solve_spectral_ws), and cut into kx sub-blocks (mhh_slab_solve_sub) issued one after the other and on one stream
per sub-block, plus the sub-block y transforms.  MHH_TDMA_PIPE=0/1 selects the sub-block kernel (read once per process).
python tools/bench_tdma_sub.py"""
import os, sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, ".")
from microhh_b200 import host as H, lib as L

SHAPES = [(512, 512, 512, 1), (512, 1024, 512, 2), (1024, 1024, 512, 4), (1024, 1024, 1024, 8)]
print("MHH_TDMA_PIPE =", os.environ.get("MHH_TDMA_PIPE", "1 (default)"))
for itot, jtot, ktot, npy in SHAPES:
    z = (np.arange(ktot) + 0.5) / ktot
    g = H.Grid(itot, jtot, ktot, 1., 1., 1., z, npy=npy, rank_y=0, device=torch.device("cuda:0"))
    L.check(g.ctx, g.lib.mhh_pres_set_values(g.ctx), "set_values")
    n = itot * (jtot // npy) * ktot
    Y = torch.rand(n, dtype=torch.float64, device="cuda")
    R = torch.rand(n, dtype=torch.float64, device="cuda")
    W = [torch.empty(n // 8, dtype=torch.float64, device="cuda") for _ in range(8)]
    main = torch.cuda.current_stream()
    streams = [torch.cuda.Stream(priority=-1) for _ in range(8)]
    def timeit(fn, reps=4):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        return round(e0.elapsed_time(e1) / reps, 3)
    def on(st): g.lib.mhh_set_stream(g.ctx, C.c_void_p(st.cuda_stream))
    def seq(B):
        for b in range(B):
            L.check(g.ctx, g.lib.mhh_slab_solve_sub(g.ctx, L.ptr(Y), L.ptr(W[0]), b, B), "solve_sub")
    def par(B):
        ev = torch.cuda.Event(); ev.record(main)
        for b in range(B):
            streams[b].wait_event(ev); on(streams[b])
            L.check(g.ctx, g.lib.mhh_slab_solve_sub(g.ctx, L.ptr(Y), L.ptr(W[b]), b, B), "solve_sub")
            e = torch.cuda.Event(); e.record(streams[b]); main.wait_event(e)
        on(main)
    def yfft(B, d):
        for b in range(B):
            if d < 0: L.check(g.ctx, g.lib.mhh_slab_fft_y_fwd_sub(g.ctx, L.ptr(R), L.ptr(Y), b, B), "y_fwd_sub")
            else:
                tab = (C.c_void_p * npy)(*[R.data_ptr() + (r * B + b) * (n // (npy * B)) * 8 for r in range(npy)])
                L.check(g.ctx, g.lib.mhh_slab_fft_y_bwd_sub(g.ctx, L.ptr(Y), tab, b, B), "y_bwd_sub")
    res = {"full": timeit(lambda: L.check(g.ctx, g.lib.mhh_pres_solve_spectral_ws(g.ctx, L.ptr(Y), L.ptr(W[0])), "solve"))}
    for B in (1, 2, 4):
        if (itot // npy) % B or npy * B > 32: continue
        res[f"sub{B}_seq"] = timeit(lambda: seq(B))
        if B > 1: res[f"sub{B}_par"] = timeit(lambda: par(B))
        res[f"yfwd{B}"] = timeit(lambda: yfft(B, -1)); res[f"ybwd{B}"] = timeit(lambda: yfft(B, +1))
    print(f"{itot}x{jtot}x{ktot} npy={npy} iblock={itot // npy} jtot={jtot} kmax={ktot}:", res)
    g.close(); del Y, R, W
    torch.cuda.empty_cache()
