"""GPU parity of the hand-written Stockham FFTs (fft.cu generic mixed radix, fft_pow2.cu register-resident)
against (a) numpy.fft restated in FFTW's half-complex order and (b) the reference's FFT::exec_forward/backward
compiled in place over the oracle's DFT shim (oracle/_ref).  Tolerance 1e-13 * max|spectrum| (FP64)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def to_halfcomplex(x, axis):
    """FFTW R2HC along `axis`: r0..r(n/2), i((n+1)/2-1)..i1."""
    n = x.shape[axis]
    f = np.fft.rfft(x, axis=axis)
    out = np.empty_like(x)
    sl = [slice(None)] * x.ndim
    for k in range(n // 2 + 1):
        sl[axis] = k
        fk = np.take(f, k, axis=axis)
        out[tuple(sl)] = fk.real
        if 0 < k < (n + 1) // 2:
            sl[axis] = n - k
            out[tuple(sl)] = fk.imag
    return out


SIZES = [(8, 8, 3), (16, 32, 5), (64, 16, 4), (512, 8, 3), (32, 256, 2), (1024, 16, 2), (48, 24, 3), (20, 12, 5),
         (96, 10, 2), (16, 1, 4), (36, 14, 3), (512, 512, 10), (1024, 64, 40), (8, 16, 2),
         # y passes on the lanes-along-the-lines kernel: 4 / 16 pairs per CTA; (24, 512): lines not in groups of 16 -> staged tile
         (16, 1024, 2), (64, 256, 3), (24, 512, 2)]


@pytest.mark.parametrize("itot,jtot,ktot", SIZES)
@pytest.mark.parametrize("generic", [0, 1])
def test_fft_forward_backward(cuda_lib, itot, jtot, ktot, generic):
    import torch
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    z = (np.arange(ktot) + 0.5) / ktot
    g = H.Grid(itot, jtot, ktot, 1., 1., 1., z)
    L.check(g.ctx, g.lib.mhh_pres_set_values(g.ctx), "set_values")
    g.lib.mhh_fft_force_generic(g.ctx, generic)
    rng = np.random.default_rng(itot * 1000 + jtot)
    a = rng.uniform(-1, 1, (ktot, jtot, itot))
    want = to_halfcomplex(to_halfcomplex(a, 2), 1)
    d = torch.from_numpy(a.copy()).cuda()
    L.check(g.ctx, g.lib.mhh_fft_forward(g.ctx, L.ptr(d)), "fwd")
    got = d.cpu().numpy()
    scale = np.abs(want).max()
    assert np.abs(got - want).max() <= 1e-13 * scale
    L.check(g.ctx, g.lib.mhh_fft_backward(g.ctx, L.ptr(d)), "bwd")
    back = d.cpu().numpy()
    assert np.abs(back - a).max() <= 1e-13 * max(1., np.abs(a).max())
    g.close()


@pytest.mark.parametrize("itot,jtot,ktot", [(32, 16, 4), (24, 40, 3)])
def test_fft_matches_reference_fft_class(cuda_lib, itot, jtot, ktot):
    import torch
    from microhh_b200 import host as H
    from microhh_b200 import lib as L
    from oracle.ref import RefWorld
    z = (np.arange(ktot) + 0.5) / ktot
    W = RefWorld(itot, jtot, ktot, 1., 1., 1., z)
    g = H.Grid(itot, jtot, ktot, 1., 1., 1., z)
    L.check(g.ctx, g.lib.mhh_pres_set_values(g.ctx), "set_values")
    a = np.random.default_rng(5).uniform(-1, 1, (ktot, jtot, itot))
    ref = a.copy()
    W.call("ref_fft_forward", ref)
    d = torch.from_numpy(a.copy()).cuda()
    L.check(g.ctx, g.lib.mhh_fft_forward(g.ctx, L.ptr(d)), "fwd")
    assert np.abs(d.cpu().numpy() - ref).max() <= 1e-13 * np.abs(ref).max()
    W.close(); g.close()
