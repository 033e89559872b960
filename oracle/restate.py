"""TEST INFRASTRUCTURE ONLY — numpy restatements of the few hot-path diagnostics whose reference translation unit
cannot be compiled here (src/fields.cxx needs Boost and NetCDF).  Each function cites the reference lines it follows;
sums are pairwise (numpy) instead of sequential, so comparisons are tolerance-level (1e-12), never bit-level.
Pinned by closed-form cases in tests/test_oracle_cpu.py (uniform fields)."""
from __future__ import annotations

import numpy as np


def _interior(W):
    return (slice(W.kstart, W.kend), slice(W.jstart, W.jend), slice(W.istart, W.iend))


def _shift(W, di=0, dj=0, dk=0):
    return (slice(W.kstart + dk, W.kend + dk), slice(W.jstart + dj, W.jend + dj), slice(W.istart + di, W.iend + di))


def check_momentum(W, u, v, w):
    """calc_momentum_2nd, src/fields.cxx:144-172: sum((u_c + v_c + w_c) * dz) / (itot*jtot*zsize)."""
    c = _interior(W)
    dz = W.dz[W.kstart:W.kend, None, None]
    m = (0.5 * (u[c] + u[_shift(W, di=1)]) + 0.5 * (v[c] + v[_shift(W, dj=1)]) + 0.5 * (w[c] + w[_shift(W, dk=1)])) * dz
    return float(m.sum() / (W.itot * W.jtot * W.zsize))


def check_tke(W, u, v, w):
    """calc_tke_2nd, src/fields.cxx:175-206."""
    c = _interior(W)
    dz = W.dz[W.kstart:W.kend, None, None]
    e = (0.5 * (u[c] ** 2 + u[_shift(W, di=1)] ** 2) + 0.5 * (v[c] ** 2 + v[_shift(W, dj=1)] ** 2)
         + 0.5 * (w[c] ** 2 + w[_shift(W, dk=1)] ** 2)) * dz
    return float(0.5 * e.sum() / (W.itot * W.jtot * W.zsize))


def check_mass(W, s):
    """calc_mass, src/fields.cxx:209-234."""
    dz = W.dz[W.kstart:W.kend, None, None]
    return float((s[_interior(W)] * dz).sum() / (W.itot * W.jtot * W.zsize))


def calc_max(W, a):
    """Field3d_operators::calc_max_g (GPU-only in the reference, src/field3d_operators.cu:189-211): max over the interior."""
    return float(a[_interior(W)].max())
