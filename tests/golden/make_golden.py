#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the reference's own code (oracle/_ref, built by `make -C oracle` where
/root/reference exists).  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from tests import golden_lib as G  # noqa: E402

for tag in G.KERNEL_CASES:
    np.savez(os.path.join(G.GOLDEN_DIR, f"kernels_{tag}.npz"), **G.oracle_kernels(tag))
_, les = G.oracle_les_substep()
np.savez(os.path.join(G.GOLDEN_DIR, "les_substep.npz"), **les)
np.savez(os.path.join(G.GOLDEN_DIR, "dns4_steps.npz"), **G.oracle_dns4_steps())
np.savez(os.path.join(G.GOLDEN_DIR, "kernels_r02.npz"), **G.oracle_r02())
for f in sorted(os.listdir(G.GOLDEN_DIR)):
    if f.endswith(".npz"):
        print(f, os.path.getsize(os.path.join(G.GOLDEN_DIR, f)))
