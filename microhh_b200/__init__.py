"""microhh_b200 — B200-native (sm_100a) dynamical core behind MicroHH's Advec/Diff/Pres/Boundary/Timeloop
interfaces.  The compute path is ``libmhh_b200.so`` (hand-written CUDA, C ABI in ``include/mhh_b200.h``);
this package is the host-side mirror of the reference's operator interface."""
from . import lib  # noqa: F401

__all__ = ["lib"]
