"""rampack_b200 — B200-native Monte Carlo sweep behind RAMPACK's Packing / Interaction contracts.

The product is the CUDA library `librampack_b200.so` (rampack_b200/csrc, C ABI in include/rampack_b200.h);
this package holds its build recipe and the ctypes plumbing tests and bench.py use.
"""
from . import build  # noqa: F401
from .capi import (Device, RpError, PreconditionError, make_desc, load, PAIR_DTYPE, TRACE_DTYPE,  # noqa: F401
                   RP_HARD_NONE, RP_HARD_SPHERE, RP_HARD_SPHEROCYLINDER, RP_HARD_POLYSPHERE,
                   RP_HARD_POLYSPHEROCYLINDER, RP_HARD_XENOCOLLIDE, XcProgram, RP_SOFT_NONE, RP_SOFT_LJ, RP_SOFT_WCA, RP_SOFT_SIC,
                   RP_MOVE_TRANSLATION, RP_MOVE_ROTATION, RP_MOVE_ROTOTRANSLATION, RP_MOVE_FLIP, NativeSimulation)
