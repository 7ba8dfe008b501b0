"""dplasma_b200 -- host-side Python mirror of the C ABI in include/dplasma_b200.h.

The product is the shared library dplasma_b200/lib/libdplasma_b200.so (hand-written sm_100a
CUDA behind the reference's own dplasma_dpotrf / dposv / pdpotrf_ entry points).  This
package only loads it through ctypes and mirrors its structs, for the tests and bench.py.
There is no CPU fallback: loading fails loudly if the library is missing, and every
operation fails if no CUDA device is present.
"""
from .capi import (  # noqa: F401
    Lib, load, build, Context, BlockCyclic, LIB_PATH,
    NoTrans, Trans, ConjTrans, Upper, Lower, UpperLower, NonUnit, Unit, Left, Right, InfNorm,
    MATRIX_DOUBLE, MATRIX_TILE, MATRIX_LAPACK,
)
