"""Generates tests/golden/golden_r01_inverse.npz from the REFERENCE's own compiled code
(oracle/_ref/libdplasma_ref.so, built by oracle/build_ref.sh from /root/reference).

Run in the build container:   python tests/golden/make_golden_inverse.py
Freezes the outputs of the reference's CPU tile kernels CORE_dtrtri / CORE_dlauum / CORE_dtrmm /
CORE_dtrsm / CORE_dgemm / CORE_dsyrk (over OpenBLAS 0.3.15) run in the step order of
src/ztrtri_{L,U}.jdf and src/zlauum_{L,U}.jdf, on the reference's own test case
(tests/Testings.cmake:146-149: testing_dposv -N 378 -t 93, whose fourth check is
ZPOTRF + ZPOTRI, tests/testing_zposv.c:217-254).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.oracle import Lower, Upper, NonUnit, Unit, Reference  # noqa: E402

R = Reference()
out = {}
SEED_A = 3872  # tests/common.c:274
N, nb = 378, 93
for name, uplo in (("L", Lower), ("U", Upper)):
    A0 = R.plgsy(float(N), uplo, N, nb, SEED_A)
    F = A0.copy(order="F")
    assert R.potrf(uplo, F, nb) == 0
    T = F.copy(order="F")
    assert R.trtri(uplo, NonUnit, T, nb) == 0
    out[f"d378_trtri_{name}_s31"] = T.ravel(order="F")[::31].copy()
    Tu = F.copy(order="F")
    assert R.trtri(uplo, Unit, Tu, nb) == 0
    out[f"d378_trtri_unit_{name}_s31"] = Tu.ravel(order="F")[::31].copy()
    U = F.copy(order="F")
    R.lauum(uplo, U, nb)
    out[f"d378_lauum_{name}_s31"] = U.ravel(order="F")[::31].copy()
    P = F.copy(order="F")
    assert R.potri(uplo, P, nb) == 0
    out[f"d378_potri_{name}_s31"] = P.ravel(order="F")[::31].copy()
    out[f"d378_potri_{name}_diag"] = np.diag(P).copy()
    # an exactly singular triangle: INFO = k*mb + info (src/ztrtri_L.jdf:180-186)
    S = F.copy(order="F")
    S[200, 200] = 0.0
    out[f"d378_trtri_singular_info_{name}"] = np.array([R.trtri(uplo, NonUnit, S, nb)])

path = os.path.join(os.path.dirname(__file__), "golden_r01_inverse.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})
