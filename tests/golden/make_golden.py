"""Generates tests/golden/golden_r01.npz from the REFERENCE's own compiled code.

Run in the build container (needs /root/reference and oracle/_ref/libdplasma_ref.so, see
oracle/build_ref.sh):   python tests/golden/make_golden.py
The reference holds no golden vectors of its own (all its tests are residual thresholds,
src/dplasma_zcheck.c:119,215); these fixtures freeze the outputs of its generators
(CORE_dplgsy / CORE_zplghe / CORE_dplrnt / CORE_zplrnt, bit-exact) and of its CPU tile
kernels (CORE_[dz]potrf/trsm/syrk|herk/gemm over OpenBLAS 0.3.15) run through the DAG
order of src/zpotrf_L.jdf / zpotrf_U.jdf and src/ztrsm_LLN.jdf / ztrsm_LLT.jdf, so that the
oracle and the CUDA path can be checked against them on machines without /root/reference.
Sizes follow the reference's own ctest cases (tests/Testings.cmake:89-90: -N 378 -t 93 -K 367).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.oracle import Lower, Upper, UpperLower, Reference  # noqa: E402

R = Reference()
out = {}
SEED_A, SEED_B, SEED_B_POSV = 3872, 3873, 2354  # tests/common.c:274, testing_zpotrf.c:104, testing_zposv.c:82

# Appendix-C case: N=8, nb=4, bump=8
A = R.plgsy(8.0, UpperLower, 8, 4, SEED_A)
out["d8_A"] = A
for name, uplo in (("L", Lower), ("U", Upper)):
    F = R.plgsy(8.0, uplo, 8, 4, SEED_A)
    info = R.potrf(uplo, F, 4)
    assert info == 0
    out[f"d8_{name}"] = F
out["d8_B"] = R.plrnt(8, 1, 4, 1, SEED_B)

# reference ctest case: N=378 nb=93 (ragged last tile 6), NRHS=367 nbrhs=93
N, nb, K = 378, 93, 367
for p in "dz":
    for name, uplo in (("L", Lower), ("U", Upper)):
        A0 = R.plgsy(float(N), uplo, N, nb, SEED_A, p)
        F = A0.copy(order="F")
        assert R.potrf(uplo, F, nb, p) == 0
        out[f"{p}378_A0_{name}_s31"] = A0.ravel(order="F")[::31].copy()
        out[f"{p}378_A0_{name}_sum"] = np.array([A0.sum(), np.abs(A0).sum()])
        out[f"{p}378_F_{name}_s31"] = F.ravel(order="F")[::31].copy()
        out[f"{p}378_F_{name}_diag"] = np.diag(F).copy()
        if p == "d":
            B = R.plrnt(N, K, nb, nb, SEED_B_POSV)
            out["d378_B_s41"] = B.ravel(order="F")[::41].copy()
            X = B.copy(order="F")
            R.potrs(uplo, F, X, nb, nb)
            out[f"d378_X_{name}_s41"] = X.ravel(order="F")[::41].copy()
    out[f"{p}378_B3873_s41"] = R.plrnt(N, 5, nb, 5, SEED_B, p).ravel(order="F")[::41].copy()

# tile-size independence / block-cyclic independence of the generator: N=200 generated
# with nb=64 must equal nb=25
out["d200_A_full"] = R.plgsy(200.0, UpperLower, 200, 64, SEED_A)[::3, ::5].copy()

# not positive definite: bump=0 -> first failing leading minor (global, 1-based)
A = R.plgsy(0.0, Lower, 200, 64, SEED_A)
out["d200_nonspd_info_L"] = np.array([R.potrf(Lower, A, 64)])
A = R.plgsy(0.0, Upper, 200, 64, SEED_A)
out["d200_nonspd_info_U"] = np.array([R.potrf(Upper, A, 64)])
# SPD except one late pivot: make entry (150,150) hugely negative
A = R.plgsy(200.0, Lower, 200, 64, SEED_A)
A[150, 150] = -1.0
out["d200_pivot150_info_L"] = np.array([R.potrf(Lower, A, 64)])

path = os.path.join(os.path.dirname(__file__), "golden_r01.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})
