"""Generates tests/golden/golden_r01_blas3.npz from the REFERENCE's own compiled code
(oracle/_ref/libdplasma_ref.so, built by oracle/build_ref.sh from /root/reference).

Run in the build container:   python tests/golden/make_golden_blas3.py
Freezes outputs of the reference's CPU tile kernels CORE_dgemm / CORE_dsyrk (over OpenBLAS 0.3.15) run in the
per-tile k order of src/zgemm_NN.jdf:197-211 (and NT/TN/TT) and src/zsyrk_{L,U}{N,T}.jdf, on inputs from the
reference's generator CORE_dplrnt, at a ragged size with three different tile edges.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.oracle import Lower, Upper, NoTrans, Trans, Reference  # noqa: E402

R = Reference()
out = {}
M, N, K, mb, nb, kb = 378, 300, 367, 93, 64, 80
for ta, na in ((NoTrans, "N"), (Trans, "T")):
    for tb, nbn in ((NoTrans, "N"), (Trans, "T")):
        A = R.plrnt(*((M, K, mb, kb) if ta == NoTrans else (K, M, kb, mb)), 11)
        B = R.plrnt(*((K, N, kb, nb) if tb == NoTrans else (N, K, nb, kb)), 12)
        C = R.plrnt(M, N, mb, nb, 13)
        R.gemm(ta, tb, 0.75, A, B, -0.5, C, mb, nb, kb)
        out[f"dgemm_{na}{nbn}_s17"] = C.ravel(order="F")[::17].copy()
for uplo, nu in ((Lower, "L"), (Upper, "U")):
    for tr, nt_ in ((NoTrans, "N"), (Trans, "T")):
        A = R.plrnt(*((N, K, nb, kb) if tr == NoTrans else (K, N, kb, nb)), 21)
        C = R.plrnt(N, N, nb, nb, 22)
        R.syrk(uplo, tr, 0.5, A, -0.25, C, nb, kb)
        out[f"dsyrk_{nu}{nt_}_s17"] = C.ravel(order="F")[::17].copy()
path = os.path.join(os.path.dirname(__file__), "golden_r01_blas3.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})
