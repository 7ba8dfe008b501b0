"""TEST INFRASTRUCTURE ONLY -- ctypes front-end for the CPU oracle (oracle/liboracle.so)
and, when it was built, for the reference's own tile kernels (oracle/_ref/libdplasma_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (dplasma_b200) never does.

All matrices are numpy float64 / complex128 arrays in Fortran (column-major) order, the
storage the reference's tile kernels work on (src/cores/core_zgemm.c:90-103 takes lda).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# src/include/dplasma/constants.h:78-96
NoTrans, Trans, ConjTrans = 111, 112, 113
Upper, Lower, UpperLower = 121, 122, 123
NonUnit, Unit = 131, 132
Left, Right = 141, 142

_c_int, _c_dbl, _c_u64 = ctypes.c_int, ctypes.c_double, ctypes.c_uint64
_vp = ctypes.c_void_p


def build(force: bool = False) -> None:
    """Compile liboracle.so (always) and oracle/_ref (only where /root/reference exists)."""
    so = os.path.join(HERE, "liboracle.so")
    srcs = [os.path.join(HERE, f) for f in ("oracle.c", "oracle_impl.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(
            ["gcc", "-O2", "-fno-fast-math", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
             "-o", so, os.path.join(HERE, "oracle.c"), "-lm"])
    ref = os.path.join(HERE, "_ref", "libdplasma_ref.so")
    if os.path.isdir("/root/reference/src/cores") and (
            force or not os.path.exists(ref)
            or os.path.getmtime(os.path.join(HERE, "ref_driver.c")) > os.path.getmtime(ref)):
        subprocess.check_call(["bash", os.path.join(HERE, "build_ref.sh")])


def _ptr(a: np.ndarray):
    assert a.flags.f_contiguous or a.ndim == 1, "oracle arrays must be column-major"
    return a.ctypes.data_as(_vp)


class _Cplx(ctypes.Structure):
    _fields_ = [("re", _c_dbl), ("im", _c_dbl)]


class Oracle:
    """Plain-C restatement (oracle.c).  `p` selects precision: 'd' or 'z'."""

    def __init__(self):
        build()
        self.lib = ctypes.CDLL(os.path.join(HERE, "liboracle.so"))
        L = self.lib
        L.orc_rnd64_jump.restype = _c_u64
        L.orc_rnd64_jump.argtypes = [_c_u64, _c_u64]
        for f in ("orc_flops_dpotrf", "orc_flops_zpotrf"):
            getattr(L, f).restype = _c_dbl
            getattr(L, f).argtypes = [_c_dbl]
        L.orc_flops_dpotrs.restype = _c_dbl
        L.orc_flops_dpotrs.argtypes = [_c_dbl, _c_dbl]
        for p in "dz":
            sc = _c_dbl if p == "d" else _Cplx
            g = lambda n: getattr(L, f"orc_{p}{n}")
            g("plgsy").argtypes = [_c_dbl, _c_int, _c_int, _c_int, _c_int, _c_int, _vp, _c_int, _c_u64]
            g("plgsy").restype = None
            g("plrnt").argtypes = [_c_int, _c_int, _c_int, _c_int, _vp, _c_int, _c_u64]
            g("plrnt").restype = None
            g("plgsy_tile").argtypes = [_c_dbl, _c_int, _c_int, _vp, _c_int, _c_int, _c_int, _c_int, _c_u64]
            g("plgsy_tile").restype = None
            g("plrnt_tile").argtypes = [_c_int, _c_int, _vp, _c_int, _c_int, _c_int, _c_int, _c_u64]
            g("plrnt_tile").restype = None
            g("potrf_tile").argtypes = [_c_int, _c_int, _vp, _c_int, ctypes.POINTER(_c_int)]
            g("potrf_tile").restype = None
            g("gemm_tile").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, sc, _vp, _c_int, _vp, _c_int, sc, _vp, _c_int]
            g("gemm_tile").restype = None
            g("herk_tile").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_dbl, _vp, _c_int, _c_dbl, _vp, _c_int]
            g("herk_tile").restype = None
            g("trsm_tile").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, _c_int, sc, _vp, _c_int, _vp, _c_int]
            g("trsm_tile").restype = None
            g("potrf").argtypes = [_c_int, _c_int, _c_int, _vp, _c_int]
            g("potrf").restype = _c_int
            g("trsm").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int, sc, _vp, _c_int, _vp, _c_int]
            g("trsm").restype = _c_int
            g("potrs").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, _vp, _c_int, _vp, _c_int]
            g("potrs").restype = _c_int
            g("posv").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, _vp, _c_int, _vp, _c_int]
            g("posv").restype = _c_int
            g("lanhe_inf").argtypes = [_c_int, _c_int, _vp, _c_int]
            g("lanhe_inf").restype = _c_dbl
            g("lange_inf").argtypes = [_c_int, _c_int, _vp, _c_int]
            g("lange_inf").restype = _c_dbl
            g("check_potrf").argtypes = [_c_int, _c_int, _vp, _c_int, _vp, _c_int] + [ctypes.POINTER(_c_dbl)] * 3
            g("check_potrf").restype = _c_int
            g("check_axmb").argtypes = [_c_int, _c_int, _c_int, _vp, _c_int, _vp, _c_int, _vp, _c_int, ctypes.POINTER(_c_dbl)]
            g("check_axmb").restype = _c_int
            g("trtri_tile").argtypes = [_c_int, _c_int, _c_int, _vp, _c_int, ctypes.POINTER(_c_int)]
            g("trtri_tile").restype = None
            g("lauum_tile").argtypes = [_c_int, _c_int, _vp, _c_int]
            g("lauum_tile").restype = None
            g("trmm_tile").argtypes = [_c_int, _c_int, _c_int, _c_int, _c_int, _c_int, sc, _vp, _c_int, _vp, _c_int]
            g("trmm_tile").restype = None
            g("trtri").argtypes = [_c_int, _c_int, _c_int, _c_int, _vp, _c_int]
            g("trtri").restype = _c_int
            for f in ("lauum", "potri", "poinv"):
                g(f).argtypes = [_c_int, _c_int, _c_int, _vp, _c_int]
                g(f).restype = _c_int
            g("check_poinv").argtypes = [_c_int, _c_int, _vp, _c_int, _vp, _c_int, ctypes.POINTER(_c_dbl)]
            g("check_poinv").restype = _c_int
            g("gemm").argtypes = [_c_int] * 8 + [sc, _vp, _c_int, _vp, _c_int, sc, _vp, _c_int]
            g("gemm").restype = _c_int
            g("herk").argtypes = [_c_int] * 6 + [_c_dbl, _vp, _c_int, _c_dbl, _vp, _c_int]
            g("herk").restype = _c_int

    @staticmethod
    def _dt(p):
        return np.float64 if p == "d" else np.complex128

    def _sc(self, p, v):
        return float(v) if p == "d" else _Cplx(complex(v).real, complex(v).imag)

    def rnd64_jump(self, n, seed):
        return int(self.lib.orc_rnd64_jump(n, seed))

    def plgsy(self, bump, uplo, N, nb, seed, p="d", M=None, mb=None):
        """d: dplgsy ; z: zplghe.  Untouched triangle stays 0."""
        M = N if M is None else M
        mb = nb if mb is None else mb
        A = np.zeros((M, N), dtype=self._dt(p), order="F")
        getattr(self.lib, f"orc_{p}plgsy")(float(bump), uplo, M, N, mb, nb, _ptr(A), M, seed)
        return A

    def plrnt(self, M, N, mb, nb, seed, p="d"):
        A = np.zeros((M, N), dtype=self._dt(p), order="F")
        getattr(self.lib, f"orc_{p}plrnt")(M, N, mb, nb, _ptr(A), M, seed)
        return A

    def potrf_tile(self, uplo, A, p="d"):
        info = _c_int(0)
        getattr(self.lib, f"orc_{p}potrf_tile")(uplo, A.shape[0], _ptr(A), A.shape[0], ctypes.byref(info))
        return info.value

    def gemm_tile(self, ta, tb, alpha, A, B, beta, C, p="d"):
        m, n = C.shape
        k = A.shape[1] if ta == NoTrans else A.shape[0]
        getattr(self.lib, f"orc_{p}gemm_tile")(ta, tb, m, n, k, self._sc(p, alpha), _ptr(A), A.shape[0],
                                               _ptr(B), B.shape[0], self._sc(p, beta), _ptr(C), C.shape[0])

    def herk_tile(self, uplo, trans, alpha, A, beta, C, p="d"):
        n = C.shape[0]
        k = A.shape[1] if trans == NoTrans else A.shape[0]
        getattr(self.lib, f"orc_{p}herk_tile")(uplo, trans, n, k, float(alpha), _ptr(A), A.shape[0],
                                               float(beta), _ptr(C), C.shape[0])

    def trsm_tile(self, side, uplo, trans, diag, alpha, A, B, p="d"):
        m, n = B.shape
        getattr(self.lib, f"orc_{p}trsm_tile")(side, uplo, trans, diag, m, n, self._sc(p, alpha),
                                               _ptr(A), A.shape[0], _ptr(B), B.shape[0])

    def potrf(self, uplo, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"orc_{p}potrf")(uplo, N, nb, _ptr(A), N)

    def trsm(self, side, uplo, trans, diag, alpha, A, B, nb, nbr, p="d"):
        M, R = B.shape
        return getattr(self.lib, f"orc_{p}trsm")(side, uplo, trans, diag, M, R, nb, nbr,
                                                 self._sc(p, alpha), _ptr(A), A.shape[0], _ptr(B), M)

    def potrs(self, uplo, A, B, nb, nbr, p="d"):
        N, R = B.shape
        return getattr(self.lib, f"orc_{p}potrs")(uplo, N, R, nb, nbr, _ptr(A), N, _ptr(B), N)

    def posv(self, uplo, A, B, nb, nbr, p="d"):
        N, R = B.shape
        return getattr(self.lib, f"orc_{p}posv")(uplo, N, R, nb, nbr, _ptr(A), N, _ptr(B), N)

    def lanhe_inf(self, uplo, A, p="d"):
        return getattr(self.lib, f"orc_{p}lanhe_inf")(uplo, A.shape[0], _ptr(A), A.shape[0])

    def lange_inf(self, A, p="d"):
        return getattr(self.lib, f"orc_{p}lange_inf")(A.shape[0], A.shape[1], _ptr(A), A.shape[0])

    def check_potrf(self, uplo, A, A0, p="d"):
        """returns (suspicious(0/1), result, Rnorm, Anorm) -- src/dplasma_zcheck.c:64-134"""
        r, rn, an = _c_dbl(0), _c_dbl(0), _c_dbl(0)
        N = A.shape[0]
        bad = getattr(self.lib, f"orc_{p}check_potrf")(uplo, N, _ptr(A), N, _ptr(A0), N,
                                                       ctypes.byref(r), ctypes.byref(rn), ctypes.byref(an))
        return bad, r.value, rn.value, an.value

    def check_axmb(self, uplo, A, b, x, p="d"):
        """b is overwritten with b - A x like the reference.  returns (suspicious, result)"""
        r = _c_dbl(0)
        N, R = b.shape
        bad = getattr(self.lib, f"orc_{p}check_axmb")(uplo, N, R, _ptr(A), N, _ptr(b), N, _ptr(x), N,
                                                      ctypes.byref(r))
        return bad, r.value

    # ---- inversion: src/ztrtri_wrapper.c, zlauum_wrapper.c, zpotri_wrapper.c, zpoinv_wrapper.c
    def trtri_tile(self, uplo, diag, A, p="d"):
        info = _c_int(0)
        getattr(self.lib, f"orc_{p}trtri_tile")(uplo, diag, A.shape[0], _ptr(A), A.shape[0], ctypes.byref(info))
        return info.value

    def lauum_tile(self, uplo, A, p="d"):
        getattr(self.lib, f"orc_{p}lauum_tile")(uplo, A.shape[0], _ptr(A), A.shape[0])

    def trmm_tile(self, side, uplo, trans, diag, alpha, A, B, p="d"):
        m, n = B.shape
        getattr(self.lib, f"orc_{p}trmm_tile")(side, uplo, trans, diag, m, n, self._sc(p, alpha),
                                               _ptr(A), A.shape[0], _ptr(B), B.shape[0])

    def trtri(self, uplo, diag, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"orc_{p}trtri")(uplo, diag, N, nb, _ptr(A), N)

    def lauum(self, uplo, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"orc_{p}lauum")(uplo, N, nb, _ptr(A), N)

    def potri(self, uplo, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"orc_{p}potri")(uplo, N, nb, _ptr(A), N)

    def poinv(self, uplo, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"orc_{p}poinv")(uplo, N, nb, _ptr(A), N)

    def check_poinv(self, uplo, A, Ainv, p="d"):
        """A is read as a general (full) matrix, Ainv from its uplo triangle.  returns (suspicious, result)
        -- src/dplasma_zcheck.c:265-315"""
        r = _c_dbl(0)
        N = A.shape[0]
        bad = getattr(self.lib, f"orc_{p}check_poinv")(uplo, N, _ptr(A), N, _ptr(Ainv), N, ctypes.byref(r))
        return bad, r.value

    # ---- public BLAS-3 ops: src/zgemm_wrapper.c:645-733, src/zsyrk_wrapper.c / zherk_wrapper.c
    def gemm(self, ta, tb, alpha, A, B, beta, C, mb, nb, kb, p="d"):
        M, N = C.shape
        K = A.shape[1] if ta == NoTrans else A.shape[0]
        return getattr(self.lib, f"orc_{p}gemm")(ta, tb, M, N, K, mb, nb, kb, self._sc(p, alpha), _ptr(A), A.shape[0],
                                                 _ptr(B), B.shape[0], self._sc(p, beta), _ptr(C), M)

    def herk(self, uplo, trans, alpha, A, beta, C, nb, kb, p="d"):
        """d: dsyrk ; z: zherk"""
        N = C.shape[0]
        K = A.shape[1] if trans == NoTrans else A.shape[0]
        return getattr(self.lib, f"orc_{p}herk")(uplo, trans, N, K, nb, kb, float(alpha), _ptr(A), A.shape[0],
                                                 float(beta), _ptr(C), N)


class Reference:
    """The reference's own compiled CPU tile kernels under an OpenMP restatement of its DAG
    (oracle/_ref/libdplasma_ref.so; see oracle/ref_driver.c).  Raises if not built."""

    def __init__(self):
        path = os.path.join(HERE, "_ref", "libdplasma_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = L = ctypes.CDLL(path)
        L.ref_blas_config.restype = ctypes.c_char_p
        L.ref_blas_corename.restype = ctypes.c_char_p
        L.ref_num_threads.restype = _c_int
        for p in "dz":
            gen = "plgsy" if p == "d" else "plghe"
            getattr(L, f"ref_{p}{gen}").argtypes = [_c_dbl, _c_int, _c_int, _c_int, _c_int, _c_int, _vp, _c_int, _c_u64]
            getattr(L, f"ref_{p}{gen}").restype = None
            getattr(L, f"ref_{p}plrnt").argtypes = [_c_int, _c_int, _c_int, _c_int, _vp, _c_int, _c_u64]
            getattr(L, f"ref_{p}plrnt").restype = None
            getattr(L, f"ref_{p}potrf").argtypes = [_c_int, _c_int, _c_int, _vp, _c_int]
            getattr(L, f"ref_{p}potrf").restype = _c_int
        L.ref_dtrsm.argtypes = [_c_int] * 8 + [_c_dbl, _vp, _c_int, _vp, _c_int]
        L.ref_dpotrs.argtypes = [_c_int] * 5 + [_vp, _c_int, _vp, _c_int]
        L.ref_dposv.argtypes = [_c_int] * 5 + [_vp, _c_int, _vp, _c_int]
        if hasattr(L, "ref_dtrtri"):
            L.ref_dtrtri.argtypes = [_c_int, _c_int, _c_int, _c_int, _vp, _c_int]
            for f in ("ref_dlauum", "ref_dpotri", "ref_dpoinv"):
                getattr(L, f).argtypes = [_c_int, _c_int, _c_int, _vp, _c_int]
        if hasattr(L, "ref_dgemm"):
            L.ref_dgemm.argtypes = [_c_int] * 8 + [_c_dbl, _vp, _c_int, _vp, _c_int, _c_dbl, _vp, _c_int]
            L.ref_dsyrk.argtypes = [_c_int] * 6 + [_c_dbl, _vp, _c_int, _c_dbl, _vp, _c_int]
        L.ref_time_dpotrf.restype = _c_dbl
        L.ref_time_dpotrf.argtypes = [_c_int, _c_int, _c_int, _c_u64, ctypes.POINTER(_c_int)]

    def info(self):
        return {"blas": self.lib.ref_blas_config().decode(), "core": self.lib.ref_blas_corename().decode(),
                "threads": self.lib.ref_num_threads()}

    def plgsy(self, bump, uplo, N, nb, seed, p="d"):
        A = np.zeros((N, N), dtype=Oracle._dt(p), order="F")
        gen = "plgsy" if p == "d" else "plghe"
        getattr(self.lib, f"ref_{p}{gen}")(float(bump), uplo, N, N, nb, nb, _ptr(A), N, seed)
        return A

    def plrnt(self, M, N, mb, nb, seed, p="d"):
        A = np.zeros((M, N), dtype=Oracle._dt(p), order="F")
        getattr(self.lib, f"ref_{p}plrnt")(M, N, mb, nb, _ptr(A), M, seed)
        return A

    def potrf(self, uplo, A, nb, p="d"):
        N = A.shape[0]
        return getattr(self.lib, f"ref_{p}potrf")(uplo, N, nb, _ptr(A), N)

    def trsm(self, side, uplo, trans, diag, alpha, A, B, nb, nbr):
        M, R = B.shape
        return self.lib.ref_dtrsm(side, uplo, trans, diag, M, R, nb, nbr, float(alpha), _ptr(A),
                                  A.shape[0], _ptr(B), M)

    def potrs(self, uplo, A, B, nb, nbr):
        N, R = B.shape
        return self.lib.ref_dpotrs(uplo, N, R, nb, nbr, _ptr(A), N, _ptr(B), N)

    def posv(self, uplo, A, B, nb, nbr):
        N, R = B.shape
        return self.lib.ref_dposv(uplo, N, R, nb, nbr, _ptr(A), N, _ptr(B), N)

    def trtri(self, uplo, diag, A, nb):
        N = A.shape[0]
        return self.lib.ref_dtrtri(uplo, diag, N, nb, _ptr(A), N)

    def lauum(self, uplo, A, nb):
        N = A.shape[0]
        return self.lib.ref_dlauum(uplo, N, nb, _ptr(A), N)

    def potri(self, uplo, A, nb):
        N = A.shape[0]
        return self.lib.ref_dpotri(uplo, N, nb, _ptr(A), N)

    def poinv(self, uplo, A, nb):
        N = A.shape[0]
        return self.lib.ref_dpoinv(uplo, N, nb, _ptr(A), N)

    def gemm(self, ta, tb, alpha, A, B, beta, C, mb, nb, kb):
        M, N = C.shape
        K = A.shape[1] if ta == NoTrans else A.shape[0]
        return self.lib.ref_dgemm(ta, tb, M, N, K, mb, nb, kb, float(alpha), _ptr(A), A.shape[0], _ptr(B), B.shape[0],
                                  float(beta), _ptr(C), M)

    def syrk(self, uplo, trans, alpha, A, beta, C, nb, kb):
        N = C.shape[0]
        K = A.shape[1] if trans == NoTrans else A.shape[0]
        return self.lib.ref_dsyrk(uplo, trans, N, K, nb, kb, float(alpha), _ptr(A), A.shape[0], float(beta), _ptr(C), N)

    def time_dpotrf(self, uplo, N, nb, seed=3872):
        """seconds for the factorization only (generation excluded), and info"""
        info = _c_int(0)
        t = self.lib.ref_time_dpotrf(uplo, N, nb, seed, ctypes.byref(info))
        return t, info.value
