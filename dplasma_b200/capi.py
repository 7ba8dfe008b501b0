"""ctypes binding of include/dplasma_b200.h (struct layouts and prototypes kept in the same
order as the header)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libdplasma_b200.so")

# src/include/dplasma/constants.h:78-96
NoTrans, Trans, ConjTrans = 111, 112, 113
Upper, Lower, UpperLower = 121, 122, 123
NonUnit, Unit = 131, 132
Left, Right = 141, 142
InfNorm = 175
MATRIX_DOUBLE, MATRIX_COMPLEX_DOUBLE = 4, 6
MATRIX_LAPACK, MATRIX_TILE = 0, 1


class DataCollection(C.Structure):
    _fields_ = [("nodes", C.c_uint32), ("myrank", C.c_uint32)]


class TiledMatrix(C.Structure):
    _fields_ = [("super", DataCollection)] + [(n, C.c_int) for n in (
        "mtype", "storage", "mb", "nb", "bsiz", "lm", "ln", "lmt", "lnt", "llm", "lln", "i", "j",
        "m", "n", "mt", "nt", "slm", "sln", "nb_local_tiles")]


class Grid(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("rank", "rows", "cols", "rrank", "crank", "krows", "kcols", "ip", "jq")]


class BlockCyclicStruct(C.Structure):
    _fields_ = [("super", TiledMatrix), ("grid", Grid), ("mat", C.c_void_p),
                ("nb_elem_r", C.c_int), ("nb_elem_c", C.c_int), ("uplo", C.c_int)]


def build(verbose: bool = False) -> str:
    """Compile libdplasma_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if verbose or out.returncode:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode:
        raise RuntimeError("building libdplasma_b200.so failed")
    return LIB_PATH


class Lib:
    def __init__(self, path: str = LIB_PATH):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        self.dll = L = C.CDLL(path, mode=C.RTLD_GLOBAL)
        vp, i, d, u64 = C.c_void_p, C.c_int, C.c_double, C.c_ulonglong
        pBC = C.POINTER(BlockCyclicStruct)

        def proto(name, res, args):
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
            return f
        proto("dplasma_b200_version", C.c_char_p, [])
        proto("parsec_init", vp, [i, vp, vp])
        proto("parsec_fini", i, [C.POINTER(vp)])
        proto("dplasma_b200_get_unique_id", i, [vp])
        proto("dplasma_b200_comm_init", i, [vp, vp, i, i])
        proto("parsec_context_add_taskpool", i, [vp, vp])
        proto("parsec_context_start", i, [vp])
        proto("parsec_context_wait", i, [vp])
        proto("parsec_data_allocate", vp, [C.c_size_t])
        proto("parsec_data_free", None, [vp])
        proto("dplasma_b200_device_allocate", vp, [C.c_size_t])
        proto("dplasma_b200_device_free", None, [vp])
        proto("parsec_matrix_block_cyclic_init", None, [pBC] + [i] * 17)
        proto("parsec_matrix_sym_block_cyclic_init", None, [pBC] + [i] * 13)
        proto("parsec_matrix_block_cyclic_lapack_init", None, [pBC] + [i] * 19)
        proto("parsec_tiled_matrix_destroy", None, [pBC])
        proto("dplasma_b200_rank_of", i, [pBC, i, i])
        proto("dplasma_b200_panel_plan", i, [pBC, i, i, C.POINTER(i), C.POINTER(i), C.POINTER(i)])
        proto("dplasma_b200_rowpanel_plan", i, [pBC, i, C.POINTER(i), C.POINTER(i), C.POINTER(i)])
        proto("dplasma_b200_transpose_cap", C.c_longlong, [i, C.c_longlong])
        proto("dplasma_b200_transpose_plan", i, [i, i, i, i, C.c_longlong, i, i, C.POINTER(i), C.POINTER(i), i, C.POINTER(i),
                                                 C.POINTER(i), i, C.POINTER(i)])
        proto("dplasma_dpotrf_New", vp, [i, pBC, C.POINTER(i)])
        proto("dplasma_dpotrf_Destruct", None, [vp])
        proto("dplasma_dpotrf_Destroy", None, [vp])
        proto("dplasma_dpotrf", i, [vp, i, pBC])
        proto("dplasma_dpotrf_rec", i, [vp, i, pBC, i])
        proto("dplasma_dpotrf_setrecursive", None, [vp, i])
        proto("dplasma_dposv", i, [vp, i, pBC, pBC])
        proto("dplasma_dpotrs", i, [vp, i, pBC, pBC])
        proto("dplasma_dtrsm", i, [vp, i, i, i, i, d, pBC, pBC])
        proto("dplasma_dplgsy", i, [vp, d, i, pBC, u64])
        proto("dplasma_dplrnt", i, [vp, i, pBC, u64])
        proto("check_dpotrf", i, [vp, i, i, pBC, pBC])
        proto("check_daxmb", i, [vp, i, i, pBC, pBC, pBC])
        proto("dplasma_dtrtri", i, [vp, i, i, pBC])
        proto("dplasma_dlauum", i, [vp, i, pBC])
        proto("dplasma_dpotri", i, [vp, i, pBC])
        proto("dplasma_dpoinv", i, [vp, i, pBC])
        proto("dplasma_dpoinv_sync", i, [vp, i, pBC])
        proto("check_dpoinv", i, [vp, i, i, pBC, pBC])
        proto("dplasma_dgemm", i, [vp, i, i, d, pBC, pBC, d, pBC])
        proto("dplasma_dsyrk", i, [vp, i, i, d, pBC, d, pBC])
        proto("dplasma_dtrsm_New", vp, [i, i, i, i, d, pBC, pBC])
        proto("dplasma_dtrtri_New", vp, [i, i, pBC, C.POINTER(i)])
        proto("dplasma_dlauum_New", vp, [i, pBC])
        proto("dplasma_dpoinv_New", vp, [i, pBC, C.POINTER(i)])
        proto("dplasma_dgemm_New", vp, [i, i, d, pBC, pBC, d, pBC])
        proto("dplasma_dsyrk_New", vp, [i, i, d, pBC, d, pBC])
        for _n in ("trsm", "trtri", "lauum", "poinv", "gemm", "syrk"):
            proto(f"dplasma_d{_n}_Destruct", None, [vp])
        proto("dplasma_zpotrf_New", vp, [i, pBC, C.POINTER(i)])
        proto("dplasma_zpotrf_Destruct", None, [vp])
        proto("dplasma_zpotrf_Destroy", None, [vp])
        proto("dplasma_zpotrf", i, [vp, i, pBC])
        proto("dplasma_zplghe", i, [vp, d, i, pBC, u64])
        proto("dplasma_zplrnt", i, [vp, i, pBC, u64])
        proto("check_zpotrf", i, [vp, i, i, pBC, pBC])
        proto("dplasma_zlanhe", d, [vp, i, i, pBC])
        proto("dplasma_zlange", d, [vp, i, pBC])
        proto("dplasma_zposv", i, [vp, i, pBC, pBC])
        proto("dplasma_zpotrs", i, [vp, i, pBC, pBC])
        proto("check_zaxmb", i, [vp, i, i, pBC, pBC, pBC])
        proto("dplasma_b200_last_check_result", d, [vp])
        proto("dplasma_dlansy", d, [vp, i, i, pBC])
        proto("dplasma_dlange", d, [vp, i, pBC])
        proto("dplasma_b200_last_elapsed", d, [vp])
        proto("dplasma_b200_kernel_launches", C.c_longlong, [vp])
        proto("dplasma_b200_profile", None, [vp, i])
        proto("dplasma_b200_profile_read", i, [vp, C.POINTER(d), C.POINTER(d), C.POINTER(i), C.POINTER(d)])
        proto("dplasma_b200_profile_read_stream", i, [vp, i, C.POINTER(d), C.POINTER(d), C.POINTER(i), C.POINTER(d)])
        proto("dplasma_b200_dpotrf_tile", i, [vp, vp, i, i, vp, i, vp, i])
        proto("dplasma_b200_dpotrf_tile_noalloc", i, [vp, i, i, vp, i, vp, i])
        proto("dplasma_b200_dtrsm_tile", i, [vp, vp, i, i, i, i, i, i, d, vp, i, vp, i])
        proto("dplasma_b200_dsyrk_tile", i, [vp, vp, i, i, i, i, d, vp, i, d, vp, i])
        proto("dplasma_b200_dgemm_tile", i, [vp, vp, i, i, i, i, i, d, vp, i, vp, i, d, vp, i])
        proto("pdpotrf_", None, [C.c_char_p, C.POINTER(i), vp, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(i)])
        proto("pdgemm_", None, [C.c_char_p] * 2 + [C.POINTER(i)] * 3 + [C.POINTER(d), vp, C.POINTER(i), C.POINTER(i), C.POINTER(i),
                                vp, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(d), vp, C.POINTER(i), C.POINTER(i),
                                C.POINTER(i)])
        proto("pdtrsm_", None, [C.c_char_p] * 4 + [C.POINTER(i), C.POINTER(i), C.POINTER(d), vp, C.POINTER(i), C.POINTER(i),
                                C.POINTER(i), vp, C.POINTER(i), C.POINTER(i), C.POINTER(i)])
        proto("parsec_init_wrapper_", None, [])
        proto("parsec_fini_wrapper_", None, [])
        proto("dplasma_b200_blacs_gridinit", None, [i] * 5)
        proto("dplasma_b200_wrapper_set_context", None, [vp])
        proto("dplasma_b200_dmma_peak_tflops", d, [i, i, C.POINTER(d)])
        proto("dplasma_b200_flops_dpotrf", d, [d])
        proto("dplasma_b200_flops_dpotrs", d, [d, d])

    def __getattr__(self, name):
        return getattr(self.dll, name)


_LIB = None


def load() -> Lib:
    global _LIB
    if _LIB is None:
        _LIB = Lib()
    return _LIB


class Context:
    """parsec_context_t: one per process, bound to the current CUDA device."""

    def __init__(self, lib: Lib | None = None):
        self.lib = lib or load()
        self.ptr = self.lib.parsec_init(-1, None, None)
        if not self.ptr:
            raise RuntimeError("parsec_init failed: no CUDA device (dplasma_b200 has no CPU fallback)")

    def comm_init(self, id_bytes: bytes, rank: int, nranks: int):
        buf = C.create_string_buffer(id_bytes, 128)
        rc = self.lib.dplasma_b200_comm_init(self.ptr, buf, rank, nranks)
        if rc:
            raise RuntimeError(f"dplasma_b200_comm_init failed ({rc})")

    def unique_id(self) -> bytes:
        buf = C.create_string_buffer(128)
        rc = self.lib.dplasma_b200_get_unique_id(buf)
        if rc:
            raise RuntimeError(f"dplasma_b200_get_unique_id failed ({rc})")
        return buf.raw

    def close(self):
        if self.ptr:
            p = C.c_void_p(self.ptr)
            self.lib.parsec_fini(C.byref(p))
            self.ptr = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class BlockCyclic:
    """A parsec_matrix_block_cyclic_t (or its sym / LAPACK-storage variants) plus its buffer.

    where: "host" -> parsec_data_allocate (pinned), "device" -> dplasma_b200_device_allocate,
    or an int address of caller memory."""

    def __init__(self, lib: Lib, M, N, mb, nb, P=1, Q=1, rank=0, kind="tile", uplo=UpperLower, where="host",
                 lld=None, cplx=False):
        self.lib = lib
        self.s = BlockCyclicStruct()
        self.kind = kind
        self.cplx = cplx
        MATRIX_DOUBLE = MATRIX_COMPLEX_DOUBLE if cplx else 4  # noqa: N806 (shadows the module constant on purpose)
        if kind == "tile":
            lib.parsec_matrix_block_cyclic_init(C.byref(self.s), MATRIX_DOUBLE, MATRIX_TILE, rank, mb, nb, M, N,
                                                0, 0, M, N, P, Q, 1, 1, 0, 0)
            self.elems = self.s.super.nb_local_tiles * self.s.super.bsiz
        elif kind == "sym":
            lib.parsec_matrix_sym_block_cyclic_init(C.byref(self.s), MATRIX_DOUBLE, rank, mb, nb, M, N, 0, 0, M, N,
                                                    P, Q, uplo)
            self.elems = self.s.super.nb_local_tiles * self.s.super.bsiz
        elif kind == "lapack":
            lib.parsec_matrix_block_cyclic_init(C.byref(self.s), MATRIX_DOUBLE, MATRIX_LAPACK, rank, mb, nb, M, N,
                                                0, 0, M, N, P, Q, 1, 1, 0, 0)
            mloc, nloc = self.s.super.llm, self.s.super.lln
            lld = max(1, mloc) if lld is None else lld
            lib.parsec_matrix_block_cyclic_lapack_init(C.byref(self.s), MATRIX_DOUBLE, MATRIX_LAPACK, rank, mb, nb,
                                                       M, N, 0, 0, M, N, P, Q, 1, 1, 0, 0, lld, nloc)
            self.elems = lld * max(1, nloc)
        else:
            raise ValueError(kind)
        self.owned = None
        if cplx:
            self.elems *= 2  # counted in doubles
        nbytes = max(8, self.elems * 8)
        if where == "host":
            self.s.mat = lib.parsec_data_allocate(nbytes)
            self.owned = "host"
        elif where == "device":
            self.s.mat = lib.dplasma_b200_device_allocate(nbytes)
            self.owned = "device"
        else:
            self.s.mat = int(where)
        if not self.s.mat:
            raise MemoryError("allocation failed")

    @property
    def ref(self):
        return C.byref(self.s)

    def host_array(self):
        """numpy view of a host buffer (flat, float64)."""
        import numpy as np
        assert self.owned == "host"
        return np.ctypeslib.as_array(C.cast(self.s.mat, C.POINTER(C.c_double)), shape=(max(1, self.elems),))

    def free(self):
        if self.owned == "host":
            self.lib.parsec_data_free(self.s.mat)
        elif self.owned == "device":
            self.lib.dplasma_b200_device_free(self.s.mat)
        self.owned = None
        self.s.mat = None
