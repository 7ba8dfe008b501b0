"""Small end-to-end pass over every kernel family (for compute-sanitizer memcheck)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
def run(N, nb, K, nbr, uplo, kind="tile", z=False):
    A = dp.BlockCyclic(lib, N, N, nb, nb, kind=kind, uplo=uplo if kind == "sym" else dp.UpperLower, where="host", cplx=z)
    A0 = dp.BlockCyclic(lib, N, N, nb, nb, where="host", cplx=z)
    gen = lib.dplasma_zplghe if z else lib.dplasma_dplgsy
    assert gen(ctx.ptr, float(N), uplo, A.ref, 3872) == 0 and gen(ctx.ptr, float(N), uplo, A0.ref, 3872) == 0
    info = (lib.dplasma_zpotrf if z else lib.dplasma_dpotrf)(ctx.ptr, uplo, A.ref)
    bad = (lib.check_zpotrf if z else lib.check_dpotrf)(ctx.ptr, 0, uplo, A.ref, A0.ref) if kind == "tile" else 0
    r1 = lib.dplasma_b200_last_check_result(ctx.ptr)
    r2 = -1.0
    if not z and kind == "tile":
        B = dp.BlockCyclic(lib, N, K, nb, nbr, where="host"); X = dp.BlockCyclic(lib, N, K, nb, nbr, where="host")
        lib.dplasma_dplrnt(ctx.ptr, 0, B.ref, 7); lib.dplasma_dplrnt(ctx.ptr, 0, X.ref, 7)
        assert lib.dplasma_dpotrs(ctx.ptr, uplo, A.ref, X.ref) == 0
        bad |= lib.check_daxmb(ctx.ptr, 0, uplo, A0.ref, B.ref, X.ref); r2 = lib.dplasma_b200_last_check_result(ctx.ptr)
    print(f"N={N} nb={nb} uplo={uplo} kind={kind} z={z}: info={info} bad={bad} res={r1:.3g} solve={r2:.3g}")
    assert info == 0 and bad == 0
run(378, 93, 37, 16, dp.Lower); run(378, 93, 37, 16, dp.Upper)
run(1024, 128, 64, 64, dp.Lower); run(1024, 128, 64, 64, dp.Lower, kind="sym")
run(640, 128, 64, 64, dp.Upper); run(300, 64, 9, 9, dp.Lower, z=True); run(512, 128, 9, 9, dp.Lower, z=True)
ctx.close(); print("sanity ok")
