"""dposv = dpotrf + dpotrs on device-resident matrices; prints times and the solve residual."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dplasma_b200 as dp

N = int(sys.argv[1]); nb = int(sys.argv[2]); K = int(sys.argv[3]); nbr = int(sys.argv[4]) if len(sys.argv) > 4 else nb
uplo = dp.Upper if (len(sys.argv) > 5 and sys.argv[5] == "U") else dp.Lower
lib = dp.load(); ctx = dp.Context(lib)
A = dp.BlockCyclic(lib, N, N, nb, nb, where="device"); A0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
B = dp.BlockCyclic(lib, N, K, nb, nbr, where="device"); X = dp.BlockCyclic(lib, N, K, nb, nbr, where="device")
for rep in range(2):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, A.ref, 3872) == 0
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, A0.ref, 3872) == 0
    assert lib.dplasma_dplrnt(ctx.ptr, 0, B.ref, 2354) == 0
    assert lib.dplasma_dplrnt(ctx.ptr, 0, X.ref, 2354) == 0
    t0 = time.perf_counter(); info = lib.dplasma_dpotrf(ctx.ptr, uplo, A.ref); t1 = time.perf_counter()
    rc = lib.dplasma_dpotrs(ctx.ptr, uplo, A.ref, X.ref); t2 = time.perf_counter()
    bad = lib.check_daxmb(ctx.ptr, 0, uplo, A0.ref, B.ref, X.ref); t3 = time.perf_counter()
    fs = lib.dplasma_b200_flops_dpotrs(float(N), float(K))
    print(f"N={N} nb={nb} NRHS={K} nbr={nbr} uplo={'U' if uplo == dp.Upper else 'L'}: potrf {1e3*(t1-t0):.1f} ms, potrs {1e3*(t2-t1):.1f} ms "
          f"({fs/(t2-t1)/1e12:.2f} TF/s), check_axmb {1e3*(t3-t2):.1f} ms, info={info} rc={rc} residual={lib.dplasma_b200_last_check_result(ctx.ptr):.4g} bad={bad}")
ctx.close()
