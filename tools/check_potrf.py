"""Device-resident dplgsy -> dpotrf -> check_dpotrf at a given size; prints time and residual."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dplasma_b200 as dp

N = int(sys.argv[1]); nb = int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
z = len(sys.argv) > 4 and sys.argv[4] == "z"
uplo = dp.Lower
lib = dp.load(); ctx = dp.Context(lib)
A = dp.BlockCyclic(lib, N, N, nb, nb, where="device", cplx=z); A0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device", cplx=z)
gen = lib.dplasma_zplghe if z else lib.dplasma_dplgsy
fac = lib.dplasma_zpotrf if z else lib.dplasma_dpotrf
chk = lib.check_zpotrf if z else lib.check_dpotrf
fl = lib.dplasma_b200_flops_dpotrf(float(N)) * (4.0 if z else 1.0)   # FLOPS_ZPOTRF ~ 4 N^3/3 (src/flops.h:270-273)
assert gen(ctx.ptr, float(N), uplo, A0.ref, 3872) == 0
for r in range(reps):
    assert gen(ctx.ptr, float(N), uplo, A.ref, 3872) == 0
    info = fac(ctx.ptr, uplo, A.ref)
    t = lib.dplasma_b200_last_elapsed(ctx.ptr)
    bad = chk(ctx.ptr, 0, uplo, A.ref, A0.ref)
    print(f"{'z' if z else 'd'}potrf N={N} nb={nb} cfg={os.environ.get('DPB_GEMM_CFG','auto')} info={info} t={t*1e3:.2f} ms {fl/t/1e12:.2f} TF/s residual={lib.dplasma_b200_last_check_result(ctx.ptr):.4g} bad={bad}")
ctx.close()
