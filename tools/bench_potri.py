"""dpotrf / dtrtri / dlauum / dpoinv on a device-resident matrix: GPU seconds (CUDA events inside the
library) and TFLOP/s with the reference's flop model (src/flops.h: FLOPS_DPOTRF, FLOPS_DTRTRI = n^3/3 + 2n/3,
FLOPS_DLAUUM = FLOPS_DPOTRF), then the reference's check_dpoinv on the result.
usage: python tools/bench_potri.py N nb [L|U]      (POTRI_PROFILE=1: one repetition of potrf, trtri, lauum only --
the command profiled under ncu for profiles/launches_potri_*.csv)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dplasma_b200 as dp  # noqa: E402

N = int(sys.argv[1]); nb = int(sys.argv[2])
uplo = dp.Upper if (len(sys.argv) > 3 and sys.argv[3] == "U") else dp.Lower
lib = dp.load(); ctx = dp.Context(lib)
A = dp.BlockCyclic(lib, N, N, nb, nb, where="device"); A0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
f_potrf = lib.dplasma_b200_flops_dpotrf(float(N))
f_trtri = N ** 3 / 3.0 + 2.0 * N / 3.0
PROFILE = os.environ.get("POTRI_PROFILE") == "1"
for rep in range(1 if PROFILE else 2):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.UpperLower, A.ref, 3872) == 0
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.UpperLower, A0.ref, 3872) == 0
    l0 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    assert lib.dplasma_dpotrf(ctx.ptr, uplo, A.ref) == 0
    t_f = lib.dplasma_b200_last_elapsed(ctx.ptr); l1 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    assert lib.dplasma_dtrtri(ctx.ptr, uplo, dp.NonUnit, A.ref) == 0
    t_t = lib.dplasma_b200_last_elapsed(ctx.ptr); l2 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    assert lib.dplasma_dlauum(ctx.ptr, uplo, A.ref) == 0
    t_l = lib.dplasma_b200_last_elapsed(ctx.ptr); l3 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    if PROFILE:
        print(f"N={N} nb={nb}: potrf {t_f * 1e3:.1f} ms ({l1 - l0} launches) | trtri {t_t * 1e3:.1f} ms ({l2 - l1}) | "
              f"lauum {t_l * 1e3:.1f} ms ({l3 - l2})", flush=True)
        break
    bad = lib.check_dpoinv(ctx.ptr, 0, uplo, A0.ref, A.ref)
    res = lib.dplasma_b200_last_check_result(ctx.ptr)
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.UpperLower, A.ref, 3872) == 0
    assert lib.dplasma_dpoinv(ctx.ptr, uplo, A.ref) == 0
    t_p = lib.dplasma_b200_last_elapsed(ctx.ptr)
    tot = f_potrf + f_trtri + f_potrf
    print(f"N={N} nb={nb} uplo={'U' if uplo == dp.Upper else 'L'}: "
          f"potrf {t_f * 1e3:.1f} ms {f_potrf / t_f / 1e12:.2f} TF/s ({l1 - l0} launches) | "
          f"trtri {t_t * 1e3:.1f} ms {f_trtri / t_t / 1e12:.2f} TF/s ({l2 - l1}) | "
          f"lauum {t_l * 1e3:.1f} ms {f_potrf / t_l / 1e12:.2f} TF/s ({l3 - l2}) | "
          f"poinv {t_p * 1e3:.1f} ms {tot / t_p / 1e12:.2f} TF/s | check_dpoinv bad={bad} result={res:.3g}", flush=True)
A.free(); A0.free(); ctx.close()
