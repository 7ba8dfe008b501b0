"""Work-order A/B driver: one dpotrf at BASELINE configs[1] (N=32768, NB=512), device-resident, with the event-timed
bulk trailing-update launches.  Run it with DPB_GEMM_ORDER=0 (strip-by-strip order of round 1) and without (grouped
order); `--once` does a single factorization (for an ncu metrics pass over its first gemm_dmma launches)."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp

once = "--once" in sys.argv
N, nb = 32768, 512
lib = dp.load()
ctx = dp.Context(lib)
d = dp.BlockCyclic(lib, N, N, nb, nb, 1, 1, 0, where="device")
ts = []
for it in range(1 if once else 4):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, d.ref, 3872) == 0
    if it == 3:
        lib.dplasma_b200_profile(ctx.ptr, 1)
    assert lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref) == 0
    ts.append(lib.dplasma_b200_last_elapsed(ctx.ptr))
flops = N ** 3 / 3.0 + N * N / 2.0 + N / 6.0
msg = f"order={os.environ.get('DPB_GEMM_ORDER', 'grouped')} dpotrf N={N} NB={nb}: " + ", ".join(f"{t*1e3:.2f} ms" for t in ts)
msg += f"; best unprofiled {flops / min(ts[:3]) / 1e12:.2f} TFLOP/s"
if not once:
    ps, pf, pl, pb = C.c_double(0), C.c_double(0), C.c_int(0), C.c_double(0)
    lib.dplasma_b200_profile_read_stream(ctx.ptr, 1, C.byref(ps), C.byref(pf), C.byref(pl), C.byref(pb))
    lib.dplasma_b200_profile(ctx.ptr, 0)
    if ps.value > 0:
        msg += f"; bulk launches: {pl.value}, {pf.value / ps.value / 1e12:.2f} TFLOP/s in situ, best {pb.value:.2f}"
print(msg)
d.free()
ctx.close()
