"""e2e dpotrf on a pinned-host sym matrix, per-iteration wall time, next to the raw pinned copy bandwidth of the box."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
N, nb = 32768, 512
lib = dp.load(); ctx = dp.Context(lib)
x = torch.empty(1 << 29, dtype=torch.float64).pin_memory(); y = torch.empty(1 << 29, dtype=torch.float64, device="cuda")
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); y.copy_(x, non_blocking=True); torch.cuda.synchronize(); t1 = time.perf_counter()
    x.copy_(y, non_blocking=True); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"pinned H2D {4.295/(t1-t0):.1f} GB/s, D2H {4.295/(t2-t1):.1f} GB/s")
del x, y
h = dp.BlockCyclic(lib, N, N, nb, nb, kind="sym", uplo=dp.Lower, where="host")
for it in range(4):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, h.ref, 3872) == 0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, h.ref)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"it {it}: e2e {dt*1e3:.1f} ms ({lib.dplasma_b200_flops_dpotrf(float(N))/dt/1e12:.2f} TF/s), device part {lib.dplasma_b200_last_elapsed(ctx.ptr)*1e3:.1f} ms, info {info}")
ctx.close()
