"""Profiling driver: the DMMA GEMM engine on one large NT product shaped like a trailing update
(C[16384x16384] -= A[16384x512] B[16384x512]^T) through the C-ABI kernel seam, a few launches."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp

M = N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
K = int(sys.argv[2]) if len(sys.argv) > 2 else 512
lib = dp.load()
ctx = dp.Context(lib)
A = torch.randn(K, M, dtype=torch.float64, device="cuda")   # column-major M x K
B = torch.randn(K, N, dtype=torch.float64, device="cuda")
Cm = torch.randn(N, M, dtype=torch.float64, device="cuda")
p = lambda t: C.c_void_p(t.data_ptr())
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(5):
    if it == 2:
        ev[0].record()
    rc = lib.dplasma_b200_dgemm_tile(ctx.ptr, None, dp.NoTrans, dp.Trans, M, N, K, -1.0, p(A), M, p(B), N, 1.0, p(Cm), M)
    assert rc == 0
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / 3
print(f"dgemm_tile NT M=N={M} K={K}: {ms:.3f} ms/launch, {2.0*M*N*K/ms/1e9:.2f} TFLOP/s (default stream events incl. launch)")
ctx.close()
