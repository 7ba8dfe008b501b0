"""Accuracy of the tile Cholesky vs a float128-free reference: compares with torch.linalg.cholesky (cuSOLVER fp64)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
p = lambda t: C.c_void_p(t.data_ptr())
torch.manual_seed(1)
for n in (64, 128, 512, 1024):
    G = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
    A = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
    L = torch.linalg.cholesky(A)
    T = A.clone()   # symmetric, so row-major == col-major
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, p(T), n, p(info), 0) == 0
    torch.cuda.synchronize()
    mine = torch.tril(T.T)   # buffer is column-major: element (i,j) at T[j,i]
    err = (mine - L).abs().max().item()
    res = (mine @ mine.T - A).abs().max().item()
    res_ref = (L @ L.T - A).abs().max().item()
    print(f"n={n} mode={os.environ.get('DPB_POTRF_TILE','fused')} max|L-Lref|={err:.3e} max|LL^T-A|={res:.3e} (cusolver {res_ref:.3e}) info={int(info.item())}")
ctx.close()
