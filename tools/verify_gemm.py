"""Stress check of the DMMA engine against torch (cuBLAS) fp64 at sizes with many CTA waves."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
path = os.environ.get("DPB_LIB")
lib = dp.Lib(path) if path else dp.load()
ctx = dp.Context(lib)
p = lambda t: C.c_void_p(t.data_ptr())
torch.manual_seed(0)
for (M, N, K) in [(4096, 4096, 512), (8192, 8192, 512), (8192, 512, 512), (8192, 64, 64)]:
    A = torch.randn(K, M, dtype=torch.float64, device="cuda")
    B = torch.randn(K, N, dtype=torch.float64, device="cuda")
    C0 = torch.randn(N, M, dtype=torch.float64, device="cuda")
    want = C0 - (B.T @ A)            # (C - A B^T)^T in row-major view
    worst = 0.0
    for it in range(4):
        Cm = C0.clone()
        assert lib.dplasma_b200_dgemm_tile(ctx.ptr, None, dp.NoTrans, dp.Trans, M, N, K, -1.0, p(A), M, p(B), N, 1.0, p(Cm), M) == 0
        torch.cuda.synchronize()
        worst = max(worst, ((Cm - want).abs().max() / want.abs().max()).item())
    print(f"cfg={os.environ.get('DPB_GEMM_CFG','auto')} lib={os.path.basename(path) if path else 'default'} M={M} N={N} K={K} max rel err {worst:.3e}")
    # in-place right-multiply (the TRSM leaf form): X <- X W^T with W 64x64 lower
if True:
    M = 8192 * 4
    X0 = torch.randn(64, M, dtype=torch.float64, device="cuda")
    W = torch.tril(torch.randn(64, 64, dtype=torch.float64, device="cuda")).T.contiguous()  # col-major lower
    want = (W.T @ X0)   # row-major view of X W^T ... (X W^T)^T = W X^T
    worst = 0.0
    for it in range(4):
        X = X0.clone()
        assert lib.dplasma_b200_dgemm_tile(ctx.ptr, None, dp.NoTrans, dp.Trans, M, 64, 64, 1.0, p(X), M, p(W), 64, 0.0, p(X), M) == 0
        torch.cuda.synchronize()
        worst = max(worst, ((X - want).abs().max() / want.abs().max()).item())
    print(f"  in-place leaf form M={M}: max rel err {worst:.3e}")
ctx.close()
