import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
p = lambda t: C.c_void_p(t.data_ptr())
torch.manual_seed(0)
M = N = int(sys.argv[1]); K = int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
A = torch.randn(K, M, dtype=torch.float64, device="cuda"); B = torch.randn(K, N, dtype=torch.float64, device="cuda")
C0 = torch.randn(N, M, dtype=torch.float64, device="cuda"); want = C0 - (B.T @ A)
for it in range(reps):
    Cm = C0.clone()
    assert lib.dplasma_b200_dgemm_tile(ctx.ptr, None, dp.NoTrans, dp.Trans, M, N, K, -1.0, p(A), M, p(B), N, 1.0, p(Cm), M) == 0
    torch.cuda.synchronize()
    err = (Cm - want).abs() / want.abs().max()
    bad = (err > 1e-10).nonzero()
    print(f"cfg={os.environ.get('DPB_GEMM_CFG','auto')} M=N={M} K={K} max rel err {err.max().item():.3e} bad elems {bad.shape[0]}")
    if bad.shape[0]:
        cols = bad[:, 0]; rows = bad[:, 1]
        print("  bad rows(mod 64) hist:", torch.bincount(rows % 64, minlength=64).tolist())
        print("  bad cols(mod 64) hist:", torch.bincount(cols % 64, minlength=64).tolist())
        print("  bad CTA tiles:", len(set(((rows // 64) * 100000 + cols // 64).tolist())), "first:", bad[:5].tolist())
ctx.close()
