"""Where does the factorization error sit?  Compares dpotrf (device) with a float64 torch Cholesky per tile."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import dplasma_b200 as dp
from tests.util import local_to_full
N = int(sys.argv[1]); nb = int(sys.argv[2])
lib = dp.load(); ctx = dp.Context(lib)
d = dp.BlockCyclic(lib, N, N, nb, nb, where=1)
t = torch.zeros(d.elems, dtype=torch.float64, device="cuda"); d.s.mat = t.data_ptr()
assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.UpperLower, d.ref, 3872) == 0
torch.cuda.synchronize()
A = torch.from_numpy(local_to_full(t.cpu().numpy(), N, N, nb, nb)).cuda()
assert lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref) == 0
torch.cuda.synchronize()
L = torch.tril(torch.from_numpy(local_to_full(t.cpu().numpy(), N, N, nb, nb)).cuda())
Lref = torch.linalg.cholesky(A)
E = (L - Lref).abs()
R = (L @ L.T - A).abs()
Rref = (Lref @ Lref.T - A).abs()
nt = N // nb
print(f"mode={os.environ.get('DPB_POTRF_TILE','fused')} N={N} nb={nb}: max|L-Lref|={E.max().item():.3e}  max|LL^T-A|={R.max().item():.3e} (torch {Rref.max().item():.3e})  rowsum-inf {R.sum(1).max().item():.3e} (torch {Rref.sum(1).max().item():.3e})")
for m in range(nt):
    print("  tile row", m, " max|R| per tile:", " ".join(f"{R[m*nb:(m+1)*nb, n*nb:(n+1)*nb].max().item():.1e}" for n in range(m + 1)))
ctx.close()
