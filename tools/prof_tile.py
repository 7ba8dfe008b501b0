import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
G = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
A = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for it in range(3):
    T = A.clone()
    assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, C.c_void_p(T.data_ptr()), n, C.c_void_p(info.data_ptr()), 0) == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
T = A.clone(); e0.record()
lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, C.c_void_p(T.data_ptr()), n, C.c_void_p(info.data_ptr()), 0)
e1.record(); torch.cuda.synchronize()
print(f"dpotrf_tile n={n}: {e0.elapsed_time(e1)*1e3:.1f} us")
ctx.close()
