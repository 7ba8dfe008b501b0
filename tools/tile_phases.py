import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
G = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
A = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for it in range(3):
    T = A.clone()
    lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, C.c_void_p(T.data_ptr()), n, C.c_void_p(info.data_ptr()), 0)
torch.cuda.synchronize()
buf = (C.c_longlong * 128)()
lib.dll.dplasma_b200_debug_read(buf, 128)
a = np.array(list(buf)).reshape(16, 8)
print("per launch (last CTA), cycles: K-loop, form D/R, leaf, RX^T+store ; total")
for j in range(n // 64):
    r = a[j]
    print(j, r[1]-r[0], r[2]-r[1], r[3]-r[2], r[4]-r[3], r[4]-r[0])
ctx.close()
lib2 = dp.load()
b2 = (C.c_longlong * 8)()
lib2.dll.dplasma_b200_debug_read_leaf(b2)
l = list(b2)
print("leaf steps (cycles): L21+X11", l[1]-l[0], " syrk", l[2]-l[1], " chol32#2", l[3]-l[2], " X22+T", l[4]-l[3], "; leaf start->after chol#1:", l[0]-a[n//64-1][2])
