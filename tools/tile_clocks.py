"""Phase clocks of the 64x64 leaf inside the single-launch tile factorization (library built with -DDPB_DEBUG_CLOCKS)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dplasma_b200 as dp
lib = dp.Lib(os.path.join(os.path.dirname(dp.capi.LIB_PATH), "libdplasma_b200_dbg.so"))
ctx = dp.Context(lib)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
G = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
A = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(4):
    T = A.clone()
    e0.record()
    assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, C.c_void_p(T.data_ptr()), n, C.c_void_p(info.data_ptr()), 0) == 0
    e1.record()
torch.cuda.synchronize()
leaf = (C.c_longlong * 8)()
lib.dll.dplasma_b200_debug_read_tile_leaf.argtypes = [C.c_void_p]
assert lib.dll.dplasma_b200_debug_read_tile_leaf(leaf) == 0
MHZ = 1965.0
l = [leaf[7]] + [leaf[k] for k in range(5)]
names = ["chol(A11)", "inv(L11)|L21", "A22 update", "chol(A22)", "inv(L22)|T"]
print(f"n={n} (debug build) whole call {e0.elapsed_time(e1) * 1e3:.1f} us; last leaf, us: " +
      ", ".join(f"{nm} {(l[k + 1] - l[k]) / MHZ:.1f}" for k, nm in enumerate(names)))
steps = (C.c_longlong * 256)()
lib.dll.dplasma_b200_debug_read_tile_steps.argtypes = [C.c_void_p]
assert lib.dll.dplasma_b200_debug_read_tile_steps(steps) == 0
nblk = (n + 63) // 64
for j in range(min(nblk - 1, 31)):
    c = [steps[j * 8 + k] for k in range(5)]
    nxt = steps[(j + 1) * 8] if j + 1 < nblk - 1 else None
    print(f"  step {j:2d} (us): solve {(c[1]-c[0])/MHZ:5.1f}  arrive {(c[2]-c[1])/MHZ:4.1f}  update {(c[3]-c[2])/MHZ:5.1f}  leaf {(c[4]-c[3])/MHZ:5.1f}" +
          (f"  barrier-1 wait {(nxt-c[4])/MHZ:5.1f}" if nxt else ""))
ctx.close()
