"""Runs the kernels of the path other than the GEMM engine once each at representative sizes, for an ncu capture
(tools/run_ncu_aux.sh): generators, norms, tile factorization, fused panel solve, complex engine, tile-row transposition."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dplasma_b200 as dp
lib = dp.load(); ctx = dp.Context(lib)
N, nb = 16384, 512
A = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
B = dp.BlockCyclic(lib, N, 1024, nb, 256, where="device")
assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, A.ref, 3872) == 0          # generate_kernel (plgsy)
assert lib.dplasma_dplrnt(ctx.ptr, 0, B.ref, 2354) == 0                           # generate_kernel (plrnt)
print("lansy", lib.dplasma_dlansy(ctx.ptr, dp.InfNorm, dp.Lower, A.ref))          # rowsum_kernel (sym), absmax_kernel
print("lange", lib.dplasma_dlange(ctx.ptr, dp.InfNorm, B.ref))                    # rowsum_kernel (general)
assert lib.dplasma_dpotrf(ctx.ptr, dp.Lower, A.ref) == 0                          # potrf_tile_kernel, trsm_panel_fused_kernel
assert lib.dplasma_dpotri(ctx.ptr, dp.Lower, A.ref) == 0                          # transpose_tiles_kernel, trtri_blocks_kernel
A.free(); B.free()
Nz = 8192
Z = dp.BlockCyclic(lib, Nz, Nz, nb, nb, where="device", cplx=True)
assert lib.dplasma_zplghe(ctx.ptr, float(Nz), dp.Lower, Z.ref, 3872) == 0
assert lib.dplasma_zpotrf(ctx.ptr, dp.Lower, Z.ref) == 0                          # zgemm_dmma_kernel, zpotrf_leaf_kernel
Z.free()
ctx.close()
print("ok")
