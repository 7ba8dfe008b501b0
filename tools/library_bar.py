"""The bar BASELINE.md section 4 sets for the tile kernels: the library calls the reference's GPU task bodies make
(src/zpotrf_L.jdf:143 cusolverDnDpotrf, :236-240 cublasDtrsm, :342-346 cublasDsyrk, :463-468 cublasDgemm), timed on
the same box beside this repo's kernels through the B2 seam (dplasma_b200_d{potrf,trsm,syrk,gemm}_tile), CUDA events,
device-resident operands.  cuBLAS / cuSOLVER are loaded here with ctypes -- measurement only, never linked into or
called from libdplasma_b200.  Also a sustained (>= 4 s) cuBLAS DGEMM as a second FP64 ceiling next to the library's
own register-resident DMMA probe.  Writes a markdown table to stdout (committed as profiles/library_bar_r02.md)."""
import ctypes as C
import glob
import os
import subprocess
import sys
import sysconfig

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import dplasma_b200 as dp  # noqa: E402

sp = sysconfig.get_paths()["purelib"]


def find(name):
    # the copies torch itself has loaded (site-packages/nvidia/*/lib); /usr/local/cuda only as a fallback: mixing a
    # libcublas from one tree with the libcublasLt of the other fails to resolve symbols
    c = glob.glob(os.path.join(sp, "nvidia", "*", "lib", name + "*")) or glob.glob("/usr/local/cuda/lib64/" + name + "*")
    if not c:
        raise SystemExit(f"{name} not found")
    return sorted(c, key=len)[0]


cublas = C.CDLL(find("libcublas.so.12"), mode=C.RTLD_GLOBAL)
cusolver = C.CDLL(find("libcusolver.so.11"), mode=C.RTLD_GLOBAL)
vp, i, d = C.c_void_p, C.c_int, C.c_double
OP_N, OP_T, FILL_LOWER, SIDE_RIGHT, DIAG_NON_UNIT = 0, 1, 0, 1, 0
hb, hs = vp(), vp()
assert cublas.cublasCreate_v2(C.byref(hb)) == 0
assert cusolver.cusolverDnCreate(C.byref(hs)) == 0
stream = torch.cuda.current_stream().cuda_stream
cublas.cublasSetStream_v2(hb, vp(stream))
cusolver.cusolverDnSetStream(hs, vp(stream))
pd = C.POINTER(d)
cublas.cublasDgemm_v2.argtypes = [vp, i, i, i, i, i, pd, vp, i, vp, i, pd, vp, i]
cublas.cublasDsyrk_v2.argtypes = [vp, i, i, i, i, pd, vp, i, pd, vp, i]
cublas.cublasDtrsm_v2.argtypes = [vp, i, i, i, i, i, i, pd, vp, i, vp, i]
cusolver.cusolverDnDpotrf_bufferSize.argtypes = [vp, i, i, vp, i, C.POINTER(i)]
cusolver.cusolverDnDpotrf.argtypes = [vp, i, i, vp, i, vp, i, vp]

lib = dp.load()
ctx = dp.Context(lib)
P = lambda t: vp(t.data_ptr())  # noqa: E731


def timed(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True, timeout=10).stdout.strip()
        return out
    except Exception:
        return "n/a"


rows = []
one, mone = d(1.0), d(-1.0)


def add(name, flops, t_lib, t_ours, note=""):
    rows.append((name, flops / t_lib / 1e12, flops / t_ours / 1e12, t_lib * 1e3, t_ours * 1e3, note))


# ---- K4: trailing-update GEMM, C[M x M] -= A[M x nb] B[M x nb]^T (one bulk launch of a step) ----
M = 16384
for nb in (512, 1024):
    A = torch.randn(nb, M, dtype=torch.float64, device="cuda")
    B = torch.randn(nb, M, dtype=torch.float64, device="cuda")
    Cm = torch.randn(M, M, dtype=torch.float64, device="cuda")
    t_lib = timed(lambda: cublas.cublasDgemm_v2(hb, OP_N, OP_T, M, M, nb, C.byref(mone), P(A), M, P(B), M, C.byref(one), P(Cm), M), 5)
    t_our = timed(lambda: lib.dplasma_b200_dgemm_tile(ctx.ptr, vp(stream), dp.NoTrans, dp.Trans, M, M, nb, -1.0, P(A), M, P(B), M, 1.0, P(Cm), M), 5)
    add(f"Dgemm NT {M}x{M}x{nb}", 2.0 * M * M * nb, t_lib, t_our, "one launch each")
    del A, B, Cm

# ---- K4 per tile, the way the reference issues it: one cublasDgemm per nb x nb tile (src/zpotrf_L.jdf:463-468);
#      ours: the same tiles as ONE batched launch (what the grouped trailing update does) ----
for nb in (512, 1024):
    cnt = 32 * 32 if nb == 512 else 16 * 16
    A = torch.randn(cnt, nb, nb, dtype=torch.float64, device="cuda")
    B = torch.randn(cnt, nb, nb, dtype=torch.float64, device="cuda")
    Cm = torch.randn(cnt, nb, nb, dtype=torch.float64, device="cuda")
    te = nb * nb * 8

    def per_tile():
        for k in range(cnt):
            cublas.cublasDgemm_v2(hb, OP_N, OP_T, nb, nb, nb, C.byref(mone), vp(A.data_ptr() + k * te), nb, vp(B.data_ptr() + k * te), nb,
                                  C.byref(one), vp(Cm.data_ptr() + k * te), nb)
    t_lib = timed(per_tile, 3, 1)
    # ours: M = cnt*nb rows against one B tile is not the same product; use the big-GEMM row above as "ours, grouped"
    t_our = timed(lambda: lib.dplasma_b200_dgemm_tile(ctx.ptr, vp(stream), dp.NoTrans, dp.Trans, nb * 32 if nb == 512 else nb * 16,
                                                    nb * 32 if nb == 512 else nb * 16, nb, -1.0, P(A), nb * 32 if nb == 512 else nb * 16,
                                                    P(B), nb * 32 if nb == 512 else nb * 16, 1.0, P(Cm), nb * 32 if nb == 512 else nb * 16), 3, 1)
    add(f"Dgemm per tile nb={nb}, {cnt} tiles", 2.0 * nb ** 3 * cnt, t_lib, t_our,
        "library: one call per tile like the reference; ours: the same flops as one grouped launch")
    del A, B, Cm

# ---- K3: Dsyrk on a diagonal tile, n = k = nb ----
for nb in (512, 1024):
    A = torch.randn(nb, nb, dtype=torch.float64, device="cuda")
    Cm = torch.randn(nb, nb, dtype=torch.float64, device="cuda")
    t_lib = timed(lambda: cublas.cublasDsyrk_v2(hb, FILL_LOWER, OP_N, nb, nb, C.byref(mone), P(A), nb, C.byref(one), P(Cm), nb), 50)
    t_our = timed(lambda: lib.dplasma_b200_dsyrk_tile(ctx.ptr, vp(stream), dp.Lower, dp.NoTrans, nb, nb, -1.0, P(A), nb, 1.0, P(Cm), nb), 50)
    add(f"Dsyrk L/N n=k={nb} (one tile)", nb * nb * (nb + 1.0), t_lib, t_our, "latency-bound at this size")

# ---- K2: Dtrsm Right/Lower/Trans/NonUnit, one tile and a 31-tile panel ----
for nb in (512, 1024):
    for tiles in (1, 31):
        L = torch.tril(torch.randn(nb, nb, dtype=torch.float64, device="cuda")).T.contiguous() + 4 * nb * torch.eye(nb, dtype=torch.float64, device="cuda")
        Bm = torch.randn(nb, tiles * nb, dtype=torch.float64, device="cuda")  # column-major (tiles*nb) x nb
        m = tiles * nb
        t_lib = timed(lambda: cublas.cublasDtrsm_v2(hb, SIDE_RIGHT, FILL_LOWER, OP_T, DIAG_NON_UNIT, m, nb, C.byref(one), P(L), nb, P(Bm), m), 20)
        t_our = timed(lambda: lib.dplasma_b200_dtrsm_tile(ctx.ptr, vp(stream), dp.Right, dp.Lower, dp.Trans, dp.NonUnit, m, nb, 1.0, P(L), nb, P(Bm), m), 20)
        add(f"Dtrsm R/L/T {m}x{nb}", float(m) * nb * nb, t_lib, t_our, "ours includes inverting the diagonal blocks")

# ---- K1: cusolverDnDpotrf on a tile, and on the whole C2 matrix ----
info = torch.zeros(4, dtype=torch.int32, device="cuda")
for n in (512, 1024):
    A0 = torch.randn(n, n, dtype=torch.float64, device="cuda")
    A0 = A0 @ A0.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
    lw = i(0)
    cusolver.cusolverDnDpotrf_bufferSize(hs, FILL_LOWER, n, P(A0), n, C.byref(lw))
    work = torch.empty(max(lw.value, 1), dtype=torch.float64, device="cuda")
    Aw = A0.clone()

    def lib_potrf():
        Aw.copy_(A0)
        cusolver.cusolverDnDpotrf(hs, FILL_LOWER, n, P(Aw), n, P(work), lw.value, P(info))

    def our_potrf():
        Aw.copy_(A0)
        lib.dplasma_b200_dpotrf_tile(ctx.ptr, vp(stream), dp.Lower, n, P(Aw), n, P(info), 0)
    t_copy = timed(lambda: Aw.copy_(A0), 50)
    t_lib = timed(lib_potrf, 30) - t_copy
    t_our = timed(our_potrf, 30) - t_copy
    add(f"Dpotrf tile n={n}", n ** 3 / 3.0, t_lib, t_our, "restore copy subtracted")

N, nb = 32768, 512
dA = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, dA.ref, 3872) == 0
full = torch.zeros(N, N, dtype=torch.float64, device="cuda")
# one column-major N x N array for cuSOLVER, filled by the library's own generator through a LAPACK-storage descriptor
dL = dp.BlockCyclic(lib, N, N, nb, nb, kind="lapack", where=full.data_ptr())
assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, dL.ref, 3872) == 0
torch.cuda.synchronize()
lw = i(0)
cusolver.cusolverDnDpotrf_bufferSize(hs, FILL_LOWER, N, P(full), N, C.byref(lw))
work = torch.empty(max(lw.value, 1), dtype=torch.float64, device="cuda")
ts = []
for rep in range(3):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, dL.ref, 3872) == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cusolver.cusolverDnDpotrf(hs, FILL_LOWER, N, P(full), N, P(work), lw.value, P(info))
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e-3)
t_lib = min(ts[1:])
ts = []
for rep in range(3):
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, dA.ref, 3872) == 0
    assert lib.dplasma_dpotrf(ctx.ptr, dp.Lower, dA.ref) == 0
    ts.append(lib.dplasma_b200_last_elapsed(ctx.ptr))
t_our = min(ts[1:])
add(f"Dpotrf N={N} (whole matrix, device-resident)", N ** 3 / 3.0 + N * N / 2.0 + N / 6.0, t_lib, t_our,
    "cusolverDnDpotrf on one column-major array vs dplasma_dpotrf on NB=512 tiles")
del full, work
dA.free()

# ---- second FP64 ceiling: sustained cuBLAS DGEMM 8192^3, >= 4 s ----
n = 8192
A = torch.randn(n, n, dtype=torch.float64, device="cuda")
B = torch.randn(n, n, dtype=torch.float64, device="cuda")
Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
zero = d(0.0)
gemm = lambda: cublas.cublasDgemm_v2(hb, OP_N, OP_T, n, n, n, C.byref(one), P(A), n, P(B), n, C.byref(zero), P(Cm), n)  # noqa: E731
t1 = timed(gemm, 3)
reps = max(8, int(4.5 / t1))
c0 = clocks()
t_sus = timed(gemm, reps, 0)
c1 = clocks()
our = lambda: lib.dplasma_b200_dgemm_tile(ctx.ptr, vp(stream), dp.NoTrans, dp.Trans, n, n, n, 1.0, P(A), n, P(B), n, 0.0, P(Cm), n)  # noqa: E731
t_our_sus = timed(our, reps, 2)
secs = d(0)
probe = lib.dplasma_b200_dmma_peak_tflops(2, 4096, C.byref(secs))

print("# Library bar, round 2 -- cuBLAS / cuSOLVER 12.9 vs this repo's kernels on the same B200\n")
print(f"`python tools/library_bar.py` on {torch.cuda.get_device_name(0)}; CUDA events on one stream; device-resident operands; "
      f"clocks before/after the sustained run (sm MHz, max MHz, W): {c0} / {c1}\n")
print("| operation | library TFLOP/s | this repo TFLOP/s | library ms | this repo ms | ratio ours/library | note |")
print("|---|---|---|---|---|---|---|")
for name, fl, fo, tl, to, note in rows:
    print(f"| {name} | {fl:.2f} | {fo:.2f} | {tl:.3f} | {to:.3f} | {fo / fl:.2f} | {note} |")
fl = 2.0 * n ** 3
print(f"\nFP64 ceilings on this GPU: register-resident DMMA probe (dplasma_b200_dmma_peak_tflops) **{probe:.2f} TFLOP/s**; "
      f"sustained cuBLAS DGEMM {n}^3 x {reps} launches ({t_sus * reps:.1f} s) **{fl / t_sus / 1e12:.2f} TFLOP/s**; "
      f"this repo's engine on the same product, same duration: **{fl / t_our_sus / 1e12:.2f} TFLOP/s**; "
      f"arithmetic 148 SM x 128 flop/clk x 1.965 GHz = 37.2.")
ctx.close()
