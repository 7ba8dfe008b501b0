#!/usr/bin/env python
"""bench.py -- dpotrf TFLOP/s (N^3/3 + N^2/2 + N/6 flops, src/flops.h:86-87) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one tiled Cholesky factorization (lower) of a dplgsy(bump=N, seed=3872) matrix:
  N=1 : BASELINE.json configs[1]  dpotrf N=32768 NB=512 on 1 x B200
  N>1 : 2D block-cyclic over a P x Q grid, one process per GPU, NB=1024, N grown so that each
        GPU holds the 17.2 GB of configs[2] (N=131072 at 8 GPUs, P=2 Q=4)  -> "weak" scaling.
`value` times the factorization only, matrix resident and generated in HBM (what the reference
times: tests/common.h:258-263); `e2e` is the same factorization through the blocking
dplasma_dpotrf() C-ABI call on a HOST (pinned) sym-block-cyclic matrix, H2D + D2H inside.
--impl reference times the reference's own CPU tile kernels (oracle/_ref) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 3872
METRIC = "dpotrf TFLOP/s (N^3/3) at 1/2/4/8 B200; % of FP64 tensor peak"


def flops_dpotrf(n):
    n = float(n)
    return n * n * n / 3.0 + n * n / 2.0 + n / 6.0


def grid_for(n):
    return {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}[n]


def workload_for(n_gpus):
    if n_gpus == 1:
        return 32768, 512
    import math
    nb = 1024
    N = int(round(131072 * math.sqrt(n_gpus / 8.0) / nb)) * nb
    return N, nb


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons, pw = [], [], set(), []
        for ts, ln in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 0.3):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def config_for(n_gpus):
    """the workload both arms report (BASELINE.json configs[1] at N=1, configs[2] at N=8)"""
    P, Q = grid_for(n_gpus)
    N, nb = workload_for(n_gpus)
    local = 8.0 * N * N / n_gpus
    return {"workload": f"dpotrf N={N} NB={nb} lower on {n_gpus} x B200, 2D block-cyclic P={P} Q={Q}, "
                        f"dplgsy(bump=N, seed={SEED})",
            "N": N, "NB": nb, "P": P, "Q": Q,
            "l2": f"inputs larger than L2 ({local / 1e9:.1f} GB per GPU vs 126 MB), fresh matrix per step",
            "timing": "CUDA events on the library's stream around each factorization, max over ranks; "
                      "generation outside the timed region like tests/common.h:258-263"}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


_PROBE = r"""
import ctypes, os, sys
sys.path.insert(0, %r)
from oracle.oracle import Reference, Lower
R = Reference()
i = ctypes.c_int(0)
t = R.lib.ref_time_dpotrf(Lower, 1024, 256, 3872, ctypes.byref(i))
print("PROBE", R.info()["core"], i.value)
"""


def pick_openblas_core():
    """The image's OpenBLAS 0.3.15 is a DYNAMIC_ARCH build that does not know this CPU model and falls back
    to its SSE-era 'Prescott' kernels.  Pick the best kernel set the host can
    actually EXECUTE: each candidate runs a small factorization in a child process (an illegal
    instruction kills only the child).  Must run before the library is loaded."""
    if os.environ.get("OPENBLAS_CORETYPE"):
        return os.environ["OPENBLAS_CORETYPE"], "set by the caller"
    flags = ""
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("flags"):
                flags = ln
                break
    except OSError:
        pass
    cands = []
    if " avx512f" in flags or not flags:
        cands.append("SKYLAKEX")
    if " avx2" in flags or not flags:
        cands.append("HASWELL")
    for c in ("SKYLAKEX", "HASWELL", "SANDYBRIDGE"):   # masked CPUID: try anyway, the child tells
        if c not in cands:
            cands.append(c)
    for core in cands:
        env = dict(os.environ, OPENBLAS_CORETYPE=core, OMP_NUM_THREADS="2")
        try:
            r = subprocess.run([sys.executable, "-c", _PROBE % ROOT], env=env, capture_output=True, text=True, timeout=120)
        except subprocess.TimeoutExpired:
            continue
        if r.returncode == 0 and "PROBE" in r.stdout and r.stdout.split("PROBE", 1)[1].split()[1] == "0":
            os.environ["OPENBLAS_CORETYPE"] = core
            return core, "probed in a child process"
    return None, "no AVX kernel set runs on this host; OpenBLAS default"


class CpuReference:
    """The reference's own CPU tile kernels under an OpenMP DAG (oracle/_ref), else the oracle port.
    torch.distributed.run exports OMP_NUM_THREADS=1: the thread count is set explicitly."""

    def __init__(self):
        self.core_choice, self.core_how = pick_openblas_core()
        self.threads = host_threads()
        try:
            from oracle.oracle import Reference, Lower
            self.R = Reference()
            self.R.lib.ref_set_num_threads(self.threads)
            self.kind = "reference"
            inf = self.R.info()
            self.blas = f"{inf['blas']} core={inf['core']} ({self.core_how})"
            self.core = inf["core"]
            L = self.R.lib
            if hasattr(L, "ref_time_lapacke_dpotrf"):
                L.ref_time_lapacke_dpotrf.restype = C.c_double
                L.ref_time_lapacke_dpotrf.argtypes = [C.c_int] * 3 + [C.c_ulonglong, C.c_int, C.POINTER(C.c_int)]
        except (FileNotFoundError, OSError):
            from oracle.oracle import Oracle
            self.R = None
            self.O = Oracle()
            self.kind, self.threads, self.blas, self.core = "port", 1, "plain C loops", "n/a"
        self.Lower = 122

    def dpotrf_seconds(self, N, nb):
        if self.R is not None:
            t, info = self.R.time_dpotrf(self.Lower, N, nb, SEED)
            assert info == 0
            return t
        A = self.O.plgsy(float(N), self.Lower, N, nb, SEED)
        t0 = time.perf_counter()
        assert self.O.potrf(self.Lower, A, nb) == 0
        return time.perf_counter() - t0

    def lapacke_seconds(self, N, nb):
        if self.R is None or not hasattr(self.R.lib, "ref_time_lapacke_dpotrf"):
            return None
        i = C.c_int(0)
        t = self.R.lib.ref_time_lapacke_dpotrf(self.Lower, N, nb, SEED, self.threads, C.byref(i))
        return t if i.value == 0 and t > 0 else None

    def describe(self):
        return (f"reference CORE_d* tile kernels ({self.blas}) under an OpenMP task DAG restating src/zpotrf_L.jdf, "
                f"serial BLAS inside tasks (INSTALL.rst:642-650); {self.threads} host threads on '{cpu_model()}'")


def cpu_sample_size(cpu, runs, budget_s, n_gpus=1):
    """Bounded sample of the workload: the largest candidate N for which `runs` factorizations fit `budget_s`,
    from a calibration run at N=8192 (C1, N=4096 NB=256, is timed too and reported).  The largest candidate is the configured workload itself when
    it is C2 (N=32768 NB=512, 8.6 GB on the host); the others use NB=256 like C1."""
    t1 = cpu.dpotrf_seconds(4096, 256)
    rate = flops_dpotrf(8192) / cpu.dpotrf_seconds(8192, 256)   # N=4096 is too small a DAG to fill many cores
    Nw, nbw = workload_for(n_gpus)
    cands = [(n, 256) for n in (6144, 8192, 12288, 16384, 24576) if n < Nw]
    if Nw <= 32768:
        cands.append((Nw, nbw))
    best = (4096, 256)
    for n, nb in cands:
        est = flops_dpotrf(n) / rate + 2.5 * (n / 32768.0) ** 2   # + generation and page faults of the n x n array
        if runs * est <= budget_s:
            best = (n, nb)
    return best[0], best[1], t1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cpu = CpuReference()
    K, W = max(1, args.steps), max(0, min(args.warmup, 1))
    N, nb, t_c1 = cpu_sample_size(cpu, K + W, 240.0 if cpu.kind == "reference" else 60.0, args.gpus)
    for _ in range(W):
        cpu.dpotrf_seconds(N, nb)
    ts = [cpu.dpotrf_seconds(N, nb) for _ in range(K)]
    t = sum(ts) / len(ts)
    val = flops_dpotrf(N) / t / 1e12
    sample = (f"dpotrf N={N} NB={nb} lower, dplgsy(bump=N, seed={SEED}); {cpu.describe()}; bounded sample of the "
              f"configured workload (PaRSEC runtime itself unavailable on this box); C1 N=4096 NB=256: "
              f"{flops_dpotrf(4096) / t_c1 / 1e12:.3f} TFLOP/s")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "TFLOP/s", "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(args.gpus),
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": cpu.threads, "kind": cpu.kind, "sample": sample,
                         "blas_core": cpu.core, "cpu_model": cpu_model(), "sample_N": N, "sample_NB": nb},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel from the committed
# `ncu --set full` capture (16384 x 16384 x 512 NT update = one bulk launch of C2 at k ~ 32); see profiles/
GEMM_TRAFFIC = {"bytes": 5.39e9,
                "note": "profiles/ncu_gemm_order_r02_grouped.csv: dram read 3.29 GB + write 2.10 GB for a launch whose "
                        "algorithmic bytes are 4.43 GB (C read + write, both operand panels once) -> 1.22 x, "
                        "10 % of the HBM roof (was 7.30 GB with the strip-by-strip work order and plain C accesses)"}


def flops_zpotrf(n):
    n = float(n)   # src/flops.h:86-87,270-273: 6 FMULS + 2 FADDS
    return 6.0 * (n * n * n / 6.0 + n * n / 2.0 + n / 3.0) + 2.0 * (n * n * n / 6.0 - n / 6.0)


def _timed(torch, dist, barrier, fn):
    barrier()
    t0 = time.perf_counter()
    rc = fn()
    barrier()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return rc, t.item()


def _allmax(torch, dist, *vals):
    t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def bench_posv(lib, dp, ctx, torch, dist, barrier, rank, world, P, Q, N, nb, nrhs, nbr):
    """BASELINE.json configs[3]: dposv = dpotrf + dpotrs on device-resident block-cyclic matrices, then the two
    checks of tests/testing_zposv.c:60-135 (check_dpotrf on the factor, check_daxmb on the solution)."""
    mk = lambda m, n, a, b: dp.BlockCyclic(lib, m, n, a, b, P, Q, rank, where="device")
    A, A0, B, X = mk(N, N, nb, nb), mk(N, N, nb, nb), mk(N, nrhs, nb, nbr), mk(N, nrhs, nb, nbr)
    out = {"workload": f"dposv N={N} NRHS={nrhs} NB={nb} (RHS tiles {nb}x{nbr}) lower on {world} x B200, P={P} Q={Q}; "
                       "A=dplgsy(bump=N, seed=3872), B=dplrnt(seed=2354); device-resident; wall clock between "
                       "barriers, max over ranks"}
    best = None
    for it in range(2):   # first pass warms the solve path (communicators, workspaces)
        for d in (A, A0):
            assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, d.ref, SEED) == 0
        for d in (B, X):
            assert lib.dplasma_dplrnt(ctx.ptr, 0, d.ref, 2354) == 0
        info, t_f = _timed(torch, dist, barrier, lambda: lib.dplasma_dpotrf(ctx.ptr, dp.Lower, A.ref))
        rc, t_s = _timed(torch, dist, barrier, lambda: lib.dplasma_dpotrs(ctx.ptr, dp.Lower, A.ref, X.ref))
        assert info == 0 and rc == 0, (info, rc)
        if best is None or t_f + t_s < best[0] + best[1]:
            best = (t_f, t_s)
    t_f, t_s = best
    bad_f = lib.check_dpotrf(ctx.ptr, 0, dp.Lower, A.ref, A0.ref)
    res_f = lib.dplasma_b200_last_check_result(ctx.ptr)
    bad_s = lib.check_daxmb(ctx.ptr, 0, dp.Lower, A0.ref, B.ref, X.ref)
    res_s = lib.dplasma_b200_last_check_result(ctx.ptr)
    bad_f, res_f, bad_s, res_s = _allmax(torch, dist, bad_f, res_f, bad_s, res_s)
    ff, fs = flops_dpotrf(N), lib.dplasma_b200_flops_dpotrs(float(N), float(nrhs))
    out.update({"seconds": t_f + t_s, "tflops": (ff + fs) / (t_f + t_s) / 1e12,
                "potrf_seconds": t_f, "potrf_tflops": ff / t_f / 1e12,
                "potrs_seconds": t_s, "potrs_tflops": fs / t_s / 1e12,
                "residual_facto": res_f, "residual_solve": res_s,
                "passed": bool(bad_f == 0 and bad_s == 0 and res_f < 60 and res_s < 60)})
    for d in (A, A0, B, X):
        d.free()
    return out


def bench_zpotrf(lib, dp, ctx, torch, dist, barrier, rank, world, P, Q, N, nb):
    """BASELINE.json configs[4]: zpotrf, complex double, lower, through check_zpotrf (src/dplasma_zcheck.c:64-134)"""
    mk = lambda: dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device", cplx=True)
    A, A0 = mk(), mk()
    best = None
    for it in range(2):
        assert lib.dplasma_zplghe(ctx.ptr, float(N), dp.Lower, A.ref, SEED) == 0
        barrier()
        info = lib.dplasma_zpotrf(ctx.ptr, dp.Lower, A.ref)
        assert info == 0, info
        t = _allmax(torch, dist, lib.dplasma_b200_last_elapsed(ctx.ptr))[0]
        best = t if best is None else min(best, t)
    assert lib.dplasma_zplghe(ctx.ptr, float(N), dp.Lower, A0.ref, SEED) == 0
    bad = lib.check_zpotrf(ctx.ptr, 0, dp.Lower, A.ref, A0.ref)
    res = lib.dplasma_b200_last_check_result(ctx.ptr)
    bad, res = _allmax(torch, dist, bad, res)
    out = {"workload": f"zpotrf N={N} NB={nb} lower, complex double, on {world} x B200, P={P} Q={Q}; "
                       "zplghe(bump=N, seed=3872); device-resident; CUDA events on the library's stream, max over ranks",
           "seconds": best, "tflops": flops_zpotrf(N) / best / 1e12,
           "flops_model": "FLOPS_ZPOTRF = 6 FMULS + 2 FADDS (src/flops.h:86-87,270-273), real flops",
           "residual": res, "passed": bool(bad == 0 and res < 60)}
    A.free()
    A0.free()
    return out


def run_ours(args):
    import torch
    import dplasma_b200 as dp

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = args.gpus
    if world != n_gpus:
        if world == 1 and n_gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: dplasma_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    lib = dp.load()
    ctx = dp.Context(lib)
    if world > 1:
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
    P, Q = grid_for(n_gpus)
    N, nb = workload_for(n_gpus)
    flops = flops_dpotrf(N)
    K, W = args.steps, max(args.warmup, 3)

    # ---- device-resident matrices: as many pre-generated copies as fit comfortably ----
    d0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device")
    local_bytes = d0.elems * 8
    nbuf = max(1, min(K, int(90e9 // max(local_bytes, 1))))
    bufs = [d0] + [dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device") for _ in range(nbuf - 1)]

    def gen(d):
        rc = lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, d.ref, SEED)
        assert rc == 0, rc

    def factor(d):
        info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref)
        assert info == 0, f"dpotrf info={info}"
        return lib.dplasma_b200_last_elapsed(ctx.ptr)

    for _ in range(W):
        gen(d0)
        factor(d0)

    # ---- timed region: exactly K factorizations, barrier + synchronize on both sides ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    dev_seconds, wall_seconds, done = 0.0, 0.0, 0
    launches0 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    launches = 0
    t_first, t_last = None, None
    while done < K:
        chunk = min(nbuf, K - done)
        for b in bufs[:chunk]:
            gen(b)                       # untimed: inputs resident in HBM when the timed region starts
        barrier()
        l0 = lib.dplasma_b200_kernel_launches(ctx.ptr)
        t0 = time.perf_counter()
        t_first = t_first or time.time()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for b in bufs[:chunk]:
            dev_seconds += factor(b)     # CUDA events on the library's own stream, per factorization
        barrier()
        wall_seconds += time.perf_counter() - t0
        t_last = time.time()
        launches += lib.dplasma_b200_kernel_launches(ctx.ptr) - l0
        done += chunk
    clocks = sampler.stop(t_first, t_last) if rank == 0 else None
    tt = torch.tensor([dev_seconds, wall_seconds], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dev_seconds, wall_seconds = tt.tolist()
    value = flops * K / dev_seconds / 1e12

    # ---- residual of the last factorization: the reference's acceptance test (src/dplasma_zcheck.c:89-119),
    # on every grid: check_dpotrf is collective, its norms are all-reduced inside the library ----
    for b in bufs[1:]:
        b.free()
    bufs = bufs[:1]
    dA0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device")
    gen(dA0)
    bad = lib.check_dpotrf(ctx.ptr, 0, dp.Lower, d0.ref, dA0.ref)
    residual = lib.dplasma_b200_last_check_result(ctx.ptr)
    dA0.free()
    chk = torch.tensor([float(bad), residual], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(chk, op=dist.ReduceOp.MAX)
    bad, residual = int(chk[0].item()), chk[1].item()
    assert bad == 0 and 0.0 <= residual < 60.0, f"residual check failed: {residual}"

    # ---- roofline leg: event-time every grouped trailing-update launch of one factorization.
    # The bulk updates run one after the other on the update stream (their summed time is a share of the
    # step <= 1); the look-ahead launches on the panel stream overlap them and are reported apart ----
    gen(d0)
    barrier()
    lib.dplasma_b200_profile(ctx.ptr, 1)
    t_prof = factor(d0)
    rd = {}
    for which, name in ((1, "bulk"), (0, "lookahead"), (-1, "all")):
        ps, pf, pl, pb = C.c_double(0), C.c_double(0), C.c_int(0), C.c_double(0)
        lib.dplasma_b200_profile_read_stream(ctx.ptr, which, C.byref(ps), C.byref(pf), C.byref(pl), C.byref(pb))
        rd[name] = (ps.value, pf.value, pl.value, pb.value)
    lib.dplasma_b200_profile(ctx.ptr, 0)
    secs = C.c_double(0)
    peak = lib.dplasma_b200_dmma_peak_tflops(2, 4096, C.byref(secs))
    bs, bf, bl, bb = rd["bulk"]
    ls, lf, ll, _ = rd["lookahead"]
    achieved = bf / bs / 1e12 if bs > 0 else None
    my_flops = flops / world   # this rank's share (block-cyclic: balanced to within one tile row/column)
    roofline = {
        "bound": "tensor", "kernel": "gemm_dmma_kernel (grouped trailing update: DGEMM + DSYRK tiles), bulk launches "
                                     "on the update stream of rank 0",
        "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if achieved and peak > 0 else None,
        "traffic": GEMM_TRAFFIC["bytes"], "traffic_note": GEMM_TRAFFIC["note"],
        "launches_timed": bl, "algorithmic_flops_per_launch_avg": bf / max(bl, 1),
        "best_launch_tflops": bb, "share_of_step": bs / t_prof if t_prof > 0 else None,
        "flops_share_of_step": bf / my_flops if my_flops > 0 else None,
        "lookahead": {"launches": ll, "seconds": ls, "flops": lf,
                      "note": "look-ahead updates of tile column k+1 on the panel stream; they overlap the bulk update "
                              "of step k-1, so their event brackets are not added to the bulk sum"},
        "whole_step_frac": (value / world / peak) if peak > 0 else None,
        "peak_source": "of measured: live register-resident DMMA (mma.sync.m8n8k4.f64) loop on this GPU "
                       "(MEASURED_PEAKS.json holds no FP64 figure; arithmetic: 148 SM x 128 flop/clk x 1.965 GHz = 37.2; "
                       "cuBLAS DGEMM ceiling on the same box: profiles/library_bar_r02.md)",
    }

    # ---- e2e: blocking C-ABI call on a HOST (pinned) sym block-cyclic matrix ----
    e2e = None
    try:
        h = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, kind="sym", uplo=dp.Lower, where="host")
        h2d = d2h = h.elems * 8
        ke = min(K, 3)
        tsum = 0.0
        for it in range(ke + 1):
            rc = lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, h.ref, SEED)  # untimed input generation
            assert rc == 0
            barrier()
            t0 = time.perf_counter()
            info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, h.ref)               # H2D + factor + D2H + info read
            barrier()
            dt = time.perf_counter() - t0
            assert info == 0
            if it > 0:
                tsum += dt
        te = torch.tensor([tsum], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        tsum = te.item()
        e2e = {"value": flops * ke / tsum / 1e12, "unit": "TFLOP/s", "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": d2h * world, "steps": ke, "ms_per_step": tsum / ke * 1e3,
               "api": "dplasma_dpotrf(parsec, dplasmaLower, A) with A->mat in pinned host memory (sym block-cyclic tile storage)"}
        h.free()
    except (MemoryError, AssertionError) as ex:  # host RAM too small for the pinned copy
        e2e = {"value": None, "unit": "TFLOP/s", "error": repr(ex)}

    # ---- dplasmaUpper (what the reference's own testing_zpotrf.c:21 runs), outside the timed region ----
    upper = None
    try:
        ts = []
        for it in range(2):
            assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Upper, d0.ref, SEED) == 0
            barrier()
            info = lib.dplasma_dpotrf(ctx.ptr, dp.Upper, d0.ref)
            assert info == 0, info
            ts.append(_allmax(torch, dist, lib.dplasma_b200_last_elapsed(ctx.ptr))[0])
        dU0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device")
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Upper, dU0.ref, SEED) == 0
        ubad = lib.check_dpotrf(ctx.ptr, 0, dp.Upper, d0.ref, dU0.ref)
        ures = lib.dplasma_b200_last_check_result(ctx.ptr)
        dU0.free()
        ubad, ures = _allmax(torch, dist, ubad, ures)
        upper = {"workload": "same matrix, uplo = dplasmaUpper (lower schedule on the transposed tile grid)",
                 "seconds": min(ts), "tflops": flops / min(ts) / 1e12, "residual": ures, "passed": bool(ubad == 0 and ures < 60)}
    except AssertionError as ex:
        upper = {"error": repr(ex)}

    # ---- BASELINE.json configs[3] / configs[4], outside the timed region, on the grids they name ----
    d0.free()
    bufs = []
    extra = {}
    # (an error code that every rank sees alike is reported in the record instead of taking the headline line down)
    if n_gpus == 8 or os.environ.get("DPB_BENCH_C4"):
        try:
            extra["c4"] = bench_posv(lib, dp, ctx, torch, dist, barrier, rank, world, P, Q,
                                     *[int(x) for x in os.environ.get("DPB_BENCH_C4", "65536,1024,1024,256").split(",")])
        except AssertionError as ex:
            extra["c4"] = {"error": repr(ex), "passed": False}
    if n_gpus == 4 or os.environ.get("DPB_BENCH_C5"):
        try:
            extra["c5"] = bench_zpotrf(lib, dp, ctx, torch, dist, barrier, rank, world, P, Q,
                                       *[int(x) for x in os.environ.get("DPB_BENCH_C5", "49152,512").split(",")])
        except AssertionError as ex:
            extra["c5"] = {"error": repr(ex), "passed": False}

    # ---- CPU baseline beside it (rank 0, N=1 only): the reference's own tile kernels ----
    cpu = None
    if rank == 0 and n_gpus == 1 and not args.no_cpu:
        ref = CpuReference()
        cn, cnb, t_c1 = cpu_sample_size(ref, 2, 36.0)
        t_s = ref.dpotrf_seconds(cn, cnb)
        t_l = ref.lapacke_seconds(cn, cnb)
        cpu = {"value": flops_dpotrf(cn) / t_s / 1e12, "unit": "TFLOP/s", "cores": ref.threads, "kind": ref.kind,
               "sample": f"dpotrf N={cn} NB={cnb} lower, one run of {t_s:.1f} s; {ref.describe()}",
               "blas_core": ref.core, "cpu_model": cpu_model(), "sample_N": cn, "sample_NB": cnb,
               "c1": {"workload": "BASELINE.json configs[0]: dpotrf N=4096 NB=256 lower on the host cores",
                      "value": flops_dpotrf(4096) / t_c1 / 1e12, "unit": "TFLOP/s", "seconds": t_c1},
               "lapacke_dpotrf": None if t_l is None else {
                   "workload": f"monolithic multithreaded LAPACKE_dpotrf N={cn} ({ref.threads} OpenBLAS threads)",
                   "value": flops_dpotrf(cn) / t_l / 1e12, "unit": "TFLOP/s", "seconds": t_l}}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": dev_seconds / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_for(n_gpus),
            "wall_ms_per_step": wall_seconds / K * 1e3,
            "residual": residual, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
        }
        out["upper"] = upper
        out.update(extra)
        print(json.dumps(out))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
