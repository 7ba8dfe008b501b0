#!/bin/bash
# round-2 check B (1 GPU): new GEMM engine (tensor-map TMA + pipelined consumer loop): sanity, full GPU suite, isolated GEMM,
# C2 bench, library bar, ncu of the GEMM kernel
mkdir -p gpurun_out
timeout 120 python tools/verify_gemm.py > gpurun_out/b_verify.log 2>&1; echo "verify rc=$?" >> gpurun_out/b_verify.log
tail -8 gpurun_out/b_verify.log
if ! grep -q "verify rc=0" gpurun_out/b_verify.log; then echo "GEMM sanity failed; stopping"; exit 1; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -5 gpurun_out/b_pytest.log
timeout 120 python tools/prof_gemm.py > gpurun_out/b_prof_gemm.log 2>&1; tail -2 gpurun_out/b_prof_gemm.log
timeout 120 python tools/prof_gemm.py 16384 1024 >> gpurun_out/b_prof_gemm.log 2>&1; tail -1 gpurun_out/b_prof_gemm.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/b_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/b_bench1.log
tail -2 gpurun_out/b_bench1.log | cut -c1-1800
timeout 600 python tools/library_bar.py > gpurun_out/b_library_bar.md 2> gpurun_out/b_library_bar.err; echo "libbar rc=$?"; cat gpurun_out/b_library_bar.md; tail -3 gpurun_out/b_library_bar.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_dmma -s 2 -c 1 -o gpurun_out/prof_gemm_r02a python tools/prof_gemm.py > gpurun_out/b_ncu.log 2>&1; tail -2 gpurun_out/b_ncu.log
