#!/bin/bash
# round-2 check D (1 GPU): sign-flip prologue/epilogue variant of the GEMM engine
mkdir -p gpurun_out
timeout 120 python tools/verify_gemm.py > gpurun_out/d_verify.log 2>&1; echo "verify rc=$?" >> gpurun_out/d_verify.log; tail -3 gpurun_out/d_verify.log
timeout 120 python tools/prof_gemm.py > gpurun_out/d_prof_gemm.log 2>&1; tail -1 gpurun_out/d_prof_gemm.log
timeout 120 python tools/prof_gemm.py 16384 1024 >> gpurun_out/d_prof_gemm.log 2>&1; tail -1 gpurun_out/d_prof_gemm.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log; tail -3 gpurun_out/d_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/d_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/d_bench1.log; tail -2 gpurun_out/d_bench1.log | cut -c1-400
