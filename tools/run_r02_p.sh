#!/bin/bash
# round-2 check P (1 GPU): auxiliary-kernel ncu metrics, library bar, full suite, bench
mkdir -p gpurun_out
bash tools/run_ncu_aux.sh
timeout 900 python tools/library_bar.py > gpurun_out/p_library_bar.md 2> gpurun_out/p_library_bar.err; echo "libbar rc=$?"; head -22 gpurun_out/p_library_bar.md | cut -c1-220
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/p_pytest_all.log; tail -3 gpurun_out/p_pytest_all.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/p_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/p_bench1.log; grep "^{" gpurun_out/p_bench1.log > gpurun_out/p_bench1.json; tail -2 gpurun_out/p_bench1.log | cut -c1-250
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/p_ref1.log 2>&1; tail -1 gpurun_out/p_ref1.log | cut -c1-300
