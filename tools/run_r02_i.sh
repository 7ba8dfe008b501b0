#!/bin/bash
mkdir -p gpurun_out
for n in 64 128 512 1024; do timeout 60 python tools/prof_tile.py $n 2>&1 | tail -1; done | tee gpurun_out/i_tile.log
timeout 300 python tools/e2e_probe.py > gpurun_out/i_e2e.log 2>&1; tail -3 gpurun_out/i_e2e.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest_all.log; tail -3 gpurun_out/i_pytest_all.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/i_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/i_bench1.log; tail -2 gpurun_out/i_bench1.log | cut -c1-300
