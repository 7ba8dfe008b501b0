#!/bin/bash
# round-2 check F (1 GPU): (1) arrival-aware staging with the old tile kernel, (2) single-launch tile kernel
mkdir -p gpurun_out
export DPB_POTRF_TILE=col
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "host_matrix or pdpotrf or alpha_zero or 65535 or info_semantics" > gpurun_out/f_pytest_stage.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest_stage.log; tail -4 gpurun_out/f_pytest_stage.log
timeout 300 python tools/e2e_probe.py > gpurun_out/f_e2e.log 2>&1; cat gpurun_out/f_e2e.log
DPB_NO_PIPELINE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/f_e2e_nopipe.log 2>&1; tail -2 gpurun_out/f_e2e_nopipe.log
unset DPB_POTRF_TILE
for n in 64 200 512 1024; do timeout 60 python tools/prof_tile.py $n 2>&1 | tail -1; done | tee gpurun_out/f_tile.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "dpotrf_tile or dpotrf_378 or fast_path or tiny" > gpurun_out/f_pytest_tile.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest_tile.log; tail -4 gpurun_out/f_pytest_tile.log
if grep -q "pytest rc=0" gpurun_out/f_pytest_tile.log; then
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest_all.log; tail -3 gpurun_out/f_pytest_all.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/f_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/f_bench1.log; tail -2 gpurun_out/f_bench1.log | cut -c1-300
fi
