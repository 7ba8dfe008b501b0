#!/bin/bash
# round-2 run ZZ (2 GPUs): solved row as an M-major operand (transposed) in the grid and single-GPU solves
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q -k "posv_grid or poinv" > gpurun_out/zz_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/zz_multi.log; tail -4 gpurun_out/zz_multi.log | cut -c1-300
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "trsm or potrs or posv" > gpurun_out/zz_single.log 2>&1; echo "single rc=$?" >> gpurun_out/zz_single.log; tail -3 gpurun_out/zz_single.log | cut -c1-300
DPB_TRACE_SOLVE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523 tools/posv_probe.py 32768 1024 1024 256 > gpurun_out/zz_posv2.log 2>&1; echo "rc=$?" >> gpurun_out/zz_posv2.log
grep "^\[solve\]" gpurun_out/zz_posv2.log | tail -4 > gpurun_out/zz_trace.txt; grep "^{" gpurun_out/zz_posv2.log | cut -c1-700 >> gpurun_out/zz_trace.txt; tail -1 gpurun_out/zz_posv2.log >> gpurun_out/zz_trace.txt
cat gpurun_out/zz_trace.txt | cut -c1-400
timeout 120 python tools/bench_posv.py 32768 1024 1024 256 2>&1 | tail -1 | tee gpurun_out/zz_posv1.log | cut -c1-300
