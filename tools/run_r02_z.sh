#!/bin/bash
# round-2 run Z (2 GPUs): grid solves with large inverse blocks built one step ahead -- tests, phases
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gpu_multi.py -x -q -k "posv_grid or poinv" > gpurun_out/z_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/z_multi.log; tail -5 gpurun_out/z_multi.log | cut -c1-300
export DPB_TRACE_SOLVE=1
for ibs in 256 1024; do
  DPB_SOLVE_IBS=$ibs timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/posv_probe.py 32768 1024 1024 256 > gpurun_out/z_posv2.log 2>&1; echo "rc=$?" >> gpurun_out/z_posv2.log
  echo "== ibs=$ibs" >> gpurun_out/z_trace.txt; grep "^\[solve\]" gpurun_out/z_posv2.log | tail -4 >> gpurun_out/z_trace.txt; grep "^{" gpurun_out/z_posv2.log | cut -c1-700 >> gpurun_out/z_trace.txt; tail -1 gpurun_out/z_posv2.log >> gpurun_out/z_trace.txt
done
cat gpurun_out/z_trace.txt | cut -c1-400
