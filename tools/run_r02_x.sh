#!/bin/bash
# round-2 run X (2 GPUs): phases of the grid dpotrs (DPB_TRACE_SOLVE)
mkdir -p gpurun_out
export DPB_TRACE_SOLVE=1
for cfg in "32768 1024 1024 256" "32768 1024 1024 512"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 tools/posv_probe.py $cfg > gpurun_out/x_posv2.log 2>&1; echo "rc=$?" >> gpurun_out/x_posv2.log
  echo "== $cfg" >> gpurun_out/x_trace.txt; grep "^\[solve\]" gpurun_out/x_posv2.log >> gpurun_out/x_trace.txt; grep "^{" gpurun_out/x_posv2.log | cut -c1-700 >> gpurun_out/x_trace.txt; tail -1 gpurun_out/x_posv2.log >> gpurun_out/x_trace.txt
done
cat gpurun_out/x_trace.txt | cut -c1-400
