#!/bin/bash
# round-2 run W (2 GPUs): grid solves with the all-to-all transposition and the look-ahead; dposv probe
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gpu_multi.py -x -q -k "posv_grid or poinv" > gpurun_out/w_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/w_multi.log; tail -5 gpurun_out/w_multi.log | cut -c1-300
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/posv_probe.py > gpurun_out/w_posv2.log 2>&1; echo "rc=$?" >> gpurun_out/w_posv2.log
grep "^{" gpurun_out/w_posv2.log > gpurun_out/w_posv2.json; cat gpurun_out/w_posv2.json; tail -3 gpurun_out/w_posv2.log | cut -c1-300
