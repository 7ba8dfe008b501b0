#!/bin/bash
# round-2 check U (2 GPUs): dplasmaUpper on grids through the local transposition onto the swapped grid
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -k "upper or (posv_grid_matches and U) or (test_pdpotrf_grid and U) or rsrc" > gpurun_out/u_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/u_multi.log; tail -5 gpurun_out/u_multi.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/upper_probe.py > gpurun_out/u_upper2.log 2>&1; echo "rc=$?" >> gpurun_out/u_upper2.log
grep "^{" gpurun_out/u_upper2.log > gpurun_out/u_upper2.json; cat gpurun_out/u_upper2.json; tail -3 gpurun_out/u_upper2.log
