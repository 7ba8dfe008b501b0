#!/bin/bash
# round-2 check M (8 GPUs): C3 bench with residual + c4, per-phase trace, two grid tests
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/m_bench8.log 2>&1; echo "rc=$?" >> gpurun_out/m_bench8.log
grep "^{" gpurun_out/m_bench8.log > gpurun_out/m_bench8.json; tail -3 gpurun_out/m_bench8.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 tools/trace_grid.py > gpurun_out/m_trace8.log 2>&1; echo "rc=$?" >> gpurun_out/m_trace8.log; tail -30 gpurun_out/m_trace8.log | cut -c1-200
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -k "rsrc_csrc and 2-4 or (host_matrix and 2-4 and sym)" > gpurun_out/m_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/m_multi.log; tail -3 gpurun_out/m_multi.log
