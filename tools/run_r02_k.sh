#!/bin/bash
# round-2 check K (2 GPUs): grid paths after the arrival-aware staging, solve prefetch, RSRC/CSRC + sub(A)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/k_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/k_multi.log; tail -4 gpurun_out/k_multi.log
DPB_BENCH_C4=32768,1024,1024,256 DPB_BENCH_C5=16384,512 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/k_bench2.log 2>&1; echo "rc=$?" >> gpurun_out/k_bench2.log
tail -2 gpurun_out/k_bench2.log | cut -c1-600
