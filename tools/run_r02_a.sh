#!/bin/bash
# round-2 check A (2 GPUs): multi-GPU tests, bench at N=2 with small c4/c5 records, reference arm under torchrun, bench at N=1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/a_gpus.txt 2>&1
nproc >> gpurun_out/a_gpus.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/a_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/a_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/a_multi.log
DPB_BENCH_C4=16384,1024,1024,256 DPB_BENCH_C5=16384,512 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/a_bench2.log 2>&1; echo "rc=$?" >> gpurun_out/a_bench2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/a_ref2.log 2>&1; echo "rc=$?" >> gpurun_out/a_ref2.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/a_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/a_bench1.log
tail -3 gpurun_out/a_multi.log; tail -2 gpurun_out/a_bench2.log | cut -c1-1500; tail -2 gpurun_out/a_ref2.log | cut -c1-600; tail -2 gpurun_out/a_bench1.log | cut -c1-1500
