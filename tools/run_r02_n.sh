#!/bin/bash
# round-2 check N (4 GPUs): bench at N=4 with the c5 record (zpotrf N=49152 NB=512 P=2 Q=2)
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --steps 2 --warmup 3 > gpurun_out/n_bench4.log 2>&1; echo "rc=$?" >> gpurun_out/n_bench4.log
grep "^{" gpurun_out/n_bench4.log > gpurun_out/n_bench4.json; tail -2 gpurun_out/n_bench4.log | cut -c1-200
