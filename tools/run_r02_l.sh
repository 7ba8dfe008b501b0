#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -k "host_matrix" > gpurun_out/l_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/l_multi.log; tail -3 gpurun_out/l_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/l_bench2.log 2>&1; echo "rc=$?" >> gpurun_out/l_bench2.log
python - <<'PY'
import json
for ln in open('gpurun_out/l_bench2.log'):
    if ln.startswith('{'):
        d=json.loads(ln); print('value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'resid',d['residual'])
PY
