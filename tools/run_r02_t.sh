#!/bin/bash
# round-2 final check T (2 GPUs): grid tests with the final tile kernel, bench at N=2 with the upper record
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "not large and not config5 and not poinv" > gpurun_out/t_multi.log 2>&1; echo "multi rc=$?" >> gpurun_out/t_multi.log; tail -3 gpurun_out/t_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/t_bench2.log 2>&1; echo "rc=$?" >> gpurun_out/t_bench2.log
grep "^{" gpurun_out/t_bench2.log > gpurun_out/t_bench2.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/t_bench2.json'))
print('value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],'resid',d['residual'],'upper',d.get('upper'))
PY
