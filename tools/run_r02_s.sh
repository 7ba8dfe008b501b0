#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s_pytest_all.log; tail -3 gpurun_out/s_pytest_all.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/s_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/s_bench1.log; grep "^{" gpurun_out/s_bench1.log > gpurun_out/s_bench1.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/s_bench1.json')); r=d['roofline']
print('value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'frac',r['frac'],'whole',r['whole_step_frac'],'peak',r['peak'])
PY
timeout 300 python tools/library_bar.py 2>/dev/null | grep -E "Dpotrf|Dtrsm" 
