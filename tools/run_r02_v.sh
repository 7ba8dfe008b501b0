#!/bin/bash
# round-2 run V (1 GPU): grouped work order of the GEMM engine -- tests, A/B timing, DRAM traffic, library bar, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v_pytest_all.log; tail -3 gpurun_out/v_pytest_all.log
for o in 0 grouped; do
  if [ $o = 0 ]; then export DPB_GEMM_ORDER=0; else unset DPB_GEMM_ORDER; fi
  timeout 200 python tools/prof_gemm.py 16384 512 2>&1 | tail -1 | sed "s/^/order=$o /" | tee -a gpurun_out/v_order.log
  timeout 200 python tools/prof_gemm.py 16384 1024 2>&1 | tail -1 | sed "s/^/order=$o /" | tee -a gpurun_out/v_order.log
  timeout 300 python tools/prof_order.py 2>&1 | tail -1 | tee -a gpurun_out/v_order.log
done
unset DPB_GEMM_ORDER
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"
for o in 0 grouped; do
  if [ $o = 0 ]; then export DPB_GEMM_ORDER=0; else unset DPB_GEMM_ORDER; fi
  timeout 300 ncu --metrics $M --clock-control none -k regex:gemm_dmma -s 2 -c 1 --csv --log-file gpurun_out/v_ncu_single_$o.csv python tools/prof_gemm.py 16384 512 > gpurun_out/v_ncu_single_$o.log 2>&1
  timeout 400 ncu --metrics $M --clock-control none -k regex:gemm_dmma -c 10 --csv --log-file gpurun_out/v_ncu_trailing_$o.csv python tools/prof_order.py --once > gpurun_out/v_ncu_trailing_$o.log 2>&1
done
unset DPB_GEMM_ORDER
timeout 400 python tools/library_bar.py > gpurun_out/v_library_bar.md 2> gpurun_out/v_library_bar.err; tail -18 gpurun_out/v_library_bar.md | cut -c1-160
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/v_bench1.log 2>&1; echo "rc=$?" >> gpurun_out/v_bench1.log; grep "^{" gpurun_out/v_bench1.log > gpurun_out/v_bench1.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/v_bench1.json')); r=d['roofline']
print('value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'frac',r['frac'],'whole',r['whole_step_frac'],'peak',r['peak'],'cpu',d.get('cpu_baseline'))
PY
