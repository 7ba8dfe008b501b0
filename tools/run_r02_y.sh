#!/bin/bash
# round-2 run Y (1 GPU): large inverse blocks in the Left tile solves -- tests, dposv timing per block size; DRAM traffic
# of the bulk trailing-update launches under both work orders
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/y_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/y_pytest_all.log; tail -3 gpurun_out/y_pytest_all.log | cut -c1-300
for ibs in 64 128 256 512 1024; do
  DPB_SOLVE_IBS=$ibs timeout 200 python tools/bench_posv.py 32768 1024 1024 256 2>&1 | tail -1 | sed "s/^/ibs=$ibs /" | tee -a gpurun_out/y_posv1.log
done
DPB_SOLVE_IBS=256 timeout 200 python tools/bench_posv.py 32768 1024 1024 256 U 2>&1 | tail -1 | sed "s/^/ibs=256 /" | tee -a gpurun_out/y_posv1.log
DPB_SOLVE_IBS=256 timeout 200 python tools/bench_posv.py 32768 512 1024 256 2>&1 | tail -1 | sed "s/^/ibs=256 /" | tee -a gpurun_out/y_posv1.log
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct"
for o in 0 grouped; do
  if [ $o = 0 ]; then export DPB_GEMM_ORDER=0; else unset DPB_GEMM_ORDER; fi
  timeout 400 ncu --metrics $M --clock-control none -k regex:gemm_dmma -c 40 --csv --log-file gpurun_out/y_ncu_trailing_$o.csv python tools/prof_order.py --once > gpurun_out/y_ncu_trailing_$o.log 2>&1
done
unset DPB_GEMM_ORDER
python - <<'PY'
import csv, io
for o in ("0", "grouped"):
    lines = [l for l in open(f"gpurun_out/y_ncu_trailing_{o}.csv").read().splitlines() if l.startswith('"')]
    by = {}
    for r in csv.DictReader(io.StringIO("\n".join(lines))):
        by.setdefault(r["ID"], {"grid": r["Grid Size"]})[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    big = sorted(by.values(), key=lambda v: -v.get("gpu__time_duration.sum", 0))[:3]
    for v in big:
        print(o, v)
PY
