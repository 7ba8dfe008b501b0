#!/bin/bash
# round-2 check E (1 GPU): arrival-aware pipelined staging
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "host_matrix or pdpotrf or alpha_zero or 65535 or info_semantics" > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log; tail -4 gpurun_out/e_pytest.log
timeout 300 python tools/e2e_probe.py > gpurun_out/e_e2e.log 2>&1; cat gpurun_out/e_e2e.log
DPB_NO_PIPELINE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/e_e2e_nopipe.log 2>&1; tail -2 gpurun_out/e_e2e_nopipe.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest_all.log; tail -3 gpurun_out/e_pytest_all.log
