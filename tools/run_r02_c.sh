#!/bin/bash
# round-2 check C (1 GPU): full GPU suite, library bar, smoke
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -6 gpurun_out/c_pytest.log
timeout 900 python tools/library_bar.py > gpurun_out/c_library_bar.md 2> gpurun_out/c_library_bar.err; echo "libbar rc=$?"; cat gpurun_out/c_library_bar.md; tail -3 gpurun_out/c_library_bar.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
