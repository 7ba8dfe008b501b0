#!/bin/bash
mkdir -p gpurun_out
for n in 64 512; do timeout 60 python tools/tile_clocks.py $n 2>&1 | tail -1; done | tee gpurun_out/j_tile_clocks.log
for n in 64 128 512 1024; do timeout 60 python tools/prof_tile.py $n 2>&1 | tail -1; done | tee gpurun_out/j_tile.log
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "dpotrf_tile or tiny or info_semantics or fast_path" 2>&1 | tail -2
