#!/bin/bash
# ncu metrics of the kernels other than the real GEMM engine (one pass, selected metrics, first launches of each kernel)
mkdir -p gpurun_out
timeout 300 python tools/aux_kernels.py > gpurun_out/aux_plain.log 2>&1 || { tail -5 gpurun_out/aux_plain.log; exit 1; }
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,sm__throughput.avg.pct_of_peak_sustained_elapsed
timeout 1200 ncu --metrics $M --clock-control none -k regex:"generate_kernel|rowsum_kernel|absmax|potrf_tile_kernel|trsm_panel_fused|zgemm_dmma|transpose_tiles|trtri_blocks|zpotrf_leaf|identity|copy_tri" -c 400 --csv --log-file gpurun_out/ncu_aux_r02.csv python tools/aux_kernels.py > gpurun_out/aux_ncu.log 2>&1
tail -2 gpurun_out/aux_ncu.log; wc -l gpurun_out/ncu_aux_r02.csv
