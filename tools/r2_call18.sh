#!/bin/bash
o=gpurun_out
for rows in 1 0; do
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:"roi_pool_fwd_rmq_kernel" -s 2 -c 1 python tools/bench_ops.py --config c2 --iters 2 --tune roi_pool_fwd_rows=$rows 2>&1 | grep -E "inst_executed|time_duration|issue_active|lsu_wavefronts"
echo ---
done
