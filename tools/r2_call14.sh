#!/bin/bash
# round 2, GPU call 14: launch list of the default bench command (kernel shares of the step) + ncu --set full of the two
# backward kernels (lane-owns-49 vs 4-wide)
o=gpurun_out
python bench.py --steps 3 --warmup 3 --no-train > $o/r2n_bench_c2.json 2> $o/r2n_bench_c2.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/r2n_c2_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-train --no-graph > $o/r2n_ncu_list.log 2>&1; echo "list rc=$?"
python tools/bench_ops.py --config c2 --iters 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_bwd_fold7" -s 2 -c 1 -o $o/r2n_bwd_fold python tools/bench_ops.py --config c2 --iters 2 > $o/r2n_ncu_bwd.log 2>&1
python tools/bench_ops.py --config c2 --iters 2 --tune roi_pool_bwd_algo=3 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_bwd_vec4" -s 2 -c 1 -o $o/r2n_bwd_vec4 python tools/bench_ops.py --config c2 --iters 2 --tune roi_pool_bwd_algo=3 > $o/r2n_ncu_bwd4.log 2>&1
ls -la $o | grep r2n
