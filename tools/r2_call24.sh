#!/bin/bash
o=gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/r2w_c4_launches.csv \
  python bench.py --config c4 --steps 1 --warmup 3 --no-graph --no-cpu-baseline > $o/r2w_ncu_list.log 2>&1; echo "list rc=$?"
