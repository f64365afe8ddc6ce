#!/bin/bash
o=gpurun_out
python tools/train_probe.py 10 > $o/r2r_train.log 2>&1; tail -2 $o/r2r_train.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $o/r2r_train_launches.csv python tools/train_probe.py 1 > $o/r2r_ncu.log 2>&1; echo "ncu rc=$?"
