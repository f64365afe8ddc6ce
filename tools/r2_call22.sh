#!/bin/bash
o=gpurun_out
python tools/train_probe.py 10 graph > $o/r2u_train_graph.log 2>&1; tail -30 $o/r2u_train_graph.log | cut -c1-250
