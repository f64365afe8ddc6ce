#!/bin/bash
# round 2, GPU call 21: whole training step as a CUDA graph — test, then the bench train leg at N = 1
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainstep.py tests/test_gpu_head_modules.py -q -m gpu -x > $o/r2t_tests.log 2>&1; echo "tests rc=$?"; tail -3 $o/r2t_tests.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $o/r2t_bench_c2.json 2> $o/r2t_bench_c2.err; echo "bench rc=$?"; python -c "
import json; d=json.loads([l for l in open('$o/r2t_bench_c2.json') if l.startswith('{')][-1]); print(d['value'], d['ms_per_step'], d['train_images_per_sec']); print(d['train_step'])"; tail -3 $o/r2t_bench_c2.err
