#!/bin/bash
# round 2, GPU call 3: head modules, device inference, c4 step; suite; benches; NMS vs torchvision; c4 launch list
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_head_modules.py tests/test_gpu_inference.py tests/test_gpu_infer_step.py tests/test_gpu_nms.py -m gpu -q > $o/r2c_new.log 2>&1; echo "new rc=$?" >> $o/r2c_new.log; tail -n 30 $o/r2c_new.log
timeout 1800 python -m pytest tests -m gpu -q > $o/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2c_pytest.log; tail -n 6 $o/r2c_pytest.log
python bench.py --steps 20 > $o/r2c_bench_c2.json 2> $o/r2c_bench_c2.err; tail -c 600 $o/r2c_bench_c2.json
python bench.py --config c4 --steps 10 > $o/r2c_bench_c4.json 2> $o/r2c_bench_c4.err; tail -c 2500 $o/r2c_bench_c4.json; tail -n 5 $o/r2c_bench_c4.err
python tools/nms_vs_tv.py > $o/r2c_nms.json 2> $o/r2c_nms.err; cat $o/r2c_nms.json; tail -n 3 $o/r2c_nms.err
python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/r2c_c4_launches.csv \
    python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline > $o/r2c_ncu_list.log 2>&1
ls -la $o | tail -n 12
