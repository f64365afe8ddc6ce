#!/bin/bash
# round 2, GPU call 6: suite, default bench (the driver's command), c4, ncu --set full of NMS / fused fc6 / pooling
o=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > $o/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2f_pytest.log; tail -n 6 $o/r2f_pytest.log | cut -c1-200
python bench.py > $o/r2f_bench_c2.json 2> $o/r2f_bench_c2.err; tail -c 900 $o/r2f_bench_c2.json; tail -n 3 $o/r2f_bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > $o/r2f_bench_ref.json 2> $o/r2f_bench_ref.err; cut -c1-300 $o/r2f_bench_ref.json
python bench.py --config c4 --steps 10 > $o/r2f_bench_c4.json 2> $o/r2f_bench_c4.err; python -c "
import json; d=json.loads(open('$o/r2f_bench_c4.json').read().strip().splitlines()[-1]); print('c4', d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'])"
python tools/bench_pool_fc6.py 3 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"pool_fc6_kernel" -s 1 -c 1 -o $o/r2f_fc6 python tools/bench_pool_fc6.py 3 > $o/r2f_ncu_fc6.log 2>&1
python tools/nms_vs_tv.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"nms_segment_kernel" -s 2 -c 2 -o $o/r2f_nms python tools/nms_vs_tv.py > $o/r2f_ncu_nms.log 2>&1
python tools/bench_ops.py --config c2 --iters 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_fwd_rmq|roi_pool_bwd_vec4" -s 4 -c 2 -o $o/r2f_pool python tools/bench_ops.py --config c2 --iters 2 > $o/r2f_ncu_pool.log 2>&1
ls -la $o | tail -n 8
