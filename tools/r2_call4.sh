#!/bin/bash
# round 2, GPU call 4: banded pooling, cluster NMS, train step; suite; benches
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_nms.py tests/test_gpu_inference.py tests/test_gpu_trainstep.py tests/test_gpu_head_modules.py -m gpu -q -k "banded or keyed or nonfinite or nms or inference or train or heads or modules" > $o/r2d_new.log 2>&1; echo "new rc=$?" >> $o/r2d_new.log; tail -n 30 $o/r2d_new.log | cut -c1-200
if grep -q "Xid\|illegal memory\|unspecified launch" $o/r2d_new.log; then echo "GPU FAULT in new tests"; exit 1; fi
timeout 1800 python -m pytest tests -m gpu -q > $o/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2d_pytest.log; tail -n 8 $o/r2d_pytest.log | cut -c1-200
python tools/nms_vs_tv.py > $o/r2d_nms.json 2> $o/r2d_nms.err; cat $o/r2d_nms.json; tail -n 3 $o/r2d_nms.err
python bench.py --config c4 --steps 10 --no-cpu-baseline > $o/r2d_bench_c4.json 2> $o/r2d_bench_c4.err; tail -c 1200 $o/r2d_bench_c4.json; tail -n 5 $o/r2d_bench_c4.err
python bench.py --steps 20 --no-cpu-baseline > $o/r2d_bench_c2.json 2> $o/r2d_bench_c2.err; tail -c 1500 $o/r2d_bench_c2.json; tail -n 5 $o/r2d_bench_c2.err
