#!/bin/bash
# round 2, GPU call 15: packed 16-bit tables — parity, then timing (bf16 / fp16, c2 / c1 / c3) against the index tables
o=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_roi_pool.py tests/test_gpu_hotpath.py -q -m gpu -x > $o/r2o_tests.log 2>&1; echo "tests rc=$?"; tail -15 $o/r2o_tests.log
: > $o/r2o_ab.jsonl
for cfg in c2 c1; do for dt in bf16 fp16; do for pk in 1 0; do
  timeout 300 python tools/bench_ops.py --config $cfg --dtype $dt --iters 20 --tune roi_pool_fwd_pack16=$pk >> $o/r2o_ab.jsonl 2>> $o/r2o_ab.err
done; done; done
for pk in 1 0; do timeout 300 python tools/bench_ops.py --config c3 --dtype bf16 --iters 10 --tune roi_pool_fwd_pack16=$pk >> $o/r2o_ab.jsonl 2>> $o/r2o_ab.err; done
python - <<'PY'
import json
for l in open("gpurun_out/r2o_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
