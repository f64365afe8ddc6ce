#!/bin/bash
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "packed or banded" > $o/r2x2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $o/r2x2_tests.log
: > $o/r2x2_ab.jsonl
for pk in 2 0; do
  timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_fwd_pack16=$pk >> $o/r2x2_ab.jsonl 2>> $o/r2x2_ab.err
  timeout 300 python tools/bench_ops.py --config c2 --dtype fp16 --iters 20 --tune roi_pool_fwd_pack16=$pk >> $o/r2x2_ab.jsonl 2>> $o/r2x2_ab.err
done
timeout 300 python tools/bench_ops.py --config c3 --dtype bf16 --iters 10 --tune roi_pool_fwd_pack16=2 >> $o/r2x2_ab.jsonl 2>> $o/r2x2_ab.err
python - <<'PY'
import json
for l in open("gpurun_out/r2x2_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
