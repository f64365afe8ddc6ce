#!/bin/bash
# round 2, GPU call 17: row-mapped main loop of the range-max kernel — parity + A/B
o=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_roi_pool.py tests/test_gpu_hotpath.py tests/test_gpu_wsl_pool.py -q -m gpu -x > $o/r2q_tests.log 2>&1; echo "tests rc=$?"; tail -5 $o/r2q_tests.log
: > $o/r2q_ab.jsonl
for cfg in c2 c1 c3; do for rows in 1 0; do
  timeout 300 python tools/bench_ops.py --config $cfg --iters 20 --tune roi_pool_fwd_rows=$rows >> $o/r2q_ab.jsonl 2>> $o/r2q_ab.err
done; done
for rows in 1 0; do timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_fwd_rows=$rows >> $o/r2q_ab.jsonl 2>> $o/r2q_ab.err; done
python - <<'PY'
import json
for l in open("gpurun_out/r2q_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
timeout 600 python bench.py --steps 30 --warmup 5 --no-train > $o/r2q_bench_c2.json 2> $o/r2q_bench_c2.err; python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2q_bench_c2.json') if l.startswith('{')][-1]); print('c2', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
