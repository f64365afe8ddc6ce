#!/bin/bash
# round 2, GPU call 10: folded 7x7 backward — parity tests, A/B timing (algo 0 = lane-owns-49, 3 = vec4), super-group sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_roi_pool.py tests/test_gpu_fullsize.py tests/test_gpu_hotpath.py -q -m gpu > gpurun_out/r2j_pool_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2j_pool_tests.log
: > gpurun_out/r2j_ab.jsonl
for cfg in c2 c1 c3; do
  for algo in 0 3; do
    timeout 300 python tools/bench_ops.py --config $cfg --iters 20 --tune roi_pool_bwd_algo=$algo >> gpurun_out/r2j_ab.jsonl 2>> gpurun_out/r2j_ab.err
  done
done
for mb in 4 8 16 64; do
  timeout 300 python tools/bench_ops.py --config c3 --iters 20 --tune roi_pool_bwd_super_mb=$mb >> gpurun_out/r2j_ab.jsonl 2>> gpurun_out/r2j_ab.err
done
for mb in 2 4 8; do
  timeout 300 python tools/bench_ops.py --config c2 --iters 20 --tune roi_pool_bwd_super_mb=$mb >> gpurun_out/r2j_ab.jsonl 2>> gpurun_out/r2j_ab.err
done
timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_bwd_algo=0 >> gpurun_out/r2j_ab.jsonl 2>> gpurun_out/r2j_ab.err
timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_bwd_algo=3 >> gpurun_out/r2j_ab.jsonl 2>> gpurun_out/r2j_ab.err
python - <<'PY'
import json
for l in open("gpurun_out/r2j_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
timeout 600 python bench.py --steps 30 --warmup 5 --no-train > gpurun_out/r2j_bench_c2.json 2> gpurun_out/r2j_bench_c2.err; python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2j_bench_c2.json') if l.startswith('{')][-1]); print('c2', d['value'], d['ms_per_step'], d['e2e']['value'])"
