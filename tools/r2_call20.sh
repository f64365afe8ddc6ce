#!/bin/bash
# round 2, GPU call 20: per-band RoI lists — parity, then c4 bench and banded 16-bit timing; train-step kernel breakdown
o=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_roi_pool.py tests/test_gpu_hotpath.py tests/test_gpu_infer_step.py -q -m gpu -x > $o/r2s_tests.log 2>&1; echo "tests rc=$?"; tail -12 $o/r2s_tests.log
for bl in 1 0; do
timeout 600 python bench.py --config c4 --steps 10 --tune roi_pool_fwd_bandlist=$bl > $o/r2s_bench_c4_bl$bl.json 2> $o/r2s_bench_c4_bl$bl.err; python -c "
import json; d=json.loads([l for l in open('$o/r2s_bench_c4_bl$bl.json') if l.startswith('{')][-1]); print('c4 bandlist=$bl', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
: > $o/r2s_ab.jsonl
for pk in 2 0; do for bl in 1 0; do
  timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_fwd_pack16=$pk roi_pool_fwd_bandlist=$bl >> $o/r2s_ab.jsonl 2>> $o/r2s_ab.err
done; done
timeout 300 python tools/bench_ops.py --config c3 --dtype bf16 --iters 10 --tune roi_pool_fwd_pack16=2 >> $o/r2s_ab.jsonl 2>> $o/r2s_ab.err
python - <<'PY'
import json
for l in open("gpurun_out/r2s_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
bash tools/r2_call19.sh
