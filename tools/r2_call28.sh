#!/bin/bash
# round 2, GPU call 28: dense bin-id band lists — parity, banded timings, c4 bench, 16-bit packed banded
o=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_roi_pool.py tests/test_gpu_hotpath.py tests/test_gpu_infer_step.py tests/test_gpu_wsl_pool.py -q -m gpu -x > $o/r2x_tests.log 2>&1; echo "tests rc=$?"; tail -8 $o/r2x_tests.log
python tools/probe_band_build.py 2>&1 | tail -1
for bl in 1 0; do
timeout 600 python bench.py --config c4 --steps 10 --no-cpu-baseline --tune roi_pool_fwd_bandlist=$bl > $o/r2x_bench_c4_bl$bl.json 2> $o/r2x_bench_c4_bl$bl.err; python -c "
import json; d=json.loads([l for l in open('$o/r2x_bench_c4_bl$bl.json') if l.startswith('{')][-1]); print('c4 bandlist=$bl', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
: > $o/r2x_ab.jsonl
for pk in 2 0; do
  timeout 300 python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune roi_pool_fwd_pack16=$pk >> $o/r2x_ab.jsonl 2>> $o/r2x_ab.err
done
timeout 300 python tools/bench_ops.py --config c3 --dtype bf16 --iters 10 --tune roi_pool_fwd_pack16=2 >> $o/r2x_ab.jsonl 2>> $o/r2x_ab.err
timeout 300 python tools/bench_ops.py --config c3 --dtype bf16 --iters 10 --tune roi_pool_fwd_pack16=0 >> $o/r2x_ab.jsonl 2>> $o/r2x_ab.err
python - <<'PY'
import json
for l in open("gpurun_out/r2x_ab.jsonl"):
    d=json.loads(l); print(d["config"], d["dtype"], d["tune"], round(d["roi_pool_fwd_us"],1), round(d["roi_pool_bwd_us"],1))
PY
