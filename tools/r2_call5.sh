#!/bin/bash
# round 2, GPU call 5: fused pool->fc6 alone (own process, bounded), then tests + benches
o=gpurun_out
timeout 300 python -m pytest tests/test_gpu_pool_fc6.py -m gpu -q -x > $o/r2e_fc6.log 2>&1; echo "fc6 rc=$?" >> $o/r2e_fc6.log; tail -n 25 $o/r2e_fc6.log | cut -c1-200
if grep -q "fc6 rc=0" $o/r2e_fc6.log; then
  timeout 600 python tools/bench_pool_fc6.py 10 > $o/r2e_fc6_bench.json 2> $o/r2e_fc6_bench.err; cat $o/r2e_fc6_bench.json; tail -n 3 $o/r2e_fc6_bench.err
fi
timeout 1800 python -m pytest tests -m gpu -q --deselect tests/test_gpu_pool_fc6.py > $o/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2e_pytest.log; tail -n 6 $o/r2e_pytest.log | cut -c1-200
for t in "roi_pool_fwd_tma=1"; do for c in c2 c1; do python tools/bench_ops.py --config $c --iters 20 --tune $t >> $o/r2e_ab.jsonl 2>> $o/r2e_ab.err; done; done; cat $o/r2e_ab.jsonl
python bench.py --steps 20 --no-cpu-baseline --no-train > $o/r2e_bench_c2.json 2> $o/r2e_bench_c2.err; python -c "
import json; d=json.loads(open('$o/r2e_bench_c2.json').read().strip().splitlines()[-1]); print('c2', d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['e2e']['value'])"
python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/r2e_c4_launches.csv \
    python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline > $o/r2e_ncu_list.log 2>&1
