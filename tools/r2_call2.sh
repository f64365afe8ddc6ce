#!/bin/bash
# round 2, GPU call 2: no-alloc entry point alone + after the suite; unified rmq kernel frame; op timings
o=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $o/r2b_noalloc_alone.log 2>&1; echo "alone rc=$?" >> $o/r2b_noalloc_alone.log; tail -3 $o/r2b_noalloc_alone.log
timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -k "keyed or nonfinite" > $o/r2b_gate.log 2>&1; echo "gate rc=$?" >> $o/r2b_gate.log; tail -3 $o/r2b_gate.log
if ! grep -q "gate rc=0" $o/r2b_gate.log; then echo "GATE FAILED"; tail -40 $o/r2b_gate.log; exit 1; fi
timeout 1800 python -m pytest tests -m gpu -q -x > $o/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2b_pytest.log; tail -6 $o/r2b_pytest.log
for t in "roi_pool_fwd_tma=1" "roi_pool_fwd_tma=0" "roi_pool_fwd_keys=1"; do
  for c in c2 c1; do python tools/bench_ops.py --config $c --iters 20 --tune $t >> $o/r2b_ab.jsonl 2>> $o/r2b_ab.err; done
  python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune $t >> $o/r2b_ab.jsonl 2>> $o/r2b_ab.err
done
cat $o/r2b_ab.jsonl
python bench.py --no-cpu-baseline --steps 20 > $o/r2b_bench_c2.json 2> $o/r2b_bench_c2.err; tail -c 1500 $o/r2b_bench_c2.json
