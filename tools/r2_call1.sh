#!/bin/bash
# round 2, GPU call 1: keyed range-max kernel -- parity gate, suite, A/B timing, ncu of the new kernel
o=gpurun_out
timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -k "keyed or nonfinite" --maxfail=5 > $o/r2a_gate.log 2>&1
echo "gate rc=$?" >> $o/r2a_gate.log; tail -5 $o/r2a_gate.log
if ! grep -q "gate rc=0" $o/r2a_gate.log; then echo "GATE FAILED"; exit 1; fi
timeout 1500 python -m pytest tests -m gpu -q > $o/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> $o/r2a_pytest.log; tail -8 $o/r2a_pytest.log
for t in "roi_pool_fwd_keys=0" "roi_pool_fwd_keys=1 roi_pool_fwd_tma=0" "roi_pool_fwd_keys=1 roi_pool_fwd_tma=1"; do
  for c in c2 c1; do
    python tools/bench_ops.py --config $c --iters 20 --tune $t >> $o/r2a_ab.jsonl 2>> $o/r2a_ab.err
  done
  python tools/bench_ops.py --config c2 --dtype bf16 --iters 20 --tune $t >> $o/r2a_ab.jsonl 2>> $o/r2a_ab.err
done
cat $o/r2a_ab.jsonl
python tools/bench_ops.py --config c2 --iters 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_fwd_rmqk" -s 2 -c 1 -o $o/r2a_rmqk \
    python tools/bench_ops.py --config c2 --iters 2 > $o/r2a_ncu.log 2>&1
ls -la $o | tail
