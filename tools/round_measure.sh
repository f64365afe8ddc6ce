#!/bin/bash
# Round-end measurements on one B200 box (run under gpurun): GPU tests, default bench, reference arm, config sweep,
# launch list + ncu --set full of the step kernels.  Usage: bash tools/round_measure.sh <tag>
tag=${1:-r1u}
o=gpurun_out
python -m pytest tests -m gpu -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log
python bench.py > $o/${tag}_bench_c2.json 2> $o/${tag}_bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
for c in c1 c3 c4 c3m; do python bench.py --config $c --no-cpu-baseline --steps 20 --warmup 3 > $o/${tag}_bench_$c.json 2> $o/${tag}_bench_$c.err; done
python bench.py --dtype bf16 --no-cpu-baseline > $o/${tag}_bench_c2_bf16.json 2> $o/${tag}_bench_c2_bf16.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > $o/${tag}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_fwd_rmq|roi_pool_bwd_vec4" -c 2 -o $o/${tag}_step_kernels \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > $o/${tag}_ncu_full.log 2>&1
python tools/bench_ops.py --ops all --tv --iters 10 > $o/${tag}_ops.json 2>&1
ls -la $o | tail -20
