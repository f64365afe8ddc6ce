#!/bin/bash
# round 2, GPU call 25: the driver's round-end sequence — GPU suite, smoke(), default bench, reference arm, c4
o=gpurun_out
timeout 1800 python -m pytest tests -q -m gpu > $o/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r2z_pytest.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $o/r2z_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $o/r2z_smoke.log
/usr/bin/time -v python bench.py > $o/r2z_bench_c2.json 2> $o/r2z_bench_c2.err; echo "bench rc=$?"; grep -E "Elapsed" $o/r2z_bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > $o/r2z_bench_ref.json 2> $o/r2z_bench_ref.err; echo "ref rc=$?"
python bench.py --config c4 --steps 10 > $o/r2z_bench_c4.json 2> $o/r2z_bench_c4.err; echo "c4 rc=$?"
python bench.py --dtype bf16 --no-cpu-baseline --no-train > $o/r2z_bench_c2_bf16.json 2> $o/r2z_bench_c2_bf16.err; echo "bf16 rc=$?"
python - <<'PY'
import json
for f in ("r2z_bench_c2","r2z_bench_ref","r2z_bench_c4","r2z_bench_c2_bf16"):
    d=json.loads([l for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1])
    print(f, d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("roofline",{}).get("frac"), d.get("train_images_per_sec"), d.get("gpu_launches"), (d.get("cpu_baseline") or {}).get("value"))
PY
