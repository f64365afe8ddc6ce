#!/bin/bash
# round 2, GPU call 27: bench lines of the other configurations with the final kernels (table of DESIGN.md section 5)
o=gpurun_out
for cfg in c1 c3 c3m c2_r4000; do
  timeout 900 python bench.py --config $cfg --steps 20 --warmup 5 --no-train > $o/r2y_bench_$cfg.json 2> $o/r2y_bench_$cfg.err; echo "$cfg rc=$?"
done
python - <<'PY'
import json
for f in ("c1","c3","c3m","c2_r4000"):
    try:
        d=json.loads([l for l in open("gpurun_out/r2y_bench_%s.json"%f) if l.startswith("{")][-1])
        print(f, d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("roofline",{}).get("frac"), d.get("roofline",{}).get("avg_launch_ms"), (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f, "ERR", e)
PY
