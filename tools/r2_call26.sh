#!/bin/bash
o=gpurun_out
s=$(date +%s); python bench.py > $o/r2z_bench_c2.json 2> $o/r2z_bench_c2.err; echo "bench rc=$? wall=$(( $(date +%s) - s ))s"
python - <<'PY'
import json
for f in ("r2z_bench_c2","r2z_bench_ref","r2z_bench_c4","r2z_bench_c2_bf16"):
    try:
        d=json.loads([l for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1])
        print(f, d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("roofline",{}).get("frac"), d.get("train_images_per_sec"), d.get("gpu_launches"), (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f, "ERR", e)
PY
