#!/bin/bash
# round 2, GPU call 23 (2 GPUs): the DDP training step captured into a CUDA graph (NCCL all-reduce inside the graph)
o=gpurun_out
N=${1:-2}
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > $o/r2v_bench_n$N.json 2> $o/r2v_bench_n$N.err; echo "n$N rc=$?"
python - <<PY
import json
for l in open("gpurun_out/r2v_bench_n$N.json"):
    l=l.strip()
    if l.startswith("{"):
        d=json.loads(l); print({k:d.get(k) for k in ("value","n_gpus","ms_per_step","train_images_per_sec","allreduce_ms")}); print(d.get("e2e",{}).get("value")); print(d.get("train_step"))
PY
tail -5 $o/r2v_bench_n$N.err | cut -c1-300
