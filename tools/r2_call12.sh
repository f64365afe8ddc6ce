#!/bin/bash
# round 2, GPU call 12 (N GPUs): the bench line at N ranks (hot-path leg + DDP training leg)
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29527 \
  bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/r2l_bench_n$N.json 2> gpurun_out/r2l_bench_n$N.err; echo "n$N rc=$?"
python - <<PY
import json
for l in open("gpurun_out/r2l_bench_n$N.json"):
    l=l.strip()
    if l.startswith("{"):
        d=json.loads(l); print({k:d.get(k) for k in ("value","n_gpus","ms_per_step","train_images_per_sec","allreduce_ms")}); print(d.get("e2e",{}).get("value")); print(d.get("train_step"))
PY
tail -3 gpurun_out/r2l_bench_n$N.err
