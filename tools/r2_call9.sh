#!/bin/bash
# round 2, GPU call 9 (2 GPUs): the data-parallel bench line (DDP training leg, NCCL all-reduce) + reference arm at N=2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2i_gpus.txt
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2i_bench_n2.json 2> gpurun_out/r2i_bench_n2.err; echo "n2 rc=$?"
grep -c "NCCL INFO" gpurun_out/r2i_bench_n2.err; grep -m3 "NVLS\|via P2P\|Connected all" gpurun_out/r2i_bench_n2.err
python - <<'PY'
import json
for l in open("gpurun_out/r2i_bench_n2.json"):
    l=l.strip()
    if l.startswith("{"):
        d=json.loads(l); print({k:d.get(k) for k in ("value","n_gpus","ms_per_step","train_images_per_sec","allreduce_ms")}); print(d.get("e2e",{}).get("value")); print(d.get("train_step"))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 \
  bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2i_ref_n2.json 2> gpurun_out/r2i_ref_n2.err; echo "ref rc=$?"; tail -c 600 gpurun_out/r2i_ref_n2.json
timeout 600 python -m pytest tests/test_gpu_trainstep.py -q -m gpu 2>&1 | tail -3
