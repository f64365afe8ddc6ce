"""N > 1 host logic on CPU: two gloo ranks run bench.py's reference arm and rank/seed plumbing.
 - `bench.py --impl reference` under torchrun semantics: rank 0 prints ONE json line, rank 1 exits 0 silently;
 - per-rank workloads differ (seed + rank) and the max-over-ranks reduction used for timing works on gloo."""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from jtsm_b200 import synth

    wl = synth.step_case("c1", sweep_index=rank, R=64, C=4)
    t = torch.tensor([float(rank + 1), float(wl["rois"].sum())], dtype=torch.float64)
    mx = t.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    gathered = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    if rank == 0:
        out.put((mx.tolist(), [g.tolist() for g in gathered]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gloo_ranks_have_distinct_shards_and_max_reduce():
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    mx, gathered = q.get()
    assert mx[0] == 2.0  # max over ranks of the per-rank "time"
    assert gathered[0][1] != gathered[1][1]  # each rank owns a different seeded image batch


def test_reference_arm_prints_on_rank0_only():
    env = dict(os.environ, WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29999")
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
           "--warmup", "1", "--config", "c1", "--cpu-sample-rois", "64"]
    r1 = subprocess.run(cmd, env=dict(env, RANK="1", LOCAL_RANK="1"), capture_output=True, text=True, timeout=300)
    assert r1.returncode == 0 and r1.stdout.strip() == ""
    r0 = subprocess.run(cmd, env=dict(env, RANK="0", LOCAL_RANK="0"), capture_output=True, text=True, timeout=300)
    assert r0.returncode == 0
    line = json.loads(r0.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "reference(pool)+port(heads)"
