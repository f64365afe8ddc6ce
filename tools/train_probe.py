"""Kernel-time breakdown of one data-parallel training step (bench.py's train leg) at N = 1: run under
`ncu --metrics gpu__time_duration.sum` (tools/r2_call19.sh) or stand-alone for the CUDA-event time."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from jtsm_b200 import synth
from jtsm_b200.trainstep import TrainStep

dev = torch.device("cuda", 0)
wl = synth.step_case("c2")
ts = TrainStep(wl["cfg"], dev, world=1, tf32=True, lr=1e-3)
batch = ts.make_batch(wl)
for _ in range(3):
    ts.step(batch)
torch.cuda.synchronize()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    ts.step(batch)
e1.record()
torch.cuda.synchronize()
print("ms_per_step", e0.elapsed_time(e1) / n)

if len(sys.argv) > 2 and sys.argv[2] == "graph":
    import traceback

    ts2 = TrainStep(wl["cfg"], dev, world=1, tf32=True, lr=1e-3, fused_sgd=True)
    b2 = ts2.make_batch(wl)
    try:
        ts2.capture(b2)
        for _ in range(3):
            ts2.replay()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            ts2.replay()
        e1.record()
        torch.cuda.synchronize()
        print("graph ms_per_step", e0.elapsed_time(e1) / n, "loss", float(ts2.graph_loss))
    except Exception:
        traceback.print_exc()
