"""Does a concurrent H2D copy slow the pooling kernels?  RoIPool forward / backward alone at c2, with and without a
16 MB pinned-host upload running on another stream.  Prints us per launch."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402

dev = torch.device("cuda:0")
cfg, feat, boxes, rois, obj = synth.pool_case("c2")
x, rois_d = feat.to(dev), rois.to(dev)
P = (cfg["P"], cfg["P"])
out, arg = ops.roi_pool_forward(x, rois_d, cfg["scale"], P, argmax_dtype=torch.int16)
g = torch.randn_like(out)
src = torch.empty(16 << 20, dtype=torch.uint8).pin_memory()
dst = torch.empty(16 << 20, dtype=torch.uint8, device=dev)
copy_stream = torch.cuda.Stream()
N = 100


def run(fn, with_copy):
    gr = torch.cuda.CUDAGraph()
    fn()
    torch.cuda.synchronize()
    with torch.cuda.graph(gr):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N):
        if with_copy:
            with torch.cuda.stream(copy_stream):
                dst.copy_(src, non_blocking=True)
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / N * 1e3


for name, fn in [("roi_pool fwd", lambda: ops.roi_pool_forward(x, rois_d, cfg["scale"], P, argmax_dtype=torch.int16)),
                 ("roi_pool bwd", lambda: ops.roi_pool_backward(g, rois_d, arg, x.shape, P))]:
    a, b = run(fn, False), run(fn, True)
    print("%-13s alone %.1f us, with a concurrent 16 MB H2D per launch %.1f us" % (name, a, b))
