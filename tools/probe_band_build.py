"""How much of a banded pooling call is table building?  Same map, RoI counts 64 / 4000: with 64 RoIs the kernel time is
(almost) only unpack + build of every (band, plane)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
res = []
for (h, w) in ((75, 100), (108, 144), (150, 200)):
    g = synth.generator(3)
    x = synth.feature_map(g, 1, 512, h, w).to(dev)
    for r in (64, 4000):
        rois = synth.rois_from_boxes([synth.proposals(g, r, h * 8, w * 8)]).to(dev)
        ts = []
        for _ in range(6):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.roi_pool_forward(x, rois, 1 / 8.0, (7, 7), argmax_dtype=torch.int16 if h * w < 65535 else torch.int32)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        res.append(dict(map=[h, w], rois=r, us=ts[len(ts) // 2]))
print(json.dumps(res))
