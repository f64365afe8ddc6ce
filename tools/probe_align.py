"""One RoIAlign forward/backward pair at a config (for ncu): python tools/probe_align.py [c2] [sampling_ratio]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
sr = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda:0")
cfg, feat, boxes, rois, obj = synth.pool_case(name)
x, rois_d = feat.to(dev), rois.to(dev)
n, c, h, w = x.shape
p = cfg["P"]
for _ in range(3):
    out = ops.roi_align_forward(x, rois_d, cfg["scale"], p, p, sr, True)
    gi = ops.roi_align_backward(out, rois_d, cfg["scale"], p, p, n, c, h, w, sr, True)
torch.cuda.synchronize()
print("ok", float(out.sum()), float(gi.sum()))
