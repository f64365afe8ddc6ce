"""Run MOIPool / ROILoopPool forward a few times at a named config (for `ncu --metrics gpu__time_duration.sum`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402

cfg, feat, boxes, rois, obj = synth.pool_case(sys.argv[1] if len(sys.argv) > 1 else "c2")
dev = torch.device("cuda:0")
x, rois_d = feat.to(dev), rois.to(dev)
gen = torch.Generator().manual_seed(5)
sp, S = synth.superpixels(gen, x.shape[0], cfg["img_h"], cfg["img_w"])
oh = synth.oh_labels(gen, rois.shape[0], S)
sp_d, oh_d = sp.to(dev), oh.to(dev)
sp2, oh2, _, _ = synth.superpixel_case(torch.Generator().manual_seed(6), boxes, cfg["img_h"], cfg["img_w"], 22)
sp2_d, oh2_d = sp2.to(dev), oh2.to(dev)
for _ in range(3):
    ops.moi_pool_forward(x, rois_d, cfg["scale"], cfg["P"], cfg["P"], oh2_d, sp2_d)  # box-consistent (dense) masks
    ops.moi_pool_forward(x, rois_d, cfg["scale"], cfg["P"], cfg["P"], oh_d, sp_d)
    ops.roi_loop_pool_forward(x, rois_d, cfg["scale"], cfg["P"], cfg["P"])
torch.cuda.synchronize()
print("ok")
