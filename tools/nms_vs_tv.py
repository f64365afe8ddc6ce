"""NMS stage timing beside torchvision CUDA: 80 000 x 20 classes (c4) and 4 000 x 1 class.
CUDA events, L2 flushed, median of --iters.  Keep lists are compared as well."""
import json
import os
import sys

import torch
import torchvision

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    g = synth.generator(5)
    res = {}
    for name, r, k in (("80000x20", 4000, 20), ("4000x1", 4000, 1), ("16000x80", 200, 80)):
        props = synth.proposals(g, r, 600, 800)
        boxes = (props[:, None, :] + torch.randn(r, k, 4, generator=g) * 4).reshape(-1, 4)
        boxes = torch.stack([boxes[:, 0].clamp(0, 800), boxes[:, 1].clamp(0, 600), boxes[:, 2].clamp(0, 800),
                             boxes[:, 3].clamp(0, 600)], 1)
        scores = torch.rand(r * k, generator=g)
        cls = torch.arange(k).repeat(r)
        bd, sd, cd = boxes.to(dev), scores.to(dev), cls.to(dev)
        ours = ops.batched_nms(bd, sd, cd, 0.3)
        tv = torchvision.ops.boxes._batched_nms_vanilla(bd, sd, cd, 0.3)
        res[name] = dict(n=r * k, classes=k, kept=int(ours.numel()), same_keep_as_torchvision_cuda=bool(torch.equal(ours, tv)),
                         ours_raw_us=timeit(lambda: ops.batched_nms_raw(bd, sd, cd, 0.3), 10, flush),
                         torchvision_cuda_vanilla_us=timeit(lambda: torchvision.ops.boxes._batched_nms_vanilla(bd, sd, cd, 0.3), 10, flush),
                         torchvision_cuda_coordinate_trick_us=timeit(lambda: torchvision.ops.boxes._batched_nms_coordinate_trick(bd, sd, cd, 0.3), 10, flush))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
