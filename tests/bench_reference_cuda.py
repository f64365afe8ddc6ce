"""Timing of the REFERENCE'S OWN CUDA OP PATH on the same GPU, for the >= 5x comparison of BASELINE.md 2.1
(not a pytest module; run by hand on the GPU box):
  pooling   = torchvision CUDA roi_pool fwd + bwd (what POOLER_TYPE "ROIPool" runs, wsl/modeling/poolers.py:183-186)
              followed by the separate objectness multiply (roi_heads_oicr.py:393-394);
  head math = the reference's eager PyTorch composition (restated in oracle/ref_heads.py and pinned against the
              reference's code) executed on CUDA tensors -- every small kernel and host sync it implies included.
Prints one JSON line with us per step for both and for jtsm_b200 (eager and CUDA-graph)."""
import json
import os
import sys

import torch
import torchvision

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import hotpath, synth  # noqa: E402
from oracle import ref_heads  # noqa: E402


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    dev = torch.device("cuda:0")
    wl = synth.step_case(name)
    cfg = wl["cfg"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    x = wl["feat"].to(dev)
    rois = wl["rois"].to(dev)
    obj = (wl["objectness"] + 1).to(dev)
    g = synth.expand_grad_tile(wl["grad_tile"].to(dev), rois.shape[0])
    P = (cfg["P"], cfg["P"])
    head_in = dict(mil_cls=wl["mil_cls"].to(dev), mil_det=wl["mil_det"].to(dev),
                   ref_logits=[t.to(dev) for t in wl["ref_logits"]], ref_deltas=[t.to(dev) for t in wl["ref_deltas"]],
                   proposal_boxes=[b.to(dev) for b in wl["boxes"]], labels=[l.to(dev) for l in wl["labels"]],
                   num_classes=cfg["K"], reg=cfg["reg"])

    def ref_pool():
        xg = x.detach().requires_grad_()
        out = torchvision.ops.roi_pool(xg, rois, P, cfg["scale"]) * obj.view(-1, 1, 1, 1)
        out.backward(g)

    def ref_heads_step():
        ref_heads.head_step(**head_in)

    step = hotpath.HotPathStep(cfg, dev)
    host = step.pin_host_inputs(wl)
    res = step.upload(host)
    out = {"config": name, "rois": int(rois.shape[0])}
    out["ref_cuda_pool_us"] = timed(ref_pool, 10, flush)
    out["ref_cuda_heads_us"] = timed(ref_heads_step, 10, flush)
    out["ref_cuda_step_us"] = out["ref_cuda_pool_us"] + out["ref_cuda_heads_us"]
    out["b200_eager_step_us"] = timed(lambda: step.run(res), 10, flush)
    step.capture(res)
    out["b200_graph_step_us"] = timed(step.replay, 20, flush)
    out["speedup_vs_ref_cuda"] = out["ref_cuda_step_us"] / out["b200_graph_step_us"]
    out["rois_per_s_ref_cuda"] = out["rois"] / out["ref_cuda_step_us"] * 1e6
    out["rois_per_s_b200"] = out["rois"] / out["b200_graph_step_us"] * 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
