"""Config 5 of BASELINE.json: RoI-op stress sweep -- 1k-16k proposals x stride 8/16/32 maps x 7x7 / 14x14 bins,
fp32 / bf16, forward + backward -- on one GPU.  Per point: op times (CUDA events, L2 flushed) of RoIPool and RoIAlign
(sampling_ratio 2) beside torchvision's CUDA ops, and a parity check against them on the same inputs
(RoIPool: values bit-equal in fp32; RoIAlign: rel 1e-5 fp32 / 1e-2 bf16; gradients by tolerance).
Usage: python tools/sweep_c5.py [--iters 5] [--quick] > sweep.jsonl"""
import argparse
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402


def timeit(fn, iters, flush):
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


def close(a, b, rtol, what):
    a, b = a.float(), b.float()
    tol = rtol * max(1.0, float(b.abs().max()))
    err = float((a - b).abs().max())
    if not err <= tol + rtol * float(b.abs().max()):
        raise AssertionError("%s: max abs err %g > %g" % (what, err, tol))
    return err


def main():
    import torchvision

    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    Rs = (1000, 4000) if a.quick else (1000, 2000, 4000, 8000, 16000)
    point = 0
    for stride in (8, 16, 32):
        h, w = math.ceil(600 / stride), math.ceil(800 / stride)
        for P in (7, 14):
            for R in Rs:
                for dt in (torch.float32, torch.bfloat16):
                    g = synth.generator(synth.seed_for(5, point))
                    point += 1
                    x = synth.feature_map(g, 1, 512, h, w).to(dev).to(dt)
                    rois = synth.rois_from_boxes([synth.proposals(g, R, 600, 800)]).to(dev)
                    scale = 1.0 / stride
                    rec = dict(stride=stride, map=[h, w], P=P, R=R, dtype=str(dt).split(".")[1])
                    # ---- RoIPool
                    adt = torch.int16
                    out, arg = ops.roi_pool_forward(x, rois, scale, (P, P), argmax_dtype=adt)
                    go = torch.randn(out.shape, generator=g).to(dev).to(dt) if R <= 4000 else torch.ones_like(out)
                    # torchvision's CUDA ops have no bf16 kernels: the reference side always runs in fp32 on the same
                    # (bf16-rounded) values; a max-pool of bf16 values is exact, so RoIPool stays bit-equal
                    fp32 = dt == torch.float32
                    xt = x.float().clone().requires_grad_()
                    ot = torchvision.ops.roi_pool(xt, rois, (P, P), scale)
                    assert torch.equal(out.float(), ot.detach()), "roi_pool values differ from torchvision CUDA"
                    gi = ops.roi_pool_backward(go, rois, arg, x.shape, (P, P))
                    (gt,) = torch.autograd.grad(ot, xt, go.float(), retain_graph=True)
                    rec["roi_pool_bwd_err"] = close(gi, gt, 1e-5 if dt == torch.float32 else 2e-2, "roi_pool backward")
                    rec["roi_pool_fwd_us"] = timeit(lambda: ops.roi_pool_forward(x, rois, scale, (P, P), argmax_dtype=adt), a.iters, flush)
                    rec["roi_pool_bwd_us"] = timeit(lambda: ops.roi_pool_backward(go, rois, arg, x.shape, (P, P)), a.iters, flush)
                    if fp32:
                        rec["tv_roi_pool_fwd_us"] = timeit(lambda: torchvision.ops.roi_pool(x, rois, (P, P), scale), a.iters, flush)
                        rec["tv_roi_pool_bwd_us"] = timeit(lambda: torch.autograd.grad(ot, xt, go, retain_graph=True), a.iters, flush)
                    del ot, gt, gi, out, arg
                    # ---- RoIAlign, sampling_ratio 2, aligned
                    oa = ops.roi_align_forward(x, rois, scale, P, P, 2, True)
                    ota = torchvision.ops.roi_align(xt, rois, (P, P), scale, 2, True)
                    rec["roi_align_fwd_err"] = close(oa, ota.detach(), 1e-5 if dt == torch.float32 else 1e-2, "roi_align forward")
                    ga = ops.roi_align_backward(go, rois, scale, P, P, 1, 512, h, w, 2, True)
                    (gta,) = torch.autograd.grad(ota, xt, go.float(), retain_graph=True)
                    rec["roi_align_bwd_err"] = close(ga, gta, 1e-5 if dt == torch.float32 else 2e-2, "roi_align backward")
                    rec["roi_align_fwd_us"] = timeit(lambda: ops.roi_align_forward(x, rois, scale, P, P, 2, True), a.iters, flush)
                    rec["roi_align_bwd_us"] = timeit(lambda: ops.roi_align_backward(go, rois, scale, P, P, 1, 512, h, w, 2, True), a.iters, flush)
                    if fp32:
                        rec["tv_roi_align_fwd_us"] = timeit(lambda: torchvision.ops.roi_align(x, rois, (P, P), scale, 2, True), a.iters, flush)
                        rec["tv_roi_align_bwd_us"] = timeit(lambda: torch.autograd.grad(ota, xt, go, retain_graph=True), a.iters, flush)
                    del ota, gta, ga, oa, go, xt
                    torch.cuda.empty_cache()
                    print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
