"""Op-level timing on one GPU (CUDA events, L2 flushed between iterations).
Usage: python tools/bench_ops.py [--config c2] [--iters 20] [--tv] [--tune key=val ...]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import _lib, ops, synth  # noqa: E402


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
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--tv", action="store_true")
    ap.add_argument("--tune", nargs="*", default=[])
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--ops", default="pool", help="pool | all")
    a = ap.parse_args()
    for kv in a.tune:
        k, v = kv.split("=")
        _lib.set_tuning(k, int(v))
    dev = torch.device("cuda:0")
    cfg, feat, boxes, rois, obj = synth.pool_case(a.config)
    dt = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    x = feat.to(dev).to(dt)
    rois_d = rois.to(dev)
    P = (cfg["P"], cfg["P"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res = {"config": a.config, "dtype": a.dtype, "tune": a.tune}
    esz = x.element_size()
    out, arg = ops.roi_pool_forward(x, rois_d, cfg["scale"], P, argmax_dtype=torch.int16)
    g = torch.randn_like(out)
    bytes_fwd = x.numel() * esz + rois.numel() * 4 + out.numel() * esz
    med, best = timeit(lambda: ops.roi_pool_forward(x, rois_d, cfg["scale"], P, argmax_dtype=torch.int16), a.iters, flush)
    res["roi_pool_fwd_us"] = med
    res["roi_pool_fwd_GBs"] = bytes_fwd / med / 1e3
    med, best = timeit(lambda: ops.roi_pool_backward(g, rois_d, arg, x.shape, P), a.iters, flush)
    res["roi_pool_bwd_us"] = med
    res["roi_pool_bwd_GBs"] = bytes_fwd / med / 1e3
    if a.tv:
        import torchvision

        xg = x.clone().requires_grad_()
        med, _ = timeit(lambda: torchvision.ops.roi_pool(x, rois_d, P, cfg["scale"]), a.iters, flush)
        res["tv_roi_pool_fwd_us"] = med
        o = torchvision.ops.roi_pool(xg, rois_d, P, cfg["scale"])
        med, _ = timeit(lambda: torch.autograd.grad(o, xg, g, retain_graph=True), a.iters, flush)
        res["tv_roi_pool_bwd_us"] = med
    if a.ops == "all":
        n, c, h, w = x.shape
        ph = pw = cfg["P"]
        for sr in (2, 0):
            med, _ = timeit(lambda: ops.roi_align_forward(x, rois_d, cfg["scale"], ph, pw, sr, True), a.iters, flush)
            res["roi_align_fwd_sr%d_us" % sr] = med
            med, _ = timeit(lambda: ops.roi_align_backward(g, rois_d, cfg["scale"], ph, pw, n, c, h, w, sr, True), a.iters, flush)
            res["roi_align_bwd_sr%d_us" % sr] = med
            if a.tv:
                import torchvision
                med, _ = timeit(lambda: torchvision.ops.roi_align(x, rois_d, P, cfg["scale"], sr, True), a.iters, flush)
                res["tv_roi_align_fwd_sr%d_us" % sr] = med
                xg = x.clone().requires_grad_()
                o = torchvision.ops.roi_align(xg, rois_d, P, cfg["scale"], sr, True)
                med, _ = timeit(lambda: torch.autograd.grad(o, xg, g, retain_graph=True), a.iters, flush)
                res["tv_roi_align_bwd_sr%d_us" % sr] = med
        med, _ = timeit(lambda: ops.roi_loop_pool_forward(x, rois_d, cfg["scale"], ph, pw), a.iters, flush)
        res["roi_loop_pool_fwd_us"] = med
        gen = torch.Generator().manual_seed(5)
        sp, S = synth.superpixels(gen, n, cfg["img_h"], cfg["img_w"])
        oh = synth.oh_labels(gen, rois.shape[0], S)
        sp_d, oh_d = sp.to(dev), oh.to(dev)
        med, _ = timeit(lambda: ops.moi_pool_forward(x, rois_d, cfg["scale"], ph, pw, oh_d, sp_d), a.iters, flush)
        res["moi_pool_fwd_us"] = med
    print(json.dumps(res))


if __name__ == "__main__":
    main()
