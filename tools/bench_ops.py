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
    ap.add_argument("--ops", default="pool", help="pool | all | tta")
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
            if sr > 0:
                med, _ = timeit(lambda: ops.roi_align_backward(g, rois_d, cfg["scale"], ph, pw, n, c, h, w, sr, True,
                                                               deterministic=True), a.iters, flush)
                res["roi_align_bwd_deterministic_sr%d_us" % sr] = med
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
        gen = torch.Generator().manual_seed(6)
        sp2, oh2, _, _ = synth.superpixel_case(gen, boxes, cfg["img_h"], cfg["img_w"], 22)
        sp2_d, oh2_d = sp2.to(dev), oh2.to(dev)
        med, _ = timeit(lambda: ops.moi_pool_forward(x, rois_d, cfg["scale"], ph, pw, oh2_d, sp2_d), a.iters, flush)
        res["moi_pool_fwd_box_consistent_masks_us"] = med
    if a.ops in ("all", "tta"):
        res.update(bench_tta(dev, a.iters, flush))
    print(json.dumps(res))


def bench_tta(dev, iters, flush):
    """c4: R = 4000, K = 20, 10 augmentations: device merge vs the reference's composition on the same GPU
    (per augmentation D2H -> numpy apply_box -> H2D, then torch.cat + torch.mean; test_time_augmentation_avg.py:364-390)."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import numpy as np

    from jtsm_b200 import ops as _ops
    from jtsm_b200 import tta

    g = synth.generator(synth.seed_for(4))
    hw = (600, 800)
    sizes = (480, 576, 688, 864, 1200)
    tf = tta.tta_transforms(hw, hw, sizes, 4000, True)
    pb = synth.proposals(g, 4000, *hw).to(dev)
    boxes, scores = [], []
    for t in tf:  # per-augmentation outputs: the proposals mapped into the augmented image, per-class jitter
        fwd = t.apply_box(pb)
        pc = fwd[:, None, :].repeat(1, 20, 1) + (torch.rand(4000, 20, 4, generator=g).to(dev) - 0.5) * 2.0
        lo, hi = torch.minimum(pc[..., :2], pc[..., 2:]), torch.maximum(pc[..., :2], pc[..., 2:]) + 1.0
        boxes.append(torch.cat([lo, hi], -1).reshape(1, 4000, 80).contiguous())
        scores.append(torch.softmax(torch.randn(4000, 21, generator=g).to(dev) * 3, 1).reshape(1, 4000, 21))

    def host_apply_box(stages, b):
        """What the reference does per augmentation on the host (fvcore apply_box per stage, numpy float32)."""
        for ty, fa, fb in stages:
            x0, y0, x1, y1 = b[:, 0].copy(), b[:, 1].copy(), b[:, 2].copy(), b[:, 3].copy()
            if ty == _ops.TTA_HFLIP:
                x0, x1 = np.float32(fa) - x0, np.float32(fa) - x1
            elif ty == _ops.TTA_SCALE:
                x0, x1, y0, y1 = x0 * np.float32(fa), x1 * np.float32(fa), y0 * np.float32(fb), y1 * np.float32(fb)
            b = np.stack([np.minimum(x0, x1), np.minimum(y0, y1), np.maximum(x0, x1), np.maximum(y0, y1)], 1)
        return b

    bd, sd = [b.to(dev) for b in boxes], [s.to(dev) for s in scores]
    out = {}
    med, _ = timeit(lambda: tta.get_augmented_boxes(bd, sd, tf), iters, flush)
    out["tta_merge_call_us"] = med  # includes the Python / ctypes call overhead (GPU idle in between)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        tta.get_augmented_boxes(bd, sd, tf)
    med, _ = timeit(gr.replay, iters, flush)
    out["tta_merge_kernel_us"] = med
    nbytes = (len(bd) + 1) * (bd[0].numel() + sd[0].numel()) * 4
    out["tta_merge_algorithmic_GBps"] = nbytes / med / 1e3

    def ref_path():
        ab = list(bd)
        for i, t in enumerate(tf):
            pbx = ab[i]
            o = host_apply_box(t.inverse().stages(), pbx.reshape(-1, 4).cpu().numpy())
            ab[i] = torch.from_numpy(o).to(dev).reshape(pbx.shape)
        return torch.mean(torch.cat(ab, 0), 0), torch.mean(torch.cat(sd, 0), 0)

    med, _ = timeit(ref_path, max(3, iters // 4), flush)
    out["tta_merge_reference_composition_us"] = med
    med, _ = timeit(lambda: tta.inference_one_image_boxes(bd, sd, tf, hw, 1e-5, 0.3, 100), iters, flush)
    out["tta_one_image_boxes_us"] = med
    return out


if __name__ == "__main__":
    main()
