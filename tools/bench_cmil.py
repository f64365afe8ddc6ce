"""Op times of the CMIL / PCL additions on the device next to the reference's own CPU code for the same call
(oracle/_ref/libref_wslcmil.so = ROIMerge_cpu.cpp + ROILabel_cpu.cpp compiled in place; the reference's heads also pay
the device -> host copies of S, the R x R IoU matrix, C and D, which are included on the CPU side as `.cpu()` calls).
Usage: python tools/bench_cmil.py [R] > cmil.json"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import heads, ops, synth  # noqa: E402
from oracle import ref_native  # noqa: E402  (CPU side of the comparison only)


def gpu_us(fn, iters=20):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def cpu_us(fn, iters=5):
    fn()
    ts = []
    for _ in range(iters):
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e6)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    r = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    k = 20
    dev = torch.device("cuda", 0)
    g = synth.generator(5)
    boxes = synth.proposals(g, r, 600, 800)
    C, D = torch.randn(r, k, generator=g), torch.randn(r, k, generator=g)
    S = torch.rand(r, 1, generator=g)
    bd, Cd, Dd, Sd = boxes.to(dev), C.to(dev), D.to(dev), S.to(dev)
    res = {"R": r, "K": k}
    J = ops.pairwise_iou(bd, bd)
    res["pairwise_iou_RxR_us"] = gpu_us(lambda: ops.pairwise_iou(bd, bd))
    res["roi_merge_fwd_us"] = gpu_us(lambda: ops.roi_merge_forward(Sd, J, Cd, Dd, 0.3))
    out = ops.roi_merge_forward(Sd, J, Cd, Dd, 0.3)
    res["roi_merge_bwd_us"] = gpu_us(lambda: ops.roi_merge_backward(out["MC"], out["MD"], out["I"], out["IC"]))
    L = torch.zeros(1, k)
    L[0, [3, 7, 11]] = 1
    CW = torch.rand(1, k, generator=g)
    Sk = torch.rand(r, k + 1, generator=g)
    Ld, CWd, Skd = L.to(dev), CW.to(dev), Sk.to(dev)
    perm = torch.randperm(r, device=dev)
    res["roi_label_us"] = gpu_us(lambda: ops.roi_label_forward(Skd, J, Ld, CWd, perm, 0.6, 0.4, 0.1, 32, 96, 1))
    centers = bd[:12].contiguous()
    res["pcl_proposal_clusters_us"] = gpu_us(lambda: ops.pcl_proposal_clusters(
        bd, centers, torch.arange(1, 13, device=dev).float(), torch.rand(12, device=dev), Skd))
    labels = [torch.tensor([3, 7, 11], device=dev)]
    img = torch.rand(1, k, device=dev)
    res["mist_labels_us"] = gpu_us(lambda: heads.refine_labels_mist(Skd, bd, bd, [r], labels, img, k, 0.15))
    res["top_k_thres_labels_us"] = gpu_us(lambda: heads.refine_labels_top_k(Skd, bd, bd, [r], labels, img, k, 0.1, 0.9))
    if ref_native.wsl_cmil_available():
        P8 = np.array([0, 1 << 30, 700, 20, 1280, 0, 0, 0], np.int32)
        P14 = np.array([0.6, 0.4, 0.1, 32, 96, 1, 0, 7, 1 << 20, 0, 0, 0, 0, 0], np.float32)

        def ref_merge():  # what fast_rcnn_wsddn.py:666 does: four device -> host copies, then the CPU op
            s, j, c, d = Sd.cpu(), J.cpu(), Cd.cpu(), Dd.cpu()
            ref_native.roi_merge_forward(s.numpy(), j.numpy(), c.numpy(), d.numpy(), P8)

        def ref_label():  # roi_heads_cmil.py:419-431
            s, u, l, cw = Skd.cpu(), J.cpu(), Ld.cpu(), CWd.cpu()
            ref_native.roi_label_forward(s.numpy(), u.numpy(), l.numpy(), cw.numpy(), P14)

        res["ref_roi_merge_cpu_incl_copies_us"] = cpu_us(ref_merge)
        res["ref_roi_label_cpu_incl_copies_us"] = cpu_us(ref_label)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
