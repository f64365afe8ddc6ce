"""CPU timing legs of bench.py (TEST / BASELINE INFRASTRUCTURE -- never imported by jtsm_b200/).

kind="port":      oracle/jtsm_oracle.c (plain C restatement), fanned out over host threads by RoI range.
kind="reference": what the reference itself executes with MODEL.DEVICE=cpu for this path:
                  torchvision CPU roi_pool fwd+bwd (projects/WSL/wsl/modeling/poolers.py:183-186) and the
                  eager-PyTorch head composition (oracle/ref_heads.py restates it; no CPU op exists).
Both run a BOUNDED sample: the first `sample_rois` proposals of the workload's first image.
"""
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from jtsm_b200 import synth

from . import capi, ref_heads


def auto_sample_rois(wl, kind, budget_s=15.0):
    cfg = wl["cfg"]
    # measured here: ~0.6 ms / RoI / 512 channels single-threaded for pool fwd+bwd at stride 8
    per_roi = 1.2e-3 * cfg["C"] / 512.0
    n = int(budget_s / per_roi)
    return max(64, min(n, cfg["R"]))


def _sample(wl, sample_rois):
    rois = wl["rois"]
    first = rois[rois[:, 0] == 0][:sample_rois].contiguous()
    return first


def time_step(wl, sample_rois, threads, kind):
    cfg = wl["cfg"]
    rois = _sample(wl, sample_rois)
    r = rois.shape[0]
    x = wl["feat"][:1].float().contiguous()
    P = (cfg["P"], cfg["P"])
    g = synth.expand_grad_tile(wl["grad_tile"], r)
    scale = (wl["objectness"][:r] + 1.0)
    desc = "first %d of %d proposals of image 0 of %s (pool fwd+bwd, MIL, %d refine branches, all gradients), %d threads" % (
        r, cfg["R"], wl["name"], cfg["Kr"], threads)
    torch.set_num_threads(threads)
    head_in = dict(mil_cls=wl["mil_cls"][:r], mil_det=wl["mil_det"][:r], ref_logits=[x[:r] for x in wl["ref_logits"]],
                   ref_deltas=[x[:r] for x in wl["ref_deltas"]], proposal_boxes=[rois[:, 1:].contiguous()],
                   labels=wl["labels"][:1], num_classes=cfg["K"], reg=cfg["reg"])
    if cfg.get("pooler", "ROIPool") == "MOIPool":
        return _time_step_moi(wl, cfg, rois, x, P, g, scale, head_in, threads, kind, desc)
    if kind == "reference":
        import torchvision

        torch.set_num_threads(threads)
        t0 = time.perf_counter()
        xg = x.clone().requires_grad_()
        out = torchvision.ops.roi_pool(xg, rois, P, cfg["scale"]) * scale.view(-1, 1, 1, 1)
        out.backward(g)
        ref_heads.head_step(**head_in)
        t = time.perf_counter() - t0
        return t, r, desc + ", torchvision CPU roi_pool + eager PyTorch heads (the reference's MODEL.DEVICE=cpu path)"
    xn, rn, gn = x.numpy(), rois.numpy(), (g * scale.view(-1, 1, 1, 1)).numpy()
    capi.lib()
    chunks = [(i * r // threads, (i + 1) * r // threads) for i in range(threads)]
    t0 = time.perf_counter()

    def work(rng):
        out, arg = capi.roi_pool_forward(xn, rn, cfg["scale"], P, roi_range=rng)
        return capi.max_pool_backward(gn[rng[0]:rng[1]], arg[rng[0]:rng[1]], rn[rng[0]:rng[1]], tuple(x.shape))

    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, chunks))
    gi = np.sum(parts, axis=0)
    ref_heads.head_step(**head_in)
    t = time.perf_counter() - t0
    return t, r, desc + ", oracle C port (pool) + eager restatement (heads)"


def _time_step_moi(wl, cfg, rois, x, P, g, scale, head_in, threads, kind, desc):
    """MOIPool configurations (JTSM).  The reference has no CPU entry point for MOIPool (MOIPool.h:74 raises), so
    kind="reference" runs the reference's own kernels compiled as serial host code (oracle/_ref/libref_wslpool.so,
    oracle/Makefile.ref) when that library is present, else the C port; both fanned out over host threads by RoI
    range.  The per-RoI mask_scale factor (a scalar multiply per RoI) is left out of the CPU legs."""
    from . import ref_native

    r = rois.shape[0]
    xn, rn, gn = x.numpy(), rois.numpy(), (g * scale.view(-1, 1, 1, 1)).numpy()
    oh = wl["oh_labels"][:r].numpy()
    sp = wl["superpixels"][:1].numpy()
    use_ref = kind == "reference" and ref_native.available()
    fwd = ref_native.moi_pool_forward if use_ref else capi.moi_pool_forward
    capi.lib()
    chunks = [(i * r // threads, (i + 1) * r // threads) for i in range(threads)]
    t0 = time.perf_counter()

    def work(rng):
        a, b = rng
        if a == b:
            return np.zeros(x.shape, np.float32)
        out, arg = fwd(xn, rn[a:b], oh[a:b], sp, cfg["scale"], P)
        return capi.max_pool_backward(gn[a:b], arg, rn[a:b], tuple(x.shape))

    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, chunks))
    np.sum(parts, axis=0)
    ref_heads.head_step(**head_in)
    t = time.perf_counter() - t0
    how = "the reference's own MOIPool kernels built as serial host code (oracle/_ref)" if use_ref else "oracle C port of MOIPool"
    return t, r, desc + ", " + how + " + eager PyTorch heads"
