"""Pin the CPU restatement of ROILoopPool / MOIPool (oracle/jtsm_oracle.c) to the reference's OWN kernels:
oracle/_ref/libref_wslpool.so is projects/WSL/wsl/layers/csrc/{ROILoopPool,MOIPool}/*_cuda.cu compiled as serial
host code from /root/reference (oracle/Makefile.ref).  Bit-exact values and argmax, including the reference's
quirks (maxval = 0 initialisation, argmax -1 where nothing beats 0, zero-initialised argmax where no thread writes).
"""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_native

pytestmark = pytest.mark.skipif(not ref_native.available(), reason="oracle/_ref not built and no /root/reference")


def _case(seed, n, c, h, w, r, img=(600, 800), signed=False):
    g = synth.generator(seed)
    x = synth.feature_map(g, n, c, h, w)
    if signed:
        x = torch.randn(n, c, h, w, generator=g)
    boxes = [synth.proposals(g, r // n, *img) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    perm = torch.randperm(rois.shape[0], generator=g)
    return g, x.numpy(), rois[perm].numpy()


@pytest.mark.parametrize("seed,n,c,h,w,r,scale,pooled,signed", [
    (11, 1, 5, 38, 50, 120, 1 / 16.0, (7, 7), False),
    (12, 2, 3, 75, 100, 200, 1 / 8.0, (7, 7), False),
    (13, 2, 4, 19, 25, 90, 1 / 32.0, (14, 14), False),
    (14, 1, 3, 38, 50, 80, 1 / 16.0, (3, 5), False),
    (15, 1, 4, 38, 50, 100, 1 / 16.0, (7, 7), True),  # negative features: the maxval = 0 start shows
])
def test_roi_loop_pool_restatement_equals_reference_kernel(oracle, seed, n, c, h, w, r, scale, pooled, signed):
    _, x, rois = _case(seed, n, c, h, w, r, signed=signed)
    out, arg = oracle.roi_loop_pool_forward(x, rois, scale, pooled)
    rout, rarg = ref_native.roi_loop_pool_forward(x, rois, scale, pooled)
    assert np.array_equal(out, rout)
    assert np.array_equal(arg, rarg)
    go = np.random.RandomState(seed).randn(*out.shape).astype(np.float32)
    gi = oracle.max_pool_backward(go, arg, np.tile(rois, (3, 1)), x.shape)
    rgi = ref_native.roi_loop_pool_backward(go, rarg, rois, x.shape, scale)
    assert np.array_equal(gi, rgi)  # both serial, same visiting order


def test_roi_loop_pool_extreme_boxes_equal_reference_kernel(oracle):
    """Degenerate, out-of-image and whole-image boxes (context rectangle 1.8x leaves the map)."""
    x = synth.feature_map(synth.generator(3), 1, 3, 38, 50).numpy()
    rois = np.array([[0, 0, 0, 799, 599], [0, 100, 100, 100, 100], [0, -50, -40, 30, 20], [0, 700, 500, 900, 700],
                     [0, 10.4, 20.6, 11.2, 300.5], [0, 300, 10, 310, 590], [0, 790, 590, 799, 599]], np.float32)
    for scale in (1 / 16.0, 1 / 8.0):
        out, arg = oracle.roi_loop_pool_forward(x, rois, scale, (7, 7))
        rout, rarg = ref_native.roi_loop_pool_forward(x, rois, scale, (7, 7))
        assert np.array_equal(out, rout) and np.array_equal(arg, rarg)


@pytest.mark.parametrize("seed,n,c,h,w,r,scale,pooled,block,p", [
    (21, 1, 4, 38, 50, 60, 1 / 16.0, (7, 7), 24, 0.35),
    (22, 2, 3, 75, 100, 80, 1 / 8.0, (7, 7), 40, 0.5),
    (23, 1, 3, 38, 50, 40, 1 / 16.0, (14, 14), 16, 0.1),
    (24, 1, 2, 38, 50, 30, 1 / 16.0, (7, 7), 24, 0.0),  # nothing switched on
    (25, 1, 2, 38, 50, 30, 1 / 16.0, (7, 7), 24, 1.0),  # everything switched on
])
def test_moi_pool_restatement_equals_reference_kernel(oracle, seed, n, c, h, w, r, scale, pooled, block, p):
    g, x, rois = _case(seed, n, c, h, w, r)
    sp, s = synth.superpixels(g, n, 600 // 4, 800 // 4, block=block)
    oh = synth.oh_labels(g, rois.shape[0], s, p=p).numpy()
    out, arg = oracle.moi_pool_forward(x, rois, oh, sp.numpy(), scale, pooled)
    rout, rarg = ref_native.moi_pool_forward(x, rois, oh, sp.numpy(), scale, pooled)
    assert np.array_equal(out, rout)
    assert np.array_equal(arg, rarg)
    go = np.random.RandomState(seed).randn(*out.shape).astype(np.float32)
    gi = oracle.max_pool_backward(go, arg, rois, x.shape)
    rgi = ref_native.moi_pool_backward(go, rarg, rois, x.shape, scale)
    assert np.array_equal(gi, rgi)


def test_moi_pool_signed_features_equal_reference_kernel(oracle):
    g, x, rois = _case(31, 1, 3, 38, 50, 50, signed=True)
    sp, s = synth.superpixels(g, 1, 150, 200)
    oh = synth.oh_labels(g, rois.shape[0], s).numpy()
    out, arg = oracle.moi_pool_forward(x, rois, oh, sp.numpy(), 1 / 16.0, (7, 7))
    rout, rarg = ref_native.moi_pool_forward(x, rois, oh, sp.numpy(), 1 / 16.0, (7, 7))
    assert np.array_equal(out, rout) and np.array_equal(arg, rarg)


# ---- RoIAlign: the reference's own ROIAlign_cpu.cpp (oracle/_ref/libref_roialign.so) ----
_align = pytest.mark.skipif(not ref_native.roi_align_available(), reason="oracle/_ref/libref_roialign.so not built")


@_align
def test_reference_roi_align_golden_vectors():
    # reference: tests/layers/test_roi_align.py:12-45 -- run through the reference's own compiled CPU op
    inp = np.arange(25, dtype=np.float32).reshape(1, 1, 5, 5)
    rois = np.array([[0, 1, 1, 3, 3]], np.float32)
    old = [[7.5, 8, 8.5, 9], [10, 10.5, 11, 11.5], [12.5, 13, 13.5, 14], [15, 15.5, 16, 16.5]]
    new = [[4.5, 5.0, 5.5, 6.0], [7.0, 7.5, 8.0, 8.5], [9.5, 10.0, 10.5, 11.0], [12.0, 12.5, 13.0, 13.5]]
    assert np.allclose(ref_native.roi_align_forward(inp, rois, 1.0, (4, 4), 0, False)[0, 0], np.array(old, np.float32))
    assert np.allclose(ref_native.roi_align_forward(inp, rois, 1.0, (4, 4), 0, True)[0, 0], np.array(new, np.float32))


@_align
@pytest.mark.parametrize("aligned", [False, True])
@pytest.mark.parametrize("sampling_ratio", [0, 2])
@pytest.mark.parametrize("scale,pooled", [(1 / 16.0, (7, 7)), (1 / 8.0, (14, 14))])
def test_roi_align_restatement_equals_reference_cpu(oracle, aligned, sampling_ratio, scale, pooled):
    _, x, rois = _case(41, 2, 4, 38, 50, 120)
    out = oracle.roi_align_forward(x, rois, scale, pooled, sampling_ratio, aligned)
    rout = ref_native.roi_align_forward(x, rois, scale, pooled, sampling_ratio, aligned)
    assert np.array_equal(out, rout)
    go = np.random.RandomState(3).randn(*out.shape).astype(np.float32)
    gi = oracle.roi_align_backward(go, rois, scale, x.shape, sampling_ratio, aligned)
    rgi = ref_native.roi_align_backward(go, rois, scale, x.shape, sampling_ratio, aligned)
    assert np.array_equal(gi, rgi)  # both serial, same visiting order


# ---- PCL loss: the reference's own pcl_loss_cpu.cpp (oracle/_ref/libref_pclloss.so) ----
_pcl = pytest.mark.skipif(not ref_native.pcl_loss_available(), reason="oracle/_ref/libref_pclloss.so not built")


@_pcl
@pytest.mark.parametrize("r,k,p", [(300, 20, 6), (2000, 20, 12), (50, 80, 1), (7, 3, 4)])
def test_pcl_loss_restatement_equals_reference_cpu(oracle, r, k, p):
    c = synth.pcl_case(synth.generator(60 + r), r, k, p)
    n = {kk: v.numpy() for kk, v in c.items()}
    out, _ = oracle.pcl_loss_forward(n["probs"], n["labels"], n["cls_loss_weights"], n["pc_labels"], n["pc_probs"],
                                     n["img_cls_loss_weights"], n["im_labels"])
    rout = ref_native.pcl_loss_forward(n["probs"], n["labels"], n["cls_loss_weights"], n["pc_labels"], n["pc_probs"],
                                       n["img_cls_loss_weights"], n["im_labels"])
    assert np.array_equal(out, rout)
    g = oracle.pcl_loss_backward(n["probs"], n["labels"], n["cls_loss_weights"], n["gt_assignment"], n["pc_labels"],
                                 n["pc_probs"], n["pc_count"], n["img_cls_loss_weights"], n["im_labels"])
    rg = ref_native.pcl_loss_backward(n["probs"], n["labels"], n["cls_loss_weights"], n["gt_assignment"],
                                      n["pc_labels"], n["pc_probs"], n["pc_count"], n["img_cls_loss_weights"],
                                      n["im_labels"])
    assert np.array_equal(g, rg / np.float32(r))
    assert (out != 0).sum() >= 2 and (g != 0).sum() > 0
