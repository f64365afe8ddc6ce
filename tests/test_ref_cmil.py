"""Pin the restatements of ROIMerge / ROILabel (oracle/ref_cmil.py) to the reference's OWN C++ (oracle/_ref/
libref_wslcmil.so = projects/WSL/wsl/layers/csrc/ROIMerge/ROIMerge_cpu.cpp + ROILabel/ROILabel_cpu.cpp compiled in
place, oracle/Makefile.ref): cluster ids, counts, merged scores, gradients, labels and weights bit for bit."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_cmil, ref_heads, ref_native

pytestmark = pytest.mark.skipif(not ref_native.wsl_cmil_available(), reason="oracle/_ref not built and no /root/reference")


def _case(seed, r, k):
    g = synth.generator(seed)
    boxes = synth.proposals(g, r, 600, 800)
    J = ref_heads.pairwise_iou(boxes, boxes).numpy()
    C = torch.randn(r, k, generator=g).numpy()
    D = torch.randn(r, k, generator=g).numpy()
    S = torch.rand(r, generator=g).numpy()
    return boxes, J, C, D, S


@pytest.mark.parametrize("seed,r,k,cur_iter", [(1, 500, 20, 0), (2, 500, 20, 700), (3, 150, 7, 2000), (4, 40, 3, 5000),
                                               (5, 1200, 20, 12000), (6, 1, 4, 10)])
def test_roi_merge_restatement_equals_reference(seed, r, k, cur_iter):
    _, J, C, D, S = _case(seed, r, k)
    P = np.array([0, 1 << 30, cur_iter, 20, 1280, 0, 0, 0], np.int32)
    mc, md, i, ic, p2 = ref_native.roi_merge_forward(S, J, C, D, P)
    lam = ref_cmil.merge_lambda(cur_iter, 1280, 20)
    mine = ref_cmil.roi_merge_forward(S, J, C, D, lam)
    assert np.array_equal(i, mine["I"]) and np.array_equal(ic, mine["IC"])
    assert np.array_equal(mc, mine["MC"]) and np.array_equal(md, mine["MD"])
    assert p2[2] == cur_iter + 1
    assert (p2[5], p2[6], p2[7]) == (mine["num_top_id"], mine["max_clique"], mine["min_clique"])
    if r >= 150 and cur_iter < 5000:
        assert mine["max_clique"] > 1  # early in training lambda is small: boxes really merge
    g = np.random.RandomState(seed)
    gmc, gmd = g.randn(*mc.shape).astype(np.float32), g.randn(*md.shape).astype(np.float32)
    gc, gd = ref_native.roi_merge_backward(C, D, gmc, gmd, i, ic)
    mgc, mgd = ref_cmil.roi_merge_backward(gmc, gmd, i, ic)
    assert np.array_equal(gc, mgc) and np.array_equal(gd, mgd)


@pytest.mark.parametrize("seed,r,k,ks,top_k,with_cw,thr", [
    (1, 1500, 20, 20, 1, True, (0.6, 0.4, 0.1)), (2, 1500, 20, 21, 1, True, (0.3, 0.3, 0.0)),
    (3, 90, 6, 6, 2, True, (0.6, 0.4, 0.1)), (4, 700, 20, 20, 1, False, (0.2, 0.2, 0.05)),
    (5, 30, 4, 4, 1, True, (0.6, 0.4, 0.1)), (6, 900, 20, 20, 3, True, (0.1, 0.5, 0.0))])
def test_roi_label_restatement_equals_reference(seed, r, k, ks, top_k, with_cw, thr):
    g = synth.generator(100 + seed)
    boxes = synth.proposals(g, r, 600, 800)
    U = ref_heads.pairwise_iou(boxes, boxes).numpy()
    S = torch.rand(r, ks, generator=g).numpy()
    L = np.zeros(k, np.float32)
    L[torch.randperm(k, generator=g)[:3].numpy()] = 1
    CW = torch.rand(k, generator=g).numpy() if with_cw else None
    P = np.array([thr[0], thr[1], thr[2], 32, 96, top_k, 0, 7, 1 << 20, 0, 0, 0, 0, 0], np.float32)
    perm = ref_native.roi_label_perm(r)
    assert sorted(perm.tolist()) == list(range(r))
    rl, rw, p2 = ref_native.roi_label_forward(S, U, L, CW, P)
    mine = ref_cmil.roi_label_forward(S, U, L, CW, perm, thr[0], thr[1], thr[2], 32, 96, top_k)
    assert np.array_equal(rl, mine["RL"]) and np.array_equal(rw, mine["RW"])
    assert (int(p2[10]), int(p2[11])) == (mine["fg_rois"], mine["bg_rois"])
    assert p2[12] == mine["fg_weight"] and p2[13] == mine["bg_weight"]
    if thr[0] < 0.5:  # a cap binds (num_pos + 1 / num_neg + 1 accepted): the visiting order matters
        assert mine["bg_rois"] == 97 or mine["fg_rois"] == 33, (mine["fg_rois"], mine["bg_rois"])
