"""Pin oracle/ref_pcl.py (`_get_proposal_clusters`) to the reference's own third_party/pcl.py executed in place."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_loader, ref_pcl

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="needs /root/reference")


@pytest.mark.parametrize("seed,r,k,g", [(1, 800, 20, 7), (2, 300, 5, 1), (3, 2000, 20, 12)])
def test_proposal_clusters_restatement_equals_reference(seed, r, k, g):
    ref = ref_pcl.reference_module()
    gen = synth.generator(seed)
    rois = synth.proposals(gen, r, 600, 800).numpy()
    centers = rois[torch.randperm(r, generator=gen)[:g].numpy()].copy()
    labels = (torch.randint(0, k, (g,), generator=gen).numpy() + 1).astype(np.int32)
    scores = torch.rand(g, generator=gen).numpy()
    prob = torch.rand(r, k + 1, generator=gen).numpy()
    props = {"gt_boxes": centers, "gt_classes": labels.reshape(-1, 1), "gt_scores": scores.reshape(-1, 1)}
    want = ref._get_proposal_clusters(rois.copy(), props, np.zeros((1, k), np.float32), prob.copy())
    mine = ref_pcl.proposal_clusters(rois, centers, labels, scores, prob)
    names = ("labels", "cls_loss_weights", "gt_assignment", "pc_labels", "pc_probs", "pc_count", "img_cls_loss_weights")
    for n, w in zip(names, want):
        assert np.array_equal(np.asarray(w), mine[n], equal_nan=True), n
    assert (mine["gt_assignment"] >= 0).sum() >= g  # every centre is its own member


@pytest.mark.parametrize("seed,r,k", [(11, 600, 20), (12, 2000, 20), (13, 40, 4)])
def test_graph_centers_host_half_equals_reference(seed, r, k):
    """jtsm_b200.pcl.graph_centers (host numpy + sklearn, with the oracle's CPU IoU injected for the device kernel) vs
    the reference's `_get_graph_centers` run in place: same centres, classes and scores."""
    from jtsm_b200 import pcl
    from oracle import ref_heads

    ref = ref_pcl.reference_module()
    gen = synth.generator(seed)
    rois = synth.proposals(gen, r, 600, 800).numpy()
    prob = np.clip(torch.rand(r, k, generator=gen).numpy() ** 4, 1e-9, 1 - 1e-9)
    labels = np.zeros((1, k), np.float32)
    labels[0, torch.randperm(k, generator=gen)[:3].numpy()] = 1
    want = ref._get_graph_centers(rois.copy(), prob.copy(), labels.copy())
    cpu_iou = lambda b: ref_heads.pairwise_iou(torch.from_numpy(b), torch.from_numpy(b)).numpy()
    cb, cc, cs = pcl.graph_centers(rois, prob, labels, iou_fn=cpu_iou)
    assert np.array_equal(cb, want["gt_boxes"])
    assert np.array_equal(cc, want["gt_classes"].reshape(-1))
    assert np.array_equal(cs, want["gt_scores"].reshape(-1).astype(np.float32))
    assert len(cc) >= 3
