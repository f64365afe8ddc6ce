"""GPU parity of the PCL targets (jtsm_b200/pcl.py, `jtsm_pcl_proposal_clusters`) against oracle/ref_pcl.py, which is
pinned to the reference's own third_party/pcl.py (tests/test_ref_pcl.py): assignments, labels, weights and counts
bit-exact; the two per-cluster float sums to rel 1e-5 (numpy's pairwise summation vs a fixed-order tree)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads, ref_pcl

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,r,k,g", [(1, 800, 20, 7), (2, 300, 5, 1), (3, 4000, 20, 15), (4, 2000, 80, 40)])
def test_pcl_proposal_clusters_vs_restatement(cuda, seed, r, k, g):
    from jtsm_b200 import ops

    gen = synth.generator(seed)
    rois = synth.proposals(gen, r, 600, 800)
    centers = rois[torch.randperm(r, generator=gen)[:g]].clone()
    labels = torch.randint(0, k, (g,), generator=gen) + 1
    scores = torch.rand(g, generator=gen)
    prob = torch.rand(r, k + 1, generator=gen)
    want = ref_pcl.proposal_clusters(rois.numpy(), centers.numpy(), labels.numpy().astype(np.int32), scores.numpy(),
                                     prob.numpy())
    got = ops.pcl_proposal_clusters(rois.to(cuda), centers.to(cuda), labels.to(cuda), scores.to(cuda), prob.to(cuda))
    for n in ("labels", "cls_loss_weights", "gt_assignment", "pc_labels", "pc_count"):
        assert got[n].dtype == torch.float32
        assert np.array_equal(got[n].cpu().numpy(), want[n].astype(np.float32)), n
    for n in ("pc_probs", "img_cls_loss_weights"):
        np.testing.assert_allclose(got[n].cpu().numpy(), want[n], rtol=1e-5, atol=1e-7, err_msg=n)


def test_pcl_targets_end_to_end_feed_the_loss(cuda):
    """PCL(): host centres (device IoU graph) + device clusters == host centres with the checker's IoU + restatement;
    the dict drives PCLLoss forward + backward."""
    from jtsm_b200 import pcl
    from jtsm_b200.layers import pcl_loss

    gen = synth.generator(31)
    r, k = 1500, 20
    rois = synth.proposals(gen, r, 600, 800)
    prob = torch.rand(r, k, generator=gen) ** 4
    prob_new = torch.softmax(torch.randn(r, k + 1, generator=gen), dim=1)
    im_labels = np.zeros((1, k), np.float32)
    im_labels[0, [3, 11]] = 1
    out = pcl.PCL(rois.numpy(), prob.to(cuda), im_labels, prob_new.to(cuda))
    cpu_iou = lambda b: ref_heads.pairwise_iou(torch.from_numpy(b), torch.from_numpy(b)).numpy()
    p = np.clip(prob.numpy(), 1e-9, 1 - 1e-9)
    cb, cc, cs = pcl.graph_centers(rois.numpy(), p, im_labels, iou_fn=cpu_iou)
    want = ref_pcl.proposal_clusters(rois.numpy(), cb, cc, cs, np.clip(prob_new.numpy(), 1e-9, 1 - 1e-9))
    assert out["labels"].shape == (1, r) and out["pc_labels"].shape == (1, len(cc))
    for n in ("labels", "cls_loss_weights", "gt_assignment", "pc_labels", "pc_count"):
        assert np.array_equal(out[n].cpu().numpy().reshape(-1), want[n].astype(np.float32)), n
    np.testing.assert_allclose(out["pc_probs"].cpu().numpy().reshape(-1), want["pc_probs"], rtol=1e-5)
    assert out["im_labels_real"].shape == (1, k + 1) and float(out["im_labels_real"][0, 0]) == 1.0
    x = prob_new.to(cuda).requires_grad_()
    loss = pcl_loss(x, out["labels"], out["cls_loss_weights"], out["gt_assignment"], out["pc_labels"], out["pc_probs"],
                    out["pc_count"], out["img_cls_loss_weights"], out["im_labels_real"])
    loss.backward()
    assert torch.isfinite(loss) and float(loss) > 0 and torch.isfinite(x.grad).all() and float(x.grad.abs().sum()) > 0
