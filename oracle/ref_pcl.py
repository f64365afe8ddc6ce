"""CPU restatement of PCL's `_get_proposal_clusters` (projects/WSL/wsl/modeling/roi_heads/third_party/pcl.py:150-200).

TEST INFRASTRUCTURE ONLY.  Pinned to the reference's own function (loaded in place by `reference_module()`, needs
/root/reference) in tests/test_ref_pcl.py."""
import os

import numpy as np

from . import ref_heads, ref_loader

import torch


def reference_module():
    """The reference's pcl.py executed in place (sklearn present in this image; detectron2.structures stubbed by
    ref_loader.install())."""
    ref_loader.install()
    return ref_loader._load_file("ref_third_party_pcl", os.path.join(ref_loader.WSL, "modeling", "roi_heads",
                                                                     "third_party", "pcl.py"))


def proposal_clusters(rois, centers, center_labels, center_scores, cls_prob, fg_thresh=0.5, bg_thresh=0.1):
    """-> dict(labels int32 [R], cls_loss_weights [R], gt_assignment int64 [R], pc_labels int32 [G], pc_probs [G],
    pc_count int32 [G], img_cls_loss_weights [G]); sums in numpy's order (pairwise float32 summation)."""
    iou = ref_heads.pairwise_iou(torch.as_tensor(rois, dtype=torch.float32),
                                 torch.as_tensor(centers, dtype=torch.float32)).numpy()
    assign = iou.argmax(axis=1)
    best = iou.max(axis=1)
    labels = np.asarray(center_labels)[assign].copy()
    w = np.asarray(center_scores, np.float32)[assign].copy()
    w[best < bg_thresh] = 0.0
    labels[best < fg_thresh] = 0
    assign = assign.copy()
    assign[best < fg_thresh] = -1
    g = len(center_labels)
    icw, pp = np.zeros(g, np.float32), np.zeros(g, np.float32)
    pl, pc = np.zeros(g, np.int32), np.zeros(g, np.int32)
    for i in range(g):
        m = np.where(assign == i)[0]
        icw[i] = np.sum(w[m])
        pl[i] = center_labels[i]
        pc[i] = len(m)
        pp[i] = np.average(cls_prob[m, pl[i]]) if len(m) else np.nan
    return dict(labels=labels, cls_loss_weights=w, gt_assignment=assign, pc_labels=pl, pc_probs=pp, pc_count=pc,
                img_cls_loss_weights=icw)
