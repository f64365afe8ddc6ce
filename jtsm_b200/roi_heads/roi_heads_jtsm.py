"""`WSDDNROIHeadsB200`, `ContextLocNetROIHeadsB200` and `JTSMROIHeadsB200`: the other registered WSL heads on the sm_100a path.

* WSDDNROIHeads (projects/WSL/wsl/modeling/roi_heads/roi_heads_wsddn.py:28-330): pooling x objectness -> box head ->
  `WSDDNOutputLayers` -> BCE; no refinement branches.  = `OICRROIHeadsB200` with `refine_K = 0`, kept as its own class so
  that `MODEL.ROI_HEADS.NAME` and `from_config` read like the reference's.
* ContextLocNetROIHeads (roi_heads_contextlocnet.py:284-335): ROILoopPool gives the RoI, frame and context pools of every
  proposal ([3R, C, P, P]); all three go through the box head and `WSDDNOutputLayers(..., context=True)` scores them as
  cls(RoI), det(frame) - det(context).
* JTSMROIHeads, BOX branch (roi_heads_jtsm.py:197-752, BASELINE.json configs[2]): the OICR chain over JTSM's pooler --
  MOIPool restricted by the proposals' superpixel sets (`oh_labels`) -- with the pooled features scaled by
  `mask_scale = P^2 / (#non-empty bins + 1)` (roi_heads_jtsm.py:607-620, recomputed there from the returned arg-max with
  four eager kernels) and by `objectness_logits + 1`; both factors are the epilogue of the one MOIPool launch here
  (`jtsm_moi_pool_forward(..., row_scale, mask_scale)`) and of its backward.  `forward` takes the reference's extra
  arguments (`gt_sem_seg`, `superpixels`) and keeps `cnt_superpixels` for `ops.pairwise_iou_sp`.
  Not ported: the segmentation branch (`_forward_mask`, GrabCut pseudo-masks on the CPU, `get_pgt_sem_seg`) -- outside the
  per-proposal hot path; requesting it raises."""
from typing import Dict, List, Optional

import torch

from ..structures import Instances
from .roi_heads_oicr import OICRROIHeadsB200, get_image_level_gt


class WSDDNROIHeadsB200(OICRROIHeadsB200):
    def __init__(self, *, num_classes, box_in_features, box_pooler, box_head, box_predictor, **kwargs):
        kwargs.pop("refine_K", None)
        kwargs.pop("box_refinery", None)
        super().__init__(num_classes=num_classes, box_in_features=box_in_features, box_pooler=box_pooler,
                         box_head=box_head, box_predictor=box_predictor, refine_K=0, refine_reg=[], box_refinery=[],
                         **kwargs)

    @classmethod
    def from_config(cls, cfg, input_shape):
        ret = OICRROIHeadsB200.from_config.__func__(cls, _NoRefine(cfg), input_shape)
        for k in ("refine_K", "refine_mist", "refine_reg", "box_refinery"):
            ret.pop(k, None)
        return ret


class ContextLocNetROIHeadsB200(WSDDNROIHeadsB200):
    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        assert self.box_pooler is None or self.box_pooler.pooler_type == "ROILoopPool", "ContextLocNet pools with ROILoopPool"

    def _forward_box(self, features, proposals):
        feats = [features[f] for f in self.box_in_features]
        x = self.box_pooler(feats, [p.proposal_boxes for p in proposals])  # [3R, C, P, P]: RoI | frame | context
        scale = torch.cat([p.objectness_logits + 1 for p in proposals], dim=0)
        x = x * torch.cat([scale, scale, scale], dim=0).view(-1, 1, 1, 1)
        parts = list(torch.chunk(self.box_head(x), 3, dim=0))
        if self.training:
            predictions = self.box_predictor(parts, proposals, context=True, gt_classes_img_oh=self.gt_classes_img_oh)
            return self.box_predictor.losses(predictions, proposals, self.gt_classes_img_oh)
        predictions = self.box_predictor(parts, proposals, context=True)
        pred_instances, _, all_scores, all_boxes = self.box_predictor.inference(predictions, proposals)
        return pred_instances, all_scores, all_boxes


class CMILROIHeadsB200(OICRROIHeadsB200):
    """CMILROIHeads (roi_heads_cmil.py:36-470), the part that differs from the OICR heads: after the usual pseudo-gt +
    IoU label assignment of a branch, `ROILabel` re-labels the proposals -- seeds by class score, capped foreground /
    background slots in a random visiting order, weights = the image score of the seed's class (:419-437).  The
    reference moves S, the R x R IoU matrix, the labels and the class weights to the host for its CPU-only op on every
    branch; here the IoU matrix comes from `jtsm_pairwise_iou` and the op runs on the device (`jtsm_roi_label_forward`).
    One image per call, like the reference (its op asserts a square IoU matrix).  `box_predictor` may be a
    `WSDDNOutputLayers(cmil=True)`; the un-merged scores (third element of its training 4-tuple) seed the branches."""

    def __init__(self, *, roi_label=None, **kwargs):
        super().__init__(**kwargs)
        self.roi_label = roi_label

    @torch.no_grad()
    def _label_proposals(self, prev_scores, prev_boxes, proposals, counts, suffix=""):
        out = super()._label_proposals(prev_scores, prev_boxes, proposals, counts, suffix)
        if self.roi_label is None:
            return out
        assert len(proposals) == 1, "ROILabel takes the proposals of ONE image (square IoU matrix)"
        from .. import ops

        boxes = proposals[0].proposal_boxes.tensor
        U = ops.pairwise_iou(boxes, boxes)
        RL, RW = self.roi_label(prev_scores, U, self.gt_classes_img_oh, self.pred_class_img_logits)
        q = Instances(proposals[0].image_size)
        for name, v in out[0].get_fields().items():
            if name not in ("gt_classes", "gt_weights", "loss_weights"):
                q.set(name, v)
        q.gt_classes = RL.to(torch.int64)  # :436-437; the losses derive their row weights from gt_weights again
        q.gt_weights = RW.to(torch.float32)
        return [q]


class PCLROIHeadsB200(OICRROIHeadsB200):
    """PCLROIHeads (roi_heads_pcl.py:300-368): the refinement streams are trained with the PCL loss on proposal clusters
    built from the previous stream's scores instead of OICR's top-1 pseudo labels; inference rotates the background
    column last (`pcl_bg=True`)."""

    def _forward_predictors(self, box_features, proposals):
        if not self.training:
            predictions_K = [self.box_refinery[k](box_features) for k in range(self.refine_K)]
            pred_instances, _, all_scores, all_boxes = self.box_refinery[-1].inference(predictions_K, proposals, pcl_bg=True)
            return pred_instances, all_scores, all_boxes
        predictions = self.box_predictor(box_features, proposals, gt_classes_img_oh=self.gt_classes_img_oh)
        losses = self.box_predictor.losses(predictions, proposals, self.gt_classes_img_oh)
        self.pred_class_img_logits = self.box_predictor.predict_probs_img(predictions, proposals).detach()
        prev_scores = predictions[0].detach()
        for k in range(self.refine_K):
            predictions_k = self.box_refinery[k](box_features)
            losses.update(self.box_refinery[k].losses_pcl(predictions_k, proposals, prev_scores, self.gt_classes_img_oh))
            prev_scores = self.box_refinery[k].predict_probs(predictions_k, proposals)[0].detach()
        return losses


class _NoRefine:
    """cfg view with WSL.REFINE_NUM = 0 (the WSDDN configs do not carry the refinement keys)."""

    def __init__(self, cfg):
        self._cfg = cfg

    def __getattr__(self, name):
        v = getattr(self._cfg, name)
        if name == "WSL":
            return _Over(v, REFINE_NUM=0, REFINE_MIST=False, REFINE_REG=[])
        return v


class _Over:
    def __init__(self, node, **over):
        self._node, self._over = node, over

    def __getattr__(self, name):
        if name in self._over:
            return self._over[name]
        return getattr(self._node, name)


class JTSMROIHeadsB200(OICRROIHeadsB200):
    def __init__(self, *, mask_on: bool = False, **kwargs):
        if mask_on:
            raise NotImplementedError("JTSM's segmentation branch (_forward_mask / GrabCut) is outside the ported hot path")
        super().__init__(**kwargs)
        assert self.box_pooler is None or self.box_pooler.pooler_type == "MOIPool", "JTSM pools with MOIPool"
        self.superpixels = None
        self.cnt_superpixels = None

    def forward(self, images, features: Dict[str, torch.Tensor], proposals: List[Instances],
                targets: Optional[List[Instances]] = None, gt_sem_seg=None, superpixels=None):
        """roi_heads_jtsm.py:502-561 (box branch).  `superpixels`: [N, Hs, Ws] integer label images (tensor or an
        ImageList-like with `.tensor`); every proposal container carries `oh_labels` [R_i, S]."""
        assert superpixels is not None, "JTSM heads need the superpixel label images"
        self.proposals = proposals
        self.superpixels = superpixels
        sp = superpixels.tensor if hasattr(superpixels, "tensor") else superpixels
        self.cnt_superpixels = [torch.bincount(s.reshape(-1).long(), minlength=p.oh_labels.size(1)).view(1, 1, -1)
                                for s, p in zip(sp, proposals)]
        return super().forward(images, features, proposals, targets)

    def _pool(self, features, proposals):
        """MOIPool x mask_scale x (objectness_logits + 1) (roi_heads_jtsm.py:606-625), one launch."""
        feats = [features[f] for f in self.box_in_features]
        scale = torch.cat([x.objectness_logits + 1 for x in proposals], dim=0)
        out = self.box_pooler(feats, [x.proposal_boxes for x in proposals],
                              oh_labels_list=[x.oh_labels for x in proposals], superpixels=self.superpixels,
                              row_scale=scale, mask_scale=True)
        box_features = out[0] if isinstance(out, tuple) else out
        return box_features, scale


def register_all(registry):
    """Register the three heads in a detectron2 `ROI_HEADS_REGISTRY`."""
    for c in (OICRROIHeadsB200, WSDDNROIHeadsB200, ContextLocNetROIHeadsB200, CMILROIHeadsB200, PCLROIHeadsB200,
              JTSMROIHeadsB200):
        registry.register(c)
    return registry


__all__ = ["WSDDNROIHeadsB200", "ContextLocNetROIHeadsB200", "CMILROIHeadsB200", "PCLROIHeadsB200", "JTSMROIHeadsB200", "register_all", "get_image_level_gt"]
