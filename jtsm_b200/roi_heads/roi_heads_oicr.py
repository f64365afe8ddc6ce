"""`OICRROIHeadsB200`: the WSDDN + OICR ROI heads (projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:36-468; with
refine_K = 0 it is the plain WSDDN heads of roi_heads_wsddn.py:28-330) on the sm_100a path.

Same members (`box_pooler`, `box_head`, `box_predictor`, `box_refinery_<k>`), same `forward(images, features,
proposals, targets)` contract and return values, so a detectron2 meta-architecture (GeneralizedRCNNWSL) drives it
unchanged and a reference checkpoint loads key for key.  What changes is what runs between the Linear layers:

  reference `_forward_box` (roi_heads_oicr.py:373-468)                   here
  ---------------------------------------------------------------------  -----------------------------------------------
  box_pooler, then box_features * (objectness_logits + 1)                one pooling launch, factor fused (fwd and bwd)
  box_predictor: ~6 eager kernels per image + BCE                        jtsm_wsddn_forward (scores, image scores, BCE)
  per branch: get_pgt_top_k + label_and_sample_proposals                 jtsm_refine_assign: one launch, no host sync
     (~40 eager kernels, 3-7 `.item()` syncs per image)
  OICROutputs: weighted CE (+ box reg), _log_accuracy (4 syncs)          jtsm_refine_ce (+ jtsm_refine_box_reg)
  predict_probs / predict_boxes for the next branch                      probabilities from the CE launch; jtsm_apply_deltas
  backward through all of it                                             autograd.Functions (functions.py): 1 launch each

Logging: the reference writes fg / bg counts and accuracies to the event storage with a host sync each
(roi_heads.py:359-361, fast_rcnn_oicr.py:252-280); here the same numbers stay on the device in `self.last_counters`
(read them when a log line is due).  WSL.REFINE_MIST targets (`get_pgt_mist`) go through heads.py (that variant reads
seed counts on the host, as the reference does).  Not ported: mask / keypoint heads, CMIL, sampling
(`WSL.SAMPLING.SAMPLING_ON`), `train_on_pred_boxes`."""
from typing import Dict, List, Optional

import torch
from torch import nn

from .. import heads as head_math
from .. import ops
from ..poolers import ROIPooler
from ..structures import Boxes, Instances
from .box_head import DiscriminativeAdaptionNeck
from .fast_rcnn_oicr import OICROutputLayers
from .fast_rcnn_wsddn import WSDDNOutputLayers


@torch.no_grad()
def get_image_level_gt(targets, num_classes):
    """Image-level class lists and one-hot labels from the per-image targets (roi_heads.py:145-161)."""
    if targets is None:
        return None, None, None
    gt_classes_img = [torch.unique(t.gt_classes, sorted=True) for t in targets]
    gt_classes_img_int = [gt.to(torch.int64) for gt in gt_classes_img]
    oh = torch.zeros((len(targets), num_classes), dtype=torch.float, device=gt_classes_img[0].device)
    for i, gt in enumerate(gt_classes_img_int):
        oh[i, gt] = 1.0
    return gt_classes_img, gt_classes_img_int, oh


class OICRROIHeadsB200(nn.Module):
    def __init__(self, *, num_classes: int, box_in_features: List[str], box_pooler: ROIPooler, box_head: nn.Module,
                 box_predictor: nn.Module, refine_K: int = 3, refine_mist: bool = False,
                 refine_reg: List[bool] = (False, False, False), box_refinery: List[nn.Module] = (),
                 cls_agnostic_bbox_reg: bool = False, iou_thresholds=(0.5,), iou_labels=(0, 1),
                 train_on_pred_boxes: bool = False, fuse_objectness: bool = True, mist_top_pro: float = 0.15, **kwargs):
        super().__init__()
        if train_on_pred_boxes:
            raise NotImplementedError("MODEL.ROI_BOX_HEAD.TRAIN_ON_PRED_BOXES is not used by the WSL configs")
        self.num_classes = num_classes
        self.in_features = self.box_in_features = box_in_features
        self.box_pooler = box_pooler
        self.box_head = box_head
        self.box_predictor = box_predictor
        self.refine_K = refine_K
        self.refine_mist = refine_mist
        self.refine_reg = list(refine_reg)
        self.box_refinery = list(box_refinery)
        assert len(self.box_refinery) == refine_K
        for k in range(refine_K):
            self.add_module("box_refinery_{}".format(k), self.box_refinery[k])
        self.cls_agnostic_bbox_reg = cls_agnostic_bbox_reg
        self.iou_thresholds = tuple(iou_thresholds)
        self.iou_labels = tuple(iou_labels)
        self.fuse_objectness = fuse_objectness
        self.mist_top_pro = mist_top_pro
        self.mask_on = self.keypoint_on = False
        self.iter = 0
        self.iter_test = 0
        self.last_counters = {}

    # ------------------------------------------------------------------ construction from a detectron2-style cfg
    @classmethod
    def from_config(cls, cfg, input_shape):
        """Mirror of OICRROIHeads.from_config / _init_box_head (roi_heads_oicr.py:140-232); `cfg` is read by attribute."""
        in_features = cfg.MODEL.ROI_HEADS.IN_FEATURES
        res = cfg.MODEL.ROI_BOX_HEAD.POOLER_RESOLUTION
        scales = tuple(1.0 / input_shape[k].stride for k in in_features)
        in_channels = [input_shape[f].channels for f in in_features]
        assert len(set(in_channels)) == 1, in_channels
        pooler = ROIPooler(output_size=res, scales=scales, sampling_ratio=cfg.MODEL.ROI_BOX_HEAD.POOLER_SAMPLING_RATIO,
                           pooler_type=cfg.MODEL.ROI_BOX_HEAD.POOLER_TYPE)
        shape = type("ShapeSpec", (), dict(channels=in_channels[0], height=res, width=res, stride=None))()
        box_head = DiscriminativeAdaptionNeck(shape, fc_dims=tuple(cfg.MODEL.ROI_BOX_HEAD.DAN_DIM))
        predictor = WSDDNOutputLayers(**WSDDNOutputLayers.from_config(cfg, box_head.output_shape))
        k_r = cfg.WSL.REFINE_NUM
        refinery = [OICROutputLayers(**OICROutputLayers.from_config(cfg, box_head.output_shape, k)) for k in range(k_r)]
        return dict(num_classes=cfg.MODEL.ROI_HEADS.NUM_CLASSES, box_in_features=in_features, box_pooler=pooler,
                    box_head=box_head, box_predictor=predictor, refine_K=k_r, refine_mist=cfg.WSL.REFINE_MIST,
                    refine_reg=cfg.WSL.REFINE_REG, box_refinery=refinery,
                    cls_agnostic_bbox_reg=cfg.MODEL.ROI_BOX_HEAD.CLS_AGNOSTIC_BBOX_REG,
                    iou_thresholds=cfg.MODEL.ROI_HEADS.IOU_THRESHOLDS, iou_labels=cfg.MODEL.ROI_HEADS.IOU_LABELS)

    # ------------------------------------------------------------------ forward
    def forward(self, images, features: Dict[str, torch.Tensor], proposals: List[Instances],
                targets: Optional[List[Instances]] = None, image_level_gt=None):
        """Training: (proposals, losses).  Inference: (pred_instances, {}, all_scores, all_boxes)  (roi_heads_oicr.py:302-345).
        image_level_gt (extension): a ready `(gt_classes_img, gt_classes_img_int, gt_classes_img_oh)` triple -- what
        `get_image_level_gt(targets)` returns, whose `torch.unique` synchronises -- so that a whole training step can be
        captured into a CUDA graph (trainstep.TrainStep.capture); gt_classes_img_int may be the device-side class lists of
        `ops.image_class_lists`."""
        if image_level_gt is not None:
            self.gt_classes_img, self.gt_classes_img_int, self.gt_classes_img_oh = image_level_gt
        else:
            self.gt_classes_img, self.gt_classes_img_int, self.gt_classes_img_oh = get_image_level_gt(targets,
                                                                                                     self.num_classes)
        self.images = images
        if self.training:
            assert targets or image_level_gt is not None, "'targets' argument is required during training"
            del targets
            losses = self._forward_box(features, proposals)
            self.iter += 1
            return proposals, losses
        pred_instances, all_scores, all_boxes = self._forward_box(features, proposals)
        self.iter_test += 1
        return pred_instances, {}, all_scores, all_boxes

    def _pool(self, features, proposals):
        """box_pooler + `box_features * (objectness_logits + 1)` (roi_heads_oicr.py:391-394), the factor fused into the
        pooling kernel and its backward where the pooler supports it (RoIPool, single level)."""
        feats = [features[f] for f in self.box_in_features]
        boxes = [x.proposal_boxes for x in proposals]
        scale = torch.cat([x.objectness_logits + 1 for x in proposals], dim=0)
        fused = self.fuse_objectness and len(feats) == 1 and self.box_pooler.pooler_type == "ROIPool"
        if fused:
            return self.box_pooler(feats, boxes, row_scale=scale), scale
        return self.box_pooler(feats, boxes) * scale.view(-1, 1, 1, 1), scale

    # ------------------------------------------------------------------ pseudo-gt + label assignment, one launch
    @torch.no_grad()
    def _label_proposals(self, prev_scores, prev_boxes, proposals, counts, suffix=""):
        """get_pgt_top_k (roi_heads_oicr.py:675-811; get_pgt_mist :549-591 with WSL.REFINE_MIST) followed by
        label_and_sample_proposals (roi_heads.py:263-370): per image the pseudo-gt seeds, IoU matching, classes /
        boxes / scores / weights of every proposal.  -> list of per-image Instances like the reference's, plus
        `loss_weights` (weights with ignored rows zeroed) and the shared `_valid_count`."""
        pb = torch.cat([p.proposal_boxes.tensor for p in proposals], dim=0)
        if self.refine_mist:
            a = head_math.refine_labels_mist(prev_scores, prev_boxes, pb, counts, self.gt_classes_img_int,
                                             self.pred_class_img_logits, self.num_classes, self.mist_top_pro,
                                             self.iou_thresholds, self.iou_labels)
        else:
            a = ops.refine_assign(prev_scores, prev_boxes, pb, counts, self.gt_classes_img_int,
                                  self.pred_class_img_logits, self.num_classes, self.iou_thresholds, self.iou_labels)
        self.last_counters["assign" + suffix] = a["counters"]  # int32 [3N + 1]: (fg, bg, ignore) per image, #valid
        out, off = [], 0
        for p, n in zip(proposals, counts):
            q = Instances(p.image_size)
            sl = slice(off, off + n)
            q.proposal_boxes = p.proposal_boxes
            q.objectness_logits = p.objectness_logits
            q.gt_classes = a["gt_classes"][sl]
            q.gt_boxes = Boxes(a["gt_boxes"][sl])
            q.gt_scores = a["gt_scores"][sl]
            q.gt_weights = a["gt_weights"][sl]
            q.loss_weights = a["loss_weights"][sl]
            q._valid_count = a["valid_count"]
            out.append(q)
            off += n
        return out

    def _forward_box(self, features, proposals):
        box_features, _ = self._pool(features, proposals)
        box_features = self.box_head(box_features)
        return self._forward_predictors(box_features, proposals)

    def _forward_predictors(self, box_features, proposals):
        """Everything of `_forward_box` after the box head (roi_heads_oicr.py:404-468) on given fc7 activations."""
        counts = [len(p) for p in proposals]
        if not self.training:
            if self.refine_K == 0:
                predictions = self.box_predictor(box_features, proposals)
                pred_instances, _, all_scores, all_boxes = self.box_predictor.inference(predictions, proposals)
                return pred_instances, all_scores, all_boxes
            predictions_K = [self.box_refinery[k](box_features) for k in range(self.refine_K)]
            pred_instances, _, all_scores, all_boxes = self.box_refinery[-1].inference(predictions_K, proposals)
            return pred_instances, all_scores, all_boxes

        predictions = self.box_predictor(box_features, proposals, gt_classes_img_oh=self.gt_classes_img_oh)
        losses = self.box_predictor.losses(predictions, proposals, self.gt_classes_img_oh)
        self.pred_class_img_logits = self.box_predictor.predict_probs_img(predictions, proposals).detach()
        prev_scores = predictions[0].detach()
        prev_boxes = torch.cat([p.proposal_boxes.tensor for p in proposals], dim=0)
        for k in range(self.refine_K):
            suffix = "_r" + str(k)
            term_weight = 3 if (self.refine_mist and k == 0) else 1
            proposals_k = self._label_proposals(prev_scores, prev_boxes, proposals, counts, suffix)
            predictions_k = self.box_refinery[k](box_features)
            losses_k = self.box_refinery[k].losses(predictions_k, proposals_k)
            if term_weight != 1:
                losses_k = {n: v * term_weight for n, v in losses_k.items()}
            self.last_counters["accuracy" + suffix] = self.box_refinery[k].accuracy_counters(predictions_k)
            prev_scores = torch.cat(self.box_refinery[k].predict_probs(predictions_k, proposals_k), dim=0).detach()
            prev_boxes = torch.cat(self.box_refinery[k].predict_boxes(predictions_k, proposals_k), dim=0).detach()
            losses.update(losses_k)
        return losses


def register(registry):
    """Register the heads in a detectron2 `ROI_HEADS_REGISTRY` (MODEL.ROI_HEADS.NAME: "OICRROIHeadsB200")."""
    registry.register(OICRROIHeadsB200)
    return OICRROIHeadsB200
