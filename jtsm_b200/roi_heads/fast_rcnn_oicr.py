"""`OICROutputLayers`: one refinement branch's predictor
(projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py:436-783), same constructor arguments, methods and return values;
the loss / probability / decoding math runs in the sm_100a kernels:

  forward          cls_score (+ bbox_pred when refine_reg[refine_k]) `nn.Linear`            torch GEMMs
  losses           {"loss_cls_r<k>": sum_r w_r CE_r / #(w_r > 1e-12) [, "loss_box_reg_r<k>"]}   jtsm_refine_ce / _box_reg
                   (forward + d loss / d logits + softmax + accuracy counters in one launch)
  predict_probs    softmax rows (reuses the probabilities `losses` produced for the same logits)    jtsm_softmax_rows
  predict_boxes    Box2BoxTransform.apply_deltas                                                    jtsm_apply_deltas
  predict_*_K      means over the Kr branches (inference)
  inference        fast_rcnn_inference

`proposals` (list per image) must carry `proposal_boxes`; for `losses` also `gt_classes`, `gt_weights` and, with box
regression, `gt_boxes` -- what `label_and_sample_proposals` attaches (roi_heads.py:318-335).  When they come from
this package's fused label assignment they also carry `loss_weights` / `valid_count` (OICROutputs.__init__'s weight
preparation, fast_rcnn_oicr.py:243-247, done by the assign kernel); otherwise those are derived here."""
from typing import Dict, List, Union

import torch
from torch import nn

from .. import ops
from ..box_regression import Box2BoxTransform
from ..inference import fast_rcnn_inference
from . import functions as Fn
from .fast_rcnn_wsddn import _cat_boxes, _input_size


class OICROutputLayers(nn.Module):
    def __init__(self, input_shape, *, box2box_transform, num_classes: int, test_score_thresh: float = 0.0,
                 test_nms_thresh: float = 0.5, test_topk_per_image: int = 100, cls_agnostic_bbox_reg: bool = False,
                 smooth_l1_beta: float = 0.0, box_reg_loss_type: str = "smooth_l1",
                 loss_weight: Union[float, Dict[str, float]] = 1.0, mean_loss: bool = True, refine_k: int = None,
                 refine_reg: List[bool] = (False,)):
        super().__init__()
        self.num_classes = num_classes
        input_size = _input_size(input_shape)
        self.cls_score = nn.Linear(input_size, num_classes + 1)
        self.num_bbox_reg_classes = 1 if cls_agnostic_bbox_reg else num_classes
        self.box_dim = len(box2box_transform.weights)
        self.bbox_pred = nn.Linear(input_size, self.num_bbox_reg_classes * self.box_dim)
        nn.init.normal_(self.cls_score.weight, std=0.01)
        nn.init.normal_(self.bbox_pred.weight, std=0.001)
        for l in (self.cls_score, self.bbox_pred):
            nn.init.constant_(l.bias, 0)
        self.box2box_transform = box2box_transform
        self.smooth_l1_beta = smooth_l1_beta
        self.test_score_thresh = test_score_thresh
        self.test_nms_thresh = test_nms_thresh
        self.test_topk_per_image = test_topk_per_image
        self.box_reg_loss_type = box_reg_loss_type
        if isinstance(loss_weight, float):
            loss_weight = {"loss_cls": loss_weight, "loss_box_reg": loss_weight}
        self.loss_weight = loss_weight
        self.mean_loss = mean_loss
        self.refine_k = refine_k
        self.refine_reg = list(refine_reg)

    @classmethod
    def from_config(cls, cfg, input_shape, refine_k):
        """Same keys as the reference's from_config (fast_rcnn_oicr.py:514-532)."""
        return dict(input_shape=input_shape,
                    box2box_transform=Box2BoxTransform(weights=cfg.MODEL.ROI_BOX_HEAD.BBOX_REG_WEIGHTS),
                    num_classes=cfg.MODEL.ROI_HEADS.NUM_CLASSES,
                    cls_agnostic_bbox_reg=cfg.MODEL.ROI_BOX_HEAD.CLS_AGNOSTIC_BBOX_REG,
                    smooth_l1_beta=cfg.MODEL.ROI_BOX_HEAD.SMOOTH_L1_BETA,
                    test_score_thresh=cfg.MODEL.ROI_HEADS.SCORE_THRESH_TEST,
                    test_nms_thresh=cfg.MODEL.ROI_HEADS.NMS_THRESH_TEST,
                    test_topk_per_image=cfg.TEST.DETECTIONS_PER_IMAGE,
                    box_reg_loss_type=cfg.MODEL.ROI_BOX_HEAD.BBOX_REG_LOSS_TYPE,
                    loss_weight={"loss_box_reg": cfg.MODEL.ROI_BOX_HEAD.BBOX_REG_LOSS_WEIGHT},
                    mean_loss=cfg.WSL.MEAN_LOSS, refine_reg=cfg.WSL.REFINE_REG, refine_k=refine_k)

    @property
    def has_reg(self):
        return bool(self.refine_reg[self.refine_k])

    def forward(self, x):
        """-> (scores [R, K+1], proposal_deltas [R, 4 * num_bbox_reg_classes]; zeros without box regression)."""
        if x.dim() > 2:
            x = torch.flatten(x, start_dim=1)
        scores = self.cls_score(x)
        if self.has_reg:
            proposal_deltas = self.bbox_pred(x)
        else:
            proposal_deltas = torch.zeros(scores.shape[0], self.num_bbox_reg_classes * self.box_dim, dtype=scores.dtype,
                                          device=scores.device, requires_grad=False)
        return scores, proposal_deltas

    # ------------------------------------------------------------------ training
    @staticmethod
    def _targets(proposals):
        """(gt_classes, loss_weights, valid_count) of all images; fast_rcnn_oicr.py:243-247 when not precomputed."""
        gt = torch.cat([p.gt_classes for p in proposals], dim=0)
        if all(p.has("loss_weights") for p in proposals) and hasattr(proposals[0], "_valid_count"):
            return gt, torch.cat([p.loss_weights for p in proposals], dim=0), proposals[0]._valid_count
        w = torch.cat([p.gt_weights for p in proposals], dim=0).to(torch.float32)
        w = torch.where(gt == -1, torch.zeros_like(w), w)
        valid = (w > 1e-12).sum().to(torch.int32).reshape(1)
        return gt, w, valid

    def losses(self, predictions, proposals):
        scores, proposal_deltas = predictions
        suffix = "_r" + str(self.refine_k)
        if scores.shape[0] == 0:
            out = {"loss_cls" + suffix: 0.0 * scores.sum()}
            if self.has_reg:
                out["loss_box_reg" + suffix] = 0.0 * proposal_deltas.sum()
            return out
        gt, w, valid = self._targets(proposals)
        loss, probs, acc = Fn.refine_ce(scores, gt, w, valid)
        scores._jtsm_probs = probs      # predict_probs of the same logits: no second softmax launch
        scores._jtsm_acc = acc          # int32 [4]: #(argmax == gt), #fg, #(fg & correct), #(fg & predicted background)
        losses = {"loss_cls" + suffix: loss.reshape(())}
        if self.has_reg:
            if self.box_reg_loss_type not in ("smooth_l1", "smooth_l1_weighted"):
                raise NotImplementedError("box_reg_loss_type %r (giou needs fvcore)" % self.box_reg_loss_type)
            gt_boxes = torch.cat([(p.gt_boxes if p.has("gt_boxes") else p.proposal_boxes).tensor for p in proposals], dim=0)
            lw = w if self.box_reg_loss_type == "smooth_l1_weighted" else None
            lb = Fn.refine_box_reg(proposal_deltas, _cat_boxes(proposals), gt_boxes, gt, lw, self.num_classes,
                                   self.box2box_transform.weights, self.smooth_l1_beta)
            losses["loss_box_reg" + suffix] = lb.reshape(())
        return {k: v * self.loss_weight.get(k, 1.0) for k, v in losses.items()}

    def losses_pcl(self, predictions, proposals, last_scores, gt_classes_img_oh):
        """The PCL refinement loss (fast_rcnn_oicr.py:589-613 -> PCLOutputs.pcl_loss :917-937): proposal-cluster targets
        from the PREVIOUS stream's scores (`jtsm_b200.pcl.PCL`: centre selection on the host like the reference's
        sklearn / numpy code, proposal clusters on the device) and `pcl_loss` on this stream's softmax (background =
        column 0 in PCL's layout).  One image per call, as the reference asserts."""
        from .. import pcl
        from ..layers import pcl_loss

        scores, proposal_deltas = predictions
        assert len(proposals) == 1, "PCL: batch size shoud be equal to 1"
        if self.has_reg:
            raise NotImplementedError("PCL with WSL.REFINE_REG is not ported")
        probs = torch.softmax(scores, dim=-1)
        im_labels = gt_classes_img_oh.detach().cpu().numpy()
        t = pcl.PCL(proposals[0].proposal_boxes.tensor, last_scores.detach(), im_labels, probs.detach())
        loss = pcl_loss(probs, t["labels"], t["cls_loss_weights"], t["gt_assignment"], t["pc_labels"], t["pc_probs"],
                        t["pc_count"], t["img_cls_loss_weights"], t["im_labels_real"])
        scores._jtsm_probs = probs.detach()
        losses = {"loss_cls_r" + str(self.refine_k): loss}
        return {k: v * self.loss_weight.get(k, 1.0) for k, v in losses.items()}

    def accuracy_counters(self, predictions):
        """Device int32 [4] behind `_log_accuracy` (fast_rcnn_oicr.py:252-280), without its four host syncs; None
        before `losses` ran on these predictions."""
        return getattr(predictions[0], "_jtsm_acc", None)

    # ------------------------------------------------------------------ predictions
    def predict_probs(self, predictions, proposals):
        scores, _ = predictions
        probs = getattr(scores, "_jtsm_probs", None)
        if probs is None:
            probs = ops.softmax_rows(scores)
        return probs.split([len(p) for p in proposals], dim=0)

    def predict_boxes(self, predictions, proposals):
        if not len(proposals):
            return []
        _, proposal_deltas = predictions
        boxes = self.box2box_transform.apply_deltas(proposal_deltas.detach(), _cat_boxes(proposals))
        return boxes.split([len(p) for p in proposals])

    def predict_boxes_for_gt_classes(self, predictions, proposals):
        if not len(proposals):
            return []
        _, proposal_deltas = predictions
        pb = _cat_boxes(proposals)
        n, b = pb.shape
        boxes = self.box2box_transform.apply_deltas(proposal_deltas.detach(), pb)
        k = boxes.shape[1] // b
        if k > 1:
            gt = torch.cat([p.gt_classes for p in proposals], dim=0).clamp(0, k - 1)
            boxes = boxes.view(n, k, b)[torch.arange(n, dtype=torch.long, device=boxes.device), gt]
        return boxes.split([len(p) for p in proposals])

    def predict_probs_K(self, predictions, proposals):
        """Mean over the branches of softmax(scores_k) (fast_rcnn_oicr.py:764-783): Kr accumulating launches."""
        out = None
        for scores_k, _ in predictions:
            out = ops.softmax_rows(scores_k, out, scale=1.0 / len(predictions), accumulate=out is not None)
        return out.split([len(p) for p in proposals], dim=0)

    def predict_boxes_K(self, predictions, proposals):
        """apply_deltas(mean of the branches' deltas) (fast_rcnn_oicr.py:712-742)."""
        if not len(proposals):
            return []
        deltas = torch.zeros_like(predictions[0][1])
        for _, d in predictions:
            deltas += d
        deltas = deltas / len(predictions)
        boxes = self.box2box_transform.apply_deltas(deltas, _cat_boxes(proposals))
        return boxes.split([len(p) for p in proposals])

    def inference(self, predictions, proposals, pcl_bg=False):
        if isinstance(predictions[0], tuple):
            boxes = self.predict_boxes_K(predictions, proposals)
            scores = self.predict_probs_K(predictions, proposals)
        else:
            boxes = self.predict_boxes(predictions, proposals)
            scores = self.predict_probs(predictions, proposals)
        image_shapes = [x.image_size for x in proposals]
        if pcl_bg:
            scores = [torch.cat((s[:, 1:], s[:, :1]), 1) for s in scores]
        return fast_rcnn_inference(boxes, scores, image_shapes, self.test_score_thresh, self.test_nms_thresh,
                                   self.test_topk_per_image)
