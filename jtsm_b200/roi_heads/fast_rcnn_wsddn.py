"""`WSDDNOutputLayers`: the reference's two-stream MIL predictor
(projects/WSL/wsl/modeling/roi_heads/fast_rcnn_wsddn.py:436-852) with the same constructor arguments, methods and
return values, the math behind them on the sm_100a kernels:

  forward            cls / det `nn.Linear` (torch GEMMs) -> ONE launch: per-image softmax(C, 1) * softmax(D, 0), the
                     proposal sums, their clamp and (when the image labels are already known) the BCE loss
  losses             `{"loss_cls": BCE(clamp(sum_r scores), one-hot)}` x loss_weight (mean, or sum / N when not mean_loss)
  predict_probs_img  clamped image scores [N, K]
  predict_probs      scores with a zero background column appended, split per image
  predict_boxes      apply_deltas(zeros) = the proposals' (centre, size) round trip, split per image
  inference          fast_rcnn_inference on the above

`proposals` is a list of per-image containers with `len()`, `.proposal_boxes.tensor` and `.image_size` (detectron2
`Instances` or jtsm_b200.structures.Instances).  With `cmil=True` (WSL.CMIL) the training forward runs `forward_cmil`
(:621-667): objectness = row sums of the scores, pairwise IoU of the proposals and the ROIMerge clique merge of the C / D
rows, all on the device (the reference copies S, the R x R IoU matrix, C and D to the host for its CPU-only op), and
returns the 4-tuple of the reference.  `context=True` is the ContextLocNet variant (`forward_contextlocnet`, :600-619):
class stream on the RoI features, detection stream = det(frame) - det(context) of ROILoopPool's three pools."""
from typing import Dict, Union

import torch
from torch import nn
from torch.nn import functional as F

from ..box_regression import Box2BoxTransform
from .. import ops
from ..inference import fast_rcnn_inference
from ..layers.roi_merge import ROIMerge
from . import functions as Fn


def _input_size(input_shape):
    if isinstance(input_shape, int):
        return input_shape
    return input_shape.channels * (getattr(input_shape, "width", None) or 1) * (getattr(input_shape, "height", None) or 1)


def _counts(proposals, total):
    return [total] if proposals is None else [len(p) for p in proposals]


def _cat_boxes(proposals):
    return torch.cat([p.proposal_boxes.tensor for p in proposals], dim=0)


class WSDDNOutputLayers(nn.Module):
    def __init__(self, input_shape, *, box2box_transform, num_classes: int, test_score_thresh: float = 0.0,
                 test_nms_thresh: float = 0.5, test_topk_per_image: int = 100, cls_agnostic_bbox_reg: bool = False,
                 smooth_l1_beta: float = 0.0, box_reg_loss_type: str = "smooth_l1",
                 loss_weight: Union[float, Dict[str, float]] = 1.0, mean_loss: bool = True, cmil: bool = False,
                 display: int = 1280, max_epoch: int = 40, size_epoch: int = 5000):
        super().__init__()
        self.num_classes = num_classes
        input_size = _input_size(input_shape)
        self.num_bbox_reg_classes = 1 if cls_agnostic_bbox_reg else num_classes
        self.box_dim = len(box2box_transform.weights)
        self.cls = nn.Linear(input_size, num_classes)
        self.det = nn.Linear(input_size, num_classes)
        nn.init.xavier_uniform_(self.cls.weight)
        nn.init.xavier_uniform_(self.det.weight)
        for l in (self.cls, self.det):
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
        self.cmil = cmil
        if cmil:
            self.display, self.max_epoch, self.size_epoch = display, max_epoch, size_epoch
            self.roi_merge = ROIMerge(0, display, max_epoch, size_epoch)

    @classmethod
    def from_config(cls, cfg, input_shape):
        """Same keys as the reference's from_config (fast_rcnn_wsddn.py:526-544)."""
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
                    mean_loss=cfg.WSL.MEAN_LOSS, cmil=cfg.WSL.CMIL, display=int(1280 / cfg.SOLVER.IMS_PER_BATCH),
                    max_epoch=int(cfg.SOLVER.MAX_ITER / cfg.WSL.SIZE_EPOCH), size_epoch=cfg.WSL.SIZE_EPOCH)

    def forward(self, x, proposals=None, context=False, gt_classes_img_oh=None):
        """-> (scores [R, K], proposal_deltas [R, 4 * num_bbox_reg_classes] zeros).
        gt_classes_img_oh (extension, optional): image-level one-hot labels [N, K]; when the caller already has them
        the BCE loss comes out of the same launch and `losses()` reuses it."""
        if context:
            C, D = self.forward_contextlocnet(x, proposals=proposals)
        else:
            if x.dim() > 2:
                x = torch.flatten(x, start_dim=1)
            C = self.cls(x)
            D = self.det(x)
        merged = self.cmil and self.training
        if merged:
            C, D, Oscores, Oproposal_deltas = self.forward_cmil(C, D, proposals)
        counts = [C.shape[0]] if merged else _counts(proposals, C.shape[0])
        scores, img_sum, img_scores, loss = Fn.wsddn(C, D, counts, gt_classes_img_oh, self.mean_loss)
        # the extra results of the launch ride on the scores tensor (the reference's return type is a plain 2-tuple)
        scores._jtsm_mil = dict(img_sum=img_sum, img_scores=img_scores, counts=counts,
                                loss=loss if gt_classes_img_oh is not None else None, labels=gt_classes_img_oh)
        proposal_deltas = torch.zeros(scores.shape[0], self.num_bbox_reg_classes * self.box_dim, dtype=scores.dtype,
                                      device=scores.device, requires_grad=False)
        if merged:
            return scores, proposal_deltas, Oscores, Oproposal_deltas
        return scores, proposal_deltas

    def forward_contextlocnet(self, x, proposals=None):
        """ContextLocNet streams (fast_rcnn_wsddn.py:600-619): `x` = (RoI, frame, context) features -- the three pools of
        ROILoopPool after the box head -- -> (cls(x), det(frame) - det(context))."""
        x, Fx, Cx = x[:]
        x, Fx, Cx = (t.flatten(start_dim=1) if t.dim() > 2 else t for t in (x, Fx, Cx))
        return self.cls(x), self.det(Fx) - self.det(Cx)

    def forward_cmil(self, C, D, proposals):
        """fast_rcnn_wsddn.py:621-667: -> (MC, MD, scores of the un-merged rows, zero deltas).  One image per call, as in
        the reference (its ROIMerge asserts a square IoU matrix, ROIMerge_cpu.cpp:58)."""
        assert proposals and len(proposals) == 1, "WSL.CMIL merges the proposals of ONE image (square IoU matrix)"
        scores, _, _, _ = Fn.wsddn(C, D, [C.shape[0]], None, self.mean_loss)
        proposal_deltas = torch.zeros(scores.shape[0], self.num_bbox_reg_classes * self.box_dim, dtype=scores.dtype,
                                      device=scores.device, requires_grad=False)
        with torch.no_grad():
            obn = scores.detach().sum(dim=1, keepdim=True)
            boxes = proposals[0].proposal_boxes.tensor
            J = ops.pairwise_iou(boxes, boxes)
        MC, MD = self.roi_merge(obn, J, C, D)
        return MC, MD, scores, proposal_deltas

    # ------------------------------------------------------------------ training
    def _img_sum(self, scores, proposals):
        st = getattr(scores, "_jtsm_mil", None)
        if st is not None:
            return st["img_sum"]
        counts = _counts(proposals, scores.shape[0])  # scores that did not come from forward(): eager sums
        return torch.cat([s.sum(dim=0, keepdim=True) for s in scores.split(counts, dim=0)], dim=0)

    def losses(self, predictions, proposals, gt_classes_img_oh):
        scores, _ = predictions
        st = getattr(scores, "_jtsm_mil", None)
        if st is not None and st["loss"] is not None and st["labels"] is gt_classes_img_oh:
            loss = st["loss"].reshape(())
        else:
            p = torch.clamp(self._img_sum(scores, proposals), min=1e-6, max=1.0 - 1e-6)
            oh = gt_classes_img_oh.to(p.dtype)
            if self.mean_loss:
                loss = F.binary_cross_entropy(p, oh, reduction="mean")
            else:
                loss = F.binary_cross_entropy(p, oh, reduction="sum") / oh.size(0)
        losses = {"loss_cls": loss}
        return {k: v * self.loss_weight.get(k, 1.0) for k, v in losses.items()}

    def predict_probs_img(self, predictions, proposals):
        scores, _ = predictions
        st = getattr(scores, "_jtsm_mil", None)
        if st is not None:
            return st["img_scores"]
        return torch.clamp(self._img_sum(scores, proposals), min=1e-6, max=1.0 - 1e-6)

    # ------------------------------------------------------------------ inference
    def predict_boxes(self, predictions, proposals):
        if not len(proposals):
            return []
        _, proposal_deltas = predictions
        boxes = self.box2box_transform.apply_deltas(proposal_deltas, _cat_boxes(proposals))
        return boxes.split([len(p) for p in proposals])

    def predict_boxes_for_gt_classes(self, predictions, proposals):
        if not len(proposals):
            return []
        _, proposal_deltas = predictions
        pb = _cat_boxes(proposals)
        n, b = pb.shape
        boxes = self.box2box_transform.apply_deltas(proposal_deltas, pb)
        k = boxes.shape[1] // b
        if k > 1:
            gt = torch.cat([p.gt_classes for p in proposals], dim=0).clamp(0, k - 1)
            boxes = boxes.view(n, k, b)[torch.arange(n, dtype=torch.long, device=boxes.device), gt]
        return boxes.split([len(p) for p in proposals])

    def predict_probs(self, predictions, proposals):
        scores, _ = predictions
        bg = torch.zeros(scores.shape[0], 1, dtype=scores.dtype, device=scores.device, requires_grad=False)
        return torch.cat((scores, bg), 1).split([len(p) for p in proposals], dim=0)

    def inference(self, predictions, proposals):
        boxes = self.predict_boxes(predictions, proposals)
        scores = self.predict_probs(predictions, proposals)
        image_shapes = [x.image_size for x in proposals]
        return fast_rcnn_inference(boxes, scores, image_shapes, self.test_score_thresh, self.test_nms_thresh,
                                   self.test_topk_per_image)
