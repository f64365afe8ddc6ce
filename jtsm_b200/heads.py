"""Head math of the WSL ROI heads on top of the fused kernels (ops.py -> C ABI).

Functional mirror of what `OICRROIHeads._forward_box` does after fc6/fc7
(projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:404-449): WSDDN scores + BCE, then per refinement
branch pseudo-gt -> label assignment -> weighted CE (+ box regression) -> softmax / decoded boxes.
Everything stays on the device; no host synchronisation inside (the reference does 3-7 `.item()` /
`nonzero` syncs per branch and image)."""
import torch

from . import ops


def head_step(mil_cls, mil_det, ref_logits, ref_deltas, proposals, num_preds_per_image, gt_classes_img_int,
              num_classes, labels_onehot, mean_loss=True, reg=False, reg_weighted=True, iou_thresholds=(0.5,),
              iou_labels=(0, 1), need_grad=True, term_weights=None, smooth_l1_beta=0.0,
              bbox_reg_weights=(10.0, 10.0, 5.0, 5.0)):
    """proposals: [R,4] (all images concatenated).  Returns dict(losses, scores, img_scores, branches[...],
    grad_cls, grad_det, grad_ref_logits, grad_ref_deltas).  Launches: 1 (MIL fwd+loss) + 1 (MIL bwd)
    + per branch 2 (assign, CE fwd/bwd/softmax) [+ 2 with box regression]."""
    fwd = ops.wsddn_forward(mil_cls, mil_det, num_preds_per_image, labels_onehot, mean_loss)
    out = dict(losses={"loss_cls": fwd["loss"]}, scores=fwd["scores"], img_scores=fwd["img_scores"], branches=[])
    prev_scores, prev_boxes = fwd["scores"], proposals
    grads_l, grads_d = [], []
    boxes_noreg = None
    for k, logits in enumerate(ref_logits):
        tw = 1.0 if term_weights is None else float(term_weights[k])
        a = ops.refine_assign(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int,
                              fwd["img_scores"], num_classes, iou_thresholds, iou_labels, want_iou=True)
        ce = ops.refine_ce(logits, a["gt_classes"], a["loss_weights"], a["valid_count"], loss_scale=tw,
                           want_grad=need_grad)
        out["losses"]["loss_cls_r%d" % k] = ce["loss"]
        grads_l.append(ce["grad_logits"])
        if reg:
            br = ops.refine_box_reg(ref_deltas[k], proposals, a["gt_boxes"], a["gt_classes"],
                                    a["loss_weights"] if reg_weighted else None, num_classes, bbox_reg_weights,
                                    smooth_l1_beta, loss_scale=tw, want_grad=need_grad)
            out["losses"]["loss_box_reg_r%d" % k] = br["loss"]
            grads_d.append(br["grad_deltas"])
            boxes = ops.apply_deltas(ref_deltas[k], proposals, bbox_reg_weights)
        else:
            # zero deltas: predict_boxes still round-trips through (ctr, size) (fast_rcnn_oicr.py:684-709)
            if boxes_noreg is None:
                boxes_noreg = ops.apply_deltas(torch.zeros(proposals.shape[0], 4, device=proposals.device),
                                               proposals, bbox_reg_weights)
            boxes = boxes_noreg
        a.update(probs=ce["probs"], boxes=boxes, acc=ce["acc"])
        out["branches"].append(a)
        prev_scores, prev_boxes = ce["probs"], boxes
    if need_grad:
        out["grad_cls"], out["grad_det"] = ops.wsddn_backward(fwd)
        out["grad_ref_logits"], out["grad_ref_deltas"] = grads_l, grads_d
    return out
