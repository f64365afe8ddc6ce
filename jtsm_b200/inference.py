"""Inference-side mirror of the WSL heads (projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py):
`predict_probs_K` / `predict_boxes_K` (:712-783) and `fast_rcnn_inference(_single_image)` (:44-163), on the
sm_100a ops.  Results come back as plain dicts of tensors (the reference wraps them in `Instances`)."""
import torch

from . import ops


def predict_probs_K(scores_K):
    """Mean over the refinement branches of softmax(scores_k); fast_rcnn_oicr.py:764-783."""
    out = None
    for s in scores_K:
        out = ops.softmax_rows(s, out, scale=1.0, accumulate=out is not None)
    return out / len(scores_K)


def predict_boxes_K(deltas_K, proposal_boxes, weights=(10.0, 10.0, 5.0, 5.0)):
    """apply_deltas(mean of the branches' deltas); fast_rcnn_oicr.py:712-742."""
    deltas = torch.zeros_like(deltas_K[0])
    for d in deltas_K:
        deltas += d
    deltas = deltas / len(deltas_K)
    return ops.apply_deltas(deltas, proposal_boxes, weights)


def fast_rcnn_inference_single_image(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image):
    """boxes [R, 4] or [R, 4K], scores [R, K+1] (background last), image_shape (h, w).
    -> dict(pred_boxes, scores, pred_classes, pred_inds), kept RoI indices, all_scores, all_boxes."""
    all_scores, all_boxes = scores.clone().unsqueeze(0), boxes.clone().unsqueeze(0)
    r, k1 = scores.shape
    pred_inds = torch.arange(r, device=scores.device, dtype=torch.long).unsqueeze(1).repeat(1, k1)
    valid = torch.isfinite(boxes).all(dim=1) & torch.isfinite(scores).all(dim=1)
    if not valid.all():
        boxes, scores, pred_inds = boxes[valid], scores[valid], pred_inds[valid]
    scores = scores[:, :-1]
    nb = boxes.shape[1] // 4
    h, w = image_shape
    boxes = boxes.reshape(-1, 4).clone()
    boxes[:, 0].clamp_(min=0, max=w)  # Boxes.clip, detectron2/structures/boxes.py:195-211
    boxes[:, 1].clamp_(min=0, max=h)
    boxes[:, 2].clamp_(min=0, max=w)
    boxes[:, 3].clamp_(min=0, max=h)
    boxes = boxes.view(-1, nb, 4)
    pred_inds = pred_inds[:, :-1]
    mask = scores > score_thresh
    filter_inds = mask.nonzero()
    boxes = boxes[filter_inds[:, 0], 0] if nb == 1 else boxes[mask]
    scores = scores[mask]
    pred_inds = pred_inds[mask]
    keep = ops.batched_nms(boxes, scores, filter_inds[:, 1], nms_thresh)
    if topk_per_image >= 0:
        keep = keep[:topk_per_image]
    boxes, scores, filter_inds, pred_inds = boxes[keep], scores[keep], filter_inds[keep], pred_inds[keep]
    result = dict(pred_boxes=boxes, scores=scores, pred_classes=filter_inds[:, 1], pred_inds=pred_inds)
    return result, filter_inds[:, 0], all_scores, all_boxes


def fast_rcnn_inference(boxes, scores, image_shapes, score_thresh, nms_thresh, topk_per_image):
    per_image = [fast_rcnn_inference_single_image(b, s, shp, score_thresh, nms_thresh, topk_per_image)
                 for s, b, shp in zip(scores, boxes, image_shapes)]
    return tuple([x[i] for x in per_image] for i in range(4))
