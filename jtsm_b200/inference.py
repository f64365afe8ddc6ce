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
    -> dict(pred_boxes, scores, pred_classes, pred_inds), kept RoI indices, all_scores, all_boxes -- the reference's
    return values (fast_rcnn_oicr.py:95-163).  The whole function is ONE C-ABI call (jtsm_fast_rcnn_inference: finite
    mask, background drop, clip, threshold, ordered compaction, per-class NMS, top-k; no host sync inside); the one
    host read here is the detection count, because the reference's signature returns tensors of that length.
    `fast_rcnn_inference_device` keeps everything padded on the device."""
    all_scores, all_boxes = scores.clone().unsqueeze(0), boxes.clone().unsqueeze(0)
    d = ops.fast_rcnn_inference_raw(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image)
    n = int(d["num"].item())
    result = dict(pred_boxes=d["pred_boxes"][:n], scores=d["scores"][:n], pred_classes=d["pred_classes"][:n],
                  pred_inds=d["pred_inds"][:n])
    return result, d["rows"][:n], all_scores, all_boxes


def fast_rcnn_inference_device(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image):
    """Same computation, results left padded on the device (`num` int64 [1] says how many rows are detections):
    no host synchronisation at all, CUDA-graph capturable (bench.py's c4 inference chain)."""
    return ops.fast_rcnn_inference_raw(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image)


def fast_rcnn_inference(boxes, scores, image_shapes, score_thresh, nms_thresh, topk_per_image):
    per_image = [fast_rcnn_inference_single_image(b, s, shp, score_thresh, nms_thresh, topk_per_image)
                 for s, b, shp in zip(scores, boxes, image_shapes)]
    return tuple([x[i] for x in per_image] for i in range(4))
