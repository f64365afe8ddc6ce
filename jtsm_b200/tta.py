"""Host mirror of the reference's test-time-augmentation average
(projects/WSL/wsl/modeling/test_time_augmentation_avg.py) on the sm_100a ops -- SURVEY 8f-1.

What stays on the host is what the reference keeps on the host and is integer / shape logic: which resized shapes
and flips to run (DatasetMapperTTAAVG.__call__, :147-196; ResizeShortestEdge.get_transform,
detectron2/data/transforms/augmentation_impl.py:152-172) and the transform parameter lists.  What moves to the
device: the inverse box mapping + mean over augmentations (`_get_augmented_boxes`, :364-390 -- a D2H / numpy / H2D
round trip per augmentation in the reference), the proposal transform (`transform_proposals`, :27-70) and the
threshold + per-class NMS that follows (`_merge_detections`, :392-404 -> inference.fast_rcnn_inference_single_image).

The transform classes carry parameters only (same names and constructor arguments as fvcore.transforms /
detectron2.data.transforms); `apply_box` runs on CUDA tensors.
"""
import numpy as np
import torch

from . import inference, ops


class Transform:
    def inverse(self):
        raise NotImplementedError

    def stages(self):
        """-> [(type, a, b)] for the C ABI (include/jtsm_b200.h: JTSM_TTA_*)."""
        raise NotImplementedError

    def apply_box(self, boxes, image_shape=None):
        return TransformList([self]).apply_box(boxes, image_shape)

    def __add__(self, other):
        other = other.transforms if isinstance(other, TransformList) else [other]
        return TransformList([self] + other)


class NoOpTransform(Transform):
    def inverse(self):
        return self

    def stages(self):
        return [(ops.TTA_NOOP, 0.0, 0.0)]


class HFlipTransform(Transform):
    def __init__(self, width):
        self.width = width

    def inverse(self):
        return self

    def stages(self):
        return [(ops.TTA_HFLIP, float(self.width), 0.0)]


class ResizeTransform(Transform):
    """detectron2/data/transforms/transform.py:94-159 (coordinates only)."""

    def __init__(self, h, w, new_h, new_w, interp=None):
        self.h, self.w, self.new_h, self.new_w, self.interp = h, w, new_h, new_w, interp

    def inverse(self):
        return ResizeTransform(self.new_h, self.new_w, self.h, self.w, self.interp)

    def stages(self):
        # the reference multiplies a float32 numpy array by the Python float new_w * 1.0 / w (:150-151):
        # the factor is rounded to float32 once, the product is an fp32 product
        return [(ops.TTA_SCALE, float(np.float32(self.new_w * 1.0 / self.w)), float(np.float32(self.new_h * 1.0 / self.h)))]


class TransformList(Transform):
    def __init__(self, transforms):
        self.transforms = []
        for t in transforms:
            self.transforms.extend(t.transforms if isinstance(t, TransformList) else [t])

    def __add__(self, other):
        other = other.transforms if isinstance(other, TransformList) else [other]
        return TransformList(self.transforms + other)

    def __radd__(self, other):
        other = other.transforms if isinstance(other, TransformList) else [other]
        return TransformList(other + self.transforms)

    def __len__(self):
        return len(self.transforms)

    def inverse(self):
        return TransformList([t.inverse() for t in self.transforms[::-1]])

    def stages(self):
        out = []
        for t in self.transforms:
            out.extend(t.stages())
        return out

    def apply_box(self, boxes, image_shape=None):
        """boxes [n, 4] CUDA -> transformed [n, 4]; no clipping when image_shape is None."""
        big = float(np.finfo(np.float32).max)
        shape = image_shape if image_shape is not None else (big, big)
        b = boxes.reshape(-1, 4)
        if image_shape is None:
            # clip(min=0) must not touch the boxes: run the merge kernel with a single "augmentation" instead
            out, _ = ops.tta_merge([b], None, [self.stages()])
            return out
        out, _ = ops.tta_transform_boxes(b, self.stages(), shape)
        return out


def resize_shortest_edge_shape(h, w, size, max_size):
    """ResizeShortestEdge.get_transform, augmentation_impl.py:152-172 -> (new_h, new_w)."""
    scale = size * 1.0 / min(h, w)
    if h < w:
        newh, neww = size, scale * w
    else:
        newh, neww = scale * h, size
    if max(newh, neww) > max_size:
        scale = max_size * 1.0 / max(newh, neww)
        newh = newh * scale
        neww = neww * scale
    return int(newh + 0.5), int(neww + 0.5)


def tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip):
    """The transform list of every augmentation DatasetMapperTTAAVG.__call__ (:147-196) generates, in its order:
    for each min_size: resize, then (if flip) resize + flip; each prefixed by pre_tfm (orig -> input image resize)."""
    h, w = image_hw
    if tuple(image_hw) != tuple(orig_hw):
        pre = ResizeTransform(orig_hw[0], orig_hw[1], h, w)
    else:
        pre = NoOpTransform()
    out = []
    for size in min_sizes:
        nh, nw = resize_shortest_edge_shape(h, w, size, max_size)
        resize = ResizeTransform(h, w, nh, nw)
        out.append(pre + TransformList([resize]))
        if flip:
            out.append(pre + TransformList([resize, HFlipTransform(nw)]))
    return out


def transform_proposals(boxes, objectness_logits, image_shape, transforms, proposal_topk, min_box_size=0,
                        oh_labels=None):
    """transform_proposals / transform_proposals_seg (:27-117): transform, clip, drop empty boxes, keep the first
    proposal_topk.  boxes [n, 4] CUDA.  -> dict(proposal_boxes, objectness_logits[, oh_labels])."""
    out, keep = ops.tta_transform_boxes(boxes, transforms.stages(), image_shape, min_box_size)
    res = dict(proposal_boxes=out[keep][:proposal_topk], objectness_logits=objectness_logits[keep][:proposal_topk])
    if oh_labels is not None:
        res["oh_labels"] = oh_labels[keep][:proposal_topk]
    return res


def get_augmented_boxes(all_boxes, all_scores, tfms):
    """_get_augmented_boxes (:364-390): all_boxes / all_scores = per-augmentation outputs of model.inference
    ([1, R, 4K] / [1, R, K+1] CUDA tensors), tfms = the augmentations' transform lists.
    -> (mean boxes [R, 4K], mean scores [R, K+1]) on the device, one launch, no host round trip."""
    inv = [t.inverse().stages() for t in tfms]
    return ops.tta_merge(all_boxes, all_scores, inv)


def merge_detections(all_boxes, all_scores, shape_hw, score_thresh, nms_thresh, detections_per_image):
    """_merge_detections (:392-404)."""
    result, _, _, _ = inference.fast_rcnn_inference_single_image(all_boxes, all_scores, shape_hw, score_thresh,
                                                                 nms_thresh, detections_per_image)
    return result


def inference_one_image_boxes(all_boxes, all_scores, tfms, orig_shape, score_thresh, nms_thresh,
                              detections_per_image):
    """The box part of GeneralizedRCNNWithTTAAVG._inference_one_image (:321-334) given the augmented passes."""
    boxes, scores = get_augmented_boxes(all_boxes, all_scores, tfms)
    return merge_detections(boxes, scores, orig_shape, score_thresh, nms_thresh, detections_per_image)
