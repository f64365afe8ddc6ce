"""CPU restatement of the reference's TTA-AVG merge (TEST INFRASTRUCTURE ONLY; travels to the GPU box).

Follows projects/WSL/wsl/modeling/test_time_augmentation_avg.py:
  get_augmented_boxes   _get_augmented_boxes :364-390  (inverse transform per augmentation in numpy, torch.mean)
  transform_proposals   transform_proposals :27-70     (apply_box, Boxes.clip, Boxes.nonempty, [:topk])
  tta_transforms        DatasetMapperTTAAVG.__call__ :147-196 + ResizeShortestEdge.get_transform
                        (detectron2/data/transforms/augmentation_impl.py:152-172)
  ResizeTransform       detectron2/data/transforms/transform.py:94-159 (coordinates only)
on top of oracle/fvcore_transforms.py (fvcore restated: unpinned).  Pinned to the reference's own code run in place
by tests/test_tta.py (needs /root/reference) and by tests/golden/tta_*.npz generated from it.
"""
import numpy as np
import torch

from .fvcore_transforms import HFlipTransform, NoOpTransform, Transform, TransformList


class ResizeTransform(Transform):
    def __init__(self, h, w, new_h, new_w, interp=None):
        super().__init__()
        self._set_attributes(locals())

    def apply_coords(self, coords):
        coords[:, 0] = coords[:, 0] * (self.new_w * 1.0 / self.w)
        coords[:, 1] = coords[:, 1] * (self.new_h * 1.0 / self.h)
        return coords

    def inverse(self):
        return ResizeTransform(self.new_h, self.new_w, self.h, self.w, self.interp)


def resize_shortest_edge_shape(h, w, size, max_size):
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


def tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip, resize_cls=ResizeTransform):
    h, w = image_hw
    pre = resize_cls(orig_hw[0], orig_hw[1], h, w) if tuple(image_hw) != tuple(orig_hw) else NoOpTransform()
    out = []
    for size in min_sizes:
        nh, nw = resize_shortest_edge_shape(h, w, size, max_size)
        resize = resize_cls(h, w, nh, nw)
        out.append(pre + TransformList([resize]))
        if flip:
            out.append(pre + TransformList([resize, HFlipTransform(nw)]))
    return out


def mean_sequential(x):
    """Mean over dim 0 as ONE sequential fp32 sum in list order followed by a true division -- the arithmetic the
    device kernel is specified to do.  torch.mean(dim=0) on the CPU computes exactly this for every element of the
    vectorised body; in the scalar tail of the inner dimension (numel % (4 x SIMD width)) ATen sums the rows in four
    interleaved accumulators instead, so those few elements can differ by one ulp -- and which elements they are
    depends on the host's SIMD width.  (On CUDA torch.mean multiplies by 1/A instead of dividing.)"""
    acc = torch.zeros_like(x[0])
    for a in range(x.shape[0]):
        acc = acc + x[a]
    return acc / x.shape[0]


def get_augmented_boxes(all_boxes, all_scores, tfms, mean=None):
    """all_boxes: list of [1, R, 4K] tensors, all_scores: list of [1, R, K+1] -> (mean boxes, mean scores).
    mean: None = torch.mean as the reference calls it; `mean_sequential` = the machine-independent definition."""
    all_boxes = list(all_boxes)
    for idx, tfm in enumerate(tfms):
        pred_boxes = all_boxes[idx]
        num_img, num_pred, num_col = pred_boxes.shape
        assert num_img == 1
        orig = tfm.inverse().apply_box(pred_boxes.reshape(num_pred * num_col // 4, 4).cpu().numpy())
        all_boxes[idx] = torch.from_numpy(orig).reshape(1, num_pred, num_col)
    if mean is not None:
        return mean(torch.cat(all_boxes, dim=0)), mean(torch.cat(list(all_scores), dim=0))
    boxes = torch.mean(torch.cat(all_boxes, dim=0), dim=0, keepdim=False)
    scores = torch.mean(torch.cat(list(all_scores), dim=0), dim=0, keepdim=False)
    return boxes, scores


def transform_proposals(boxes, objectness_logits, image_shape, transforms, proposal_topk, min_box_size=0):
    b = torch.as_tensor(transforms.apply_box(boxes.cpu().numpy())).to(torch.float32).reshape(-1, 4).clone()
    h, w = image_shape
    b[:, 0].clamp_(min=0, max=w)
    b[:, 1].clamp_(min=0, max=h)
    b[:, 2].clamp_(min=0, max=w)
    b[:, 3].clamp_(min=0, max=h)
    keep = ((b[:, 2] - b[:, 0]) > min_box_size) & ((b[:, 3] - b[:, 1]) > min_box_size)
    return dict(proposal_boxes=b[keep][:proposal_topk], objectness_logits=objectness_logits[keep][:proposal_topk],
                all_boxes=b, keep=keep)


def synth_augmented_outputs(g, proposal_boxes, tfms, num_classes, jitter=2.0):
    """Synthetic per-augmentation inference outputs: per-class boxes = the proposals mapped INTO
    each augmented image plus a small per-class jitter, scores = softmax of random logits; fp32, [1, R, 4K] / [1, R, K+1]."""
    r = proposal_boxes.shape[0]
    boxes, scores = [], []
    for t in tfms:
        fwd = torch.from_numpy(t.apply_box(proposal_boxes.numpy().copy())).to(torch.float32)
        per_class = fwd[:, None, :].repeat(1, num_classes, 1)
        per_class = per_class + (torch.rand(r, num_classes, 4, generator=g) - 0.5) * jitter
        lo = torch.minimum(per_class[..., :2], per_class[..., 2:])
        hi = torch.maximum(per_class[..., :2], per_class[..., 2:]) + 1.0
        boxes.append(torch.cat([lo, hi], dim=-1).reshape(1, r, 4 * num_classes).contiguous())
        scores.append(torch.softmax(torch.randn(r, num_classes + 1, generator=g) * 3, dim=1).reshape(1, r, num_classes + 1))
    return boxes, scores
