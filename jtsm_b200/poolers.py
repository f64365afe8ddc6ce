"""ROIPooler: host-side mirror of the reference's multi-level RoI pooler
(projects/WSL/wsl/modeling/poolers.py:24-108 helpers, :119-331 ROIPooler; the detectron2 twin is
detectron2/modeling/poolers.py:147-249) built on the sm_100a ops.

Same constructor arguments, same `forward(x, box_lists, level_ids=None, oh_labels_list=None, superpixels=None)`,
same level-assignment rule and output layout (ROILoopPool: `[3M, C, P, P]`).  Boxes may be detectron2 `Boxes`
(anything with a `.tensor` attribute) or plain `[R_i, 4]` tensors.  Extensions: `row_scale=` (per-RoI factor,
`objectness_logits + 1` in the WSL heads, roi_heads_oicr.py:393-394) is fused into the RoIPool / MOIPool kernels,
`mask_scale=True` adds MOIPool's P*P / (#valid bins + 1) factor of the JTSM heads (roi_heads_jtsm.py:615-630).
"""
import math

import torch
from torch import nn

from .layers import MOIPool, ROIAlign, ROILoopPool, RoIPool

VALID_RANGE = [[0, 60], [60, 160], [160, 2000]]  # wsl/modeling/poolers.py:110-116 (use_range=True)


def _t(boxes):
    return boxes.tensor if hasattr(boxes, "tensor") else boxes


def _sizes(box_lists):
    b = torch.cat([_t(x) for x in box_lists], dim=0)
    return torch.sqrt((b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1]))


def assign_boxes_to_levels(box_lists, min_level, max_level, canonical_box_size, canonical_level, valid_range=None):
    """Level index (offset from min_level) of every box; wsl/modeling/poolers.py:24-71."""
    box_sizes = _sizes(box_lists)
    if valid_range is not None:
        levels = torch.full_like(box_sizes, -1)
        for level, (lo, hi) in enumerate(valid_range):
            levels[(box_sizes >= lo) & (box_sizes < hi)] = level
        return levels.to(torch.int64)
    # Eqn. (1) of the FPN paper
    levels = torch.floor(canonical_level + torch.log2(box_sizes / canonical_box_size + 1e-8))
    return torch.clamp(levels, min=min_level, max=max_level).to(torch.int64) - min_level


def convert_boxes_to_pooler_format(box_lists):
    """list of per-image boxes -> [M, 5] (batch index, x0, y0, x1, y1); wsl/modeling/poolers.py:81-108."""
    parts = []
    for i, boxes in enumerate(box_lists):
        b = _t(boxes)
        parts.append(torch.cat((torch.full_like(b[:, :1], i), b), dim=1))
    return torch.cat(parts, dim=0)


class ROIPooler(nn.Module):
    def __init__(self, output_size, scales, sampling_ratio, pooler_type, canonical_box_size=224, canonical_level=4,
                 use_range=False):
        super().__init__()
        if isinstance(output_size, int):
            output_size = (output_size, output_size)
        assert len(output_size) == 2 and all(isinstance(v, int) for v in output_size)
        self.output_size = output_size
        makers = {
            "ROIAlign": lambda s: ROIAlign(output_size, spatial_scale=s, sampling_ratio=sampling_ratio, aligned=False),
            "ROIAlignV2": lambda s: ROIAlign(output_size, spatial_scale=s, sampling_ratio=sampling_ratio, aligned=True),
            "ROIPool": lambda s: RoIPool(output_size, spatial_scale=s),
            "MOIPool": lambda s: MOIPool(output_size, spatial_scale=s),
            "ROILoopPool": lambda s: ROILoopPool(output_size, spatial_scale=s),
        }
        if pooler_type not in makers:  # ROIAlignRotated is outside the hot path
            raise ValueError("Unknown pooler type: {}".format(pooler_type))
        self.pooler_type = pooler_type
        self.level_poolers = nn.ModuleList(makers[pooler_type](s) for s in scales)
        min_level, max_level = -math.log2(scales[0]), -math.log2(scales[-1])
        assert math.isclose(min_level, int(min_level)) and math.isclose(max_level, int(max_level)), \
            "Featuremap stride is not power of 2!"
        self.min_level, self.max_level = int(min_level), int(max_level)
        assert 0 <= self.min_level <= self.max_level
        assert canonical_box_size > 0
        self.canonical_level = canonical_level
        self.canonical_box_size = canonical_box_size
        self.valid_range = VALID_RANGE if use_range else None

    def forward(self, x, box_lists, level_ids=None, oh_labels_list=None, superpixels=None, row_scale=None,
                mask_scale=False):
        n_levels = len(self.level_poolers)
        assert isinstance(x, list) and isinstance(box_lists, list), "Arguments to pooler must be lists"
        assert len(x) == n_levels, "num_level_assignments={}, but x is list of {} Tensors".format(n_levels, len(x))
        assert len(box_lists) == x[0].size(0), "x[0] batch dim 0 is {}, but box_list has length {}".format(
            x[0].size(0), len(box_lists))
        rois = convert_boxes_to_pooler_format(box_lists)
        labels = None
        if superpixels is not None:
            sp = _t(superpixels)
            width = max(l.size(1) for l in oh_labels_list)  # pad one-hot lists to a common length (:266-275)
            labels = torch.cat([torch.nn.functional.pad(l.to(torch.int32), (0, width - l.size(1))) for l in oh_labels_list])

        def call(level, boxes, lab=None, scale=None):
            pooler = self.level_poolers[level]
            if lab is not None:
                return pooler(x[level], boxes, lab, sp, row_scale=scale, mask_scale=mask_scale)
            if scale is not None:
                assert isinstance(pooler, RoIPool), "row_scale is fused for ROIPool and MOIPool only"
                return pooler(x[level], boxes, row_scale=scale)
            return pooler(x[level], boxes)

        if n_levels == 1:
            return call(0, rois, labels, row_scale)

        levels = assign_boxes_to_levels(box_lists, self.min_level, self.max_level, self.canonical_box_size,
                                        self.canonical_level, self.valid_range)
        if level_ids is not None:
            levels = torch.cat(level_ids).to(torch.int64)
        m, c, p = rois.size(0), x[0].shape[1], self.output_size[0]
        triple = self.pooler_type == "ROILoopPool"
        out = torch.zeros(((3 if triple else 1) * m, c, p, p), dtype=x[0].dtype, device=x[0].device)
        for level in range(n_levels):
            inds = torch.nonzero(levels == level, as_tuple=True)[0]
            res = call(level, rois[inds], labels[inds] if labels is not None else None,
                       row_scale[inds] if row_scale is not None else None)
            if labels is not None:
                res = res[0]  # like the reference (:311-318), the multi-level MOIPool path keeps the features only
            if triple:
                inds = torch.cat([inds, inds + m, inds + 2 * m], dim=0)
            out[inds] = res
        return out
