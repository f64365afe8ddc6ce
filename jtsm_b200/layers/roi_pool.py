"""RoIPool: drop-in for `torchvision.ops.RoIPool` / `roi_pool`, which is what the reference's
poolers construct for POOLER_TYPE "ROIPool" (projects/WSL/wsl/modeling/poolers.py:183-186,
detectron2/modeling/poolers.py:161-164).  Same signature, CUDA (sm_100a) only.

Extension over the reference: an optional per-RoI `row_scale` fuses
`box_features * (objectness_logits + 1)` (roi_heads_oicr.py:393-394) into the pooling
epilogue and its backward.
"""
import torch
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn.modules.utils import _pair

from .. import ops


def _boxes_to_rois(boxes):
    """torchvision accepts Tensor[K,5] or a list of Tensor[L,4] (one per image)."""
    if isinstance(boxes, (list, tuple)):
        parts = []
        for i, b in enumerate(boxes):
            idx = torch.full((b.shape[0], 1), float(i), dtype=b.dtype, device=b.device)
            parts.append(torch.cat([idx, b], dim=1))
        return torch.cat(parts, dim=0) if parts else torch.zeros((0, 5))
    return boxes


class _RoIPool(Function):
    @staticmethod
    def forward(ctx, input, roi, output_size, spatial_scale, row_scale):
        ctx.output_size = _pair(output_size)
        ctx.spatial_scale = spatial_scale
        ctx.input_shape = input.size()
        # argmax is internal to this op: keep it in 16 bits when the plane allows it
        hw = input.shape[2] * input.shape[3]
        adt = torch.int16 if hw < 65535 else torch.int32
        output, argmax = ops.roi_pool_forward(
            input, roi, spatial_scale, ctx.output_size, row_scale=row_scale, argmax_dtype=adt
        )
        ctx.save_for_backward(roi, argmax, row_scale)
        return output

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        rois, argmax, row_scale = ctx.saved_tensors
        grad_input = ops.roi_pool_backward(
            grad_output, rois, argmax, ctx.input_shape, ctx.output_size, row_scale=row_scale
        )
        return grad_input, None, None, None, None


def roi_pool(input, boxes, output_size, spatial_scale=1.0, row_scale=None):
    rois = _boxes_to_rois(boxes)
    assert rois.dim() == 2 and rois.size(1) == 5
    return _RoIPool.apply(input, rois, output_size, spatial_scale, row_scale)


class RoIPool(nn.Module):
    def __init__(self, output_size, spatial_scale):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input, rois, row_scale=None):
        return roi_pool(input, rois, self.output_size, self.spatial_scale, row_scale=row_scale)

    def __repr__(self):
        return "%s(output_size=%s, spatial_scale=%s)" % (
            self.__class__.__name__, self.output_size, self.spatial_scale)
