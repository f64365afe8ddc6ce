"""MOIPool: drop-in for `wsl.layers.MOIPool` (projects/WSL/wsl/layers/moi_pool.py:10-88).
forward(input, rois, oh_labels, superpixels) -> (output, argmax); argmax is an API output
(consumed by roi_heads_jtsm.py:615-630) and non-differentiable.
Extension: forward(..., row_scale=objectness_logits + 1, mask_scale=True) fuses the two per-RoI multiplies the
JTSM heads apply right after pooling (roi_heads_jtsm.py:615-633) into the kernel epilogue."""
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn.modules.utils import _pair

from .. import ops


class _MOIPool(Function):
    @staticmethod
    def forward(ctx, input, roi, output_size, spatial_scale, oh_labels, superpixels, row_scale=None, mask_scale=False):
        ctx.output_size = _pair(output_size)
        ctx.spatial_scale = spatial_scale
        ctx.input_shape = input.size()
        output, argmax, scales = ops.moi_pool_forward(input, roi, spatial_scale, ctx.output_size[0], ctx.output_size[1],
                                                      oh_labels, superpixels, row_scale, mask_scale, return_scales=True)
        ctx.fused = scales is not None
        if ctx.fused:
            ctx.save_for_backward(roi, argmax, scales)
        else:
            ctx.save_for_backward(roi, argmax)
        ctx.mark_non_differentiable(argmax)
        return output, argmax

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output, _grad_argmax=None):
        rois, argmax = ctx.saved_tensors[:2]
        bs, ch, h, w = ctx.input_shape
        grad_input = ops.moi_pool_backward(grad_output, rois, argmax, ctx.spatial_scale, ctx.output_size[0],
                                           ctx.output_size[1], bs, ch, h, w,
                                           row_scale=ctx.saved_tensors[2][2] if ctx.fused else None)
        return grad_input, None, None, None, None, None, None, None


moi_pool = _MOIPool.apply


class MOIPool(nn.Module):
    def __init__(self, output_size, spatial_scale):
        super(MOIPool, self).__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input, rois, oh_labels, superpixels, row_scale=None, mask_scale=False):
        assert rois.dim() == 2 and rois.size(1) == 5
        return moi_pool(input, rois, self.output_size, self.spatial_scale, oh_labels, superpixels, row_scale, mask_scale)

    def __repr__(self):
        tmpstr = self.__class__.__name__ + "("
        tmpstr += "output_size=" + str(self.output_size)
        tmpstr += ", spatial_scale=" + str(self.spatial_scale)
        tmpstr += ")"
        return tmpstr
