"""ROILoopPool: drop-in for `wsl.layers.ROILoopPool` (projects/WSL/wsl/layers/roi_loop_pool.py:9-58).
Same autograd contract: argmax saved and non-differentiable, once-differentiable backward."""
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn.modules.utils import _pair

from .. import ops


class _ROILoopPool(Function):
    @staticmethod
    def forward(ctx, input, roi, output_size, spatial_scale):
        ctx.output_size = _pair(output_size)
        ctx.spatial_scale = spatial_scale
        ctx.input_shape = input.size()
        output, argmax = ops.roi_loop_pool_forward(input, roi, spatial_scale, ctx.output_size[0], ctx.output_size[1])
        ctx.save_for_backward(roi, argmax)
        ctx.mark_non_differentiable(argmax)
        return output

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        (rois, argmax) = ctx.saved_tensors
        bs, ch, h, w = ctx.input_shape
        grad_input = ops.roi_loop_pool_backward(grad_output, rois, argmax, ctx.spatial_scale, ctx.output_size[0],
                                                ctx.output_size[1], bs, ch, h, w)
        return grad_input, None, None, None


roi_loop_pool = _ROILoopPool.apply


class ROILoopPool(nn.Module):
    def __init__(self, output_size, spatial_scale):
        super(ROILoopPool, self).__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input, rois):
        """input: NCHW images; rois: Bx5 boxes (batch index, x1, y1, x2, y2) -> [3B, C, PH, PW]."""
        assert rois.dim() == 2 and rois.size(1) == 5
        return roi_loop_pool(input, rois, self.output_size, self.spatial_scale)

    def __repr__(self):
        tmpstr = self.__class__.__name__ + "("
        tmpstr += "output_size=" + str(self.output_size)
        tmpstr += ", spatial_scale=" + str(self.spatial_scale)
        tmpstr += ")"
        return tmpstr
