"""ROIAlign: drop-in for `detectron2.layers.ROIAlign` / `roi_align`
(detectron2/layers/roi_align.py:22-59 `_ROIAlign`, :63-122 `ROIAlign`).  Same constructor, forward,
repr and autograd contract (once-differentiable backward, rois saved, sizes on ctx); CUDA (sm_100a) only."""
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn.modules.utils import _pair

from .. import ops


class _ROIAlign(Function):
    @staticmethod
    def forward(ctx, input, roi, output_size, spatial_scale, sampling_ratio, aligned):
        ctx.save_for_backward(roi)
        ctx.output_size = _pair(output_size)
        ctx.spatial_scale = spatial_scale
        ctx.sampling_ratio = sampling_ratio
        ctx.input_shape = input.size()
        ctx.aligned = aligned
        return ops.roi_align_forward(input, roi, spatial_scale, ctx.output_size[0], ctx.output_size[1],
                                     sampling_ratio, aligned)

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        (rois,) = ctx.saved_tensors
        bs, ch, h, w = ctx.input_shape
        grad_input = ops.roi_align_backward(grad_output, rois, ctx.spatial_scale, ctx.output_size[0],
                                            ctx.output_size[1], bs, ch, h, w, ctx.sampling_ratio, ctx.aligned)
        return grad_input, None, None, None, None, None


roi_align = _ROIAlign.apply


class ROIAlign(nn.Module):
    def __init__(self, output_size, spatial_scale, sampling_ratio, aligned=True):
        """See detectron2/layers/roi_align.py:64-97 for the meaning of `aligned`."""
        super(ROIAlign, self).__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale
        self.sampling_ratio = sampling_ratio
        self.aligned = aligned

    def forward(self, input, rois):
        """input: NCHW images; rois: Bx5 boxes (batch index, x1, y1, x2, y2)."""
        assert rois.dim() == 2 and rois.size(1) == 5
        return roi_align(input, rois.to(dtype=input.dtype), self.output_size, self.spatial_scale,
                         self.sampling_ratio, self.aligned)

    def __repr__(self):
        tmpstr = self.__class__.__name__ + "("
        tmpstr += "output_size=" + str(self.output_size)
        tmpstr += ", spatial_scale=" + str(self.spatial_scale)
        tmpstr += ", sampling_ratio=" + str(self.sampling_ratio)
        tmpstr += ", aligned=" + str(self.aligned)
        tmpstr += ")"
        return tmpstr
