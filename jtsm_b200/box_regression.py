"""`Box2BoxTransform` with the reference's interface (detectron2/modeling/box_regression.py:21-109) on the sm_100a
kernels: `apply_deltas` -> jtsm_apply_deltas; `get_deltas` stays a handful of eager element-wise ops (it is only
called outside the step: the training path computes the regression targets inside jtsm_refine_box_reg)."""
import math

import torch

from . import ops

_DEFAULT_SCALE_CLAMP = math.log(1000.0 / 16)


class Box2BoxTransform:
    def __init__(self, weights, scale_clamp=_DEFAULT_SCALE_CLAMP):
        self.weights = tuple(float(w) for w in weights)
        self.scale_clamp = float(scale_clamp)

    def get_deltas(self, src_boxes, target_boxes):
        """(dx, dy, dw, dh) that map src_boxes onto target_boxes; box_regression.py:38-71, operation for operation."""
        assert isinstance(src_boxes, torch.Tensor), type(src_boxes)
        assert isinstance(target_boxes, torch.Tensor), type(target_boxes)
        sw = src_boxes[:, 2] - src_boxes[:, 0]
        sh = src_boxes[:, 3] - src_boxes[:, 1]
        scx = src_boxes[:, 0] + 0.5 * sw
        scy = src_boxes[:, 1] + 0.5 * sh
        tw = target_boxes[:, 2] - target_boxes[:, 0]
        th = target_boxes[:, 3] - target_boxes[:, 1]
        tcx = target_boxes[:, 0] + 0.5 * tw
        tcy = target_boxes[:, 1] + 0.5 * th
        wx, wy, ww, wh = self.weights
        deltas = torch.stack((wx * (tcx - scx) / sw, wy * (tcy - scy) / sh, ww * torch.log(tw / sw),
                              wh * torch.log(th / sh)), dim=1)
        assert (sw > 0).all().item(), "Input boxes to Box2BoxTransform are not valid!"
        return deltas

    def apply_deltas(self, deltas, boxes):
        """[R, k*4] deltas on [R, 4] boxes -> [R, k*4] boxes (box_regression.py:73-109)."""
        if deltas.shape[0] == 0:
            return torch.zeros_like(deltas, dtype=torch.float32)
        return ops.apply_deltas(deltas, boxes, self.weights, self.scale_clamp)
