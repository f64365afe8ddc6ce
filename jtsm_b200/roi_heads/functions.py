"""`torch.autograd.Function`s over the head kernels, so that the `cls` / `det` / refinement `Linear` layers train
through them with a plain `loss.backward()`.

The reference builds these stages from eager element-wise ops and lets autograd differentiate each of them
(fast_rcnn_wsddn.py:571-584, 346-379; fast_rcnn_oicr.py:243-298, 300-380); here every stage is one launch forward
and one launch (or none) backward:

  wsddn(cls, det)                 -> scores, unclamped image sums, clamped image scores [, BCE loss]     jtsm_wsddn_forward
      backward(d scores, d image sums, d loss)                                                          jtsm_wsddn_backward
  refine_ce(logits | labels, w)   -> loss, softmax probabilities, accuracy counters                     jtsm_refine_ce
      backward: the kernel already produced d loss / d logits in the forward launch; scaled by the upstream gradient
  refine_box_reg(deltas | ...)    -> loss                                                               jtsm_refine_box_reg
      backward: likewise
No double backward (the reference's ops are `once_differentiable` too, roi_align.py:36)."""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import ops


class _WSDDN(Function):
    @staticmethod
    def forward(ctx, cls_logits, det_logits, counts, labels_onehot, mean_loss):
        fwd = ops.wsddn_forward(cls_logits, det_logits, counts, labels_onehot, mean_loss)
        ctx.set_materialize_grads(False)
        ctx.n = fwd["n"]
        ctx.in_dtypes = (cls_logits.dtype, det_logits.dtype)
        ctx.has_loss = fwd["loss"] is not None
        saved = [fwd["cls"], fwd["det"], fwd["scores"], fwd["col_stats"], fwd["seg"]]
        if ctx.has_loss:
            saved.append(fwd["grad_img"])
        ctx.save_for_backward(*saved)
        img_sum = fwd["col_stats"][:, :, 2].contiguous()
        loss = fwd["loss"] if ctx.has_loss else torch.zeros(1, dtype=torch.float32, device=fwd["scores"].device)
        ctx.mark_non_differentiable(fwd["img_scores"])
        return fwd["scores"], img_sum, fwd["img_scores"], loss

    @staticmethod
    @once_differentiable
    def backward(ctx, g_scores, g_img_sum, _g_img_scores, g_loss):
        t = ctx.saved_tensors
        fwd = dict(cls=t[0], det=t[1], scores=t[2], col_stats=t[3], seg=t[4], n=ctx.n)
        gi = None
        if ctx.has_loss and g_loss is not None:
            gi = t[5] * g_loss.to(torch.float32).reshape(1, 1)
        if g_img_sum is not None:
            gi = g_img_sum.to(torch.float32) if gi is None else gi + g_img_sum
        if gi is None and g_scores is None:
            return None, None, None, None, None
        dc, dd = ops.wsddn_backward(fwd, grad_scores=g_scores, grad_img=gi)
        return dc.to(ctx.in_dtypes[0]), dd.to(ctx.in_dtypes[1]), None, None, None


def wsddn(cls_logits, det_logits, num_preds_per_image, labels_onehot=None, mean_loss=True):
    """-> (scores [R,K], img_sum [N,K] unclamped, img_scores [N,K] clamped & detached, loss [1] or zeros)."""
    return _WSDDN.apply(cls_logits, det_logits, tuple(int(c) for c in num_preds_per_image), labels_onehot, bool(mean_loss))


class _RefineCE(Function):
    @staticmethod
    def forward(ctx, logits, gt_classes, loss_weights, valid_count, loss_scale):
        need = ctx.needs_input_grad[0]
        r = ops.refine_ce(logits, gt_classes, loss_weights, valid_count, loss_scale=loss_scale, want_probs=True,
                          want_grad=need, want_acc=True)
        ctx.set_materialize_grads(False)
        ctx.in_dtype = logits.dtype
        ctx.save_for_backward(r["grad_logits"] if need else None)
        ctx.mark_non_differentiable(r["probs"], r["acc"])
        return r["loss"], r["probs"], r["acc"]

    @staticmethod
    @once_differentiable
    def backward(ctx, g_loss, _gp, _ga):
        (g,) = ctx.saved_tensors
        if g is None or g_loss is None:
            return None, None, None, None, None
        return (g * g_loss.to(torch.float32).reshape(1, 1)).to(ctx.in_dtype), None, None, None, None


def refine_ce(logits, gt_classes, loss_weights, valid_count, loss_scale=1.0):
    """Weighted softmax CE (fast_rcnn_oicr.py:282-298) -> (loss [1], probs [R,K+1] detached, acc counters int32 [4])."""
    return _RefineCE.apply(logits, gt_classes, loss_weights, valid_count, float(loss_scale))


class _RefineBoxReg(Function):
    @staticmethod
    def forward(ctx, pred_deltas, proposals, gt_boxes, gt_classes, loss_weights, num_classes, weights, beta, loss_scale):
        need = ctx.needs_input_grad[0]
        r = ops.refine_box_reg(pred_deltas, proposals, gt_boxes, gt_classes, loss_weights, num_classes, weights, beta,
                               loss_scale=loss_scale, want_grad=need)
        ctx.set_materialize_grads(False)
        ctx.in_dtype = pred_deltas.dtype
        ctx.save_for_backward(r["grad_deltas"] if need else None)
        return r["loss"]

    @staticmethod
    @once_differentiable
    def backward(ctx, g_loss):
        (g,) = ctx.saved_tensors
        if g is None or g_loss is None:
            return (None,) * 9
        return ((g * g_loss.to(torch.float32).reshape(1, 1)).to(ctx.in_dtype),) + (None,) * 8


def refine_box_reg(pred_deltas, proposals, gt_boxes, gt_classes, loss_weights, num_classes,
                   weights=(10.0, 10.0, 5.0, 5.0), smooth_l1_beta=0.0, loss_scale=1.0):
    """OICROutputs.box_reg_loss (fast_rcnn_oicr.py:300-380; loss_weights=None -> plain "smooth_l1") -> loss [1]."""
    return _RefineBoxReg.apply(pred_deltas, proposals, gt_boxes, gt_classes, loss_weights, int(num_classes),
                               tuple(float(w) for w in weights), float(smooth_l1_beta), float(loss_scale))
