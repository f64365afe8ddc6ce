"""PCLLoss: drop-in for `wsl.layers.PCLLoss` / `pcl_loss` (projects/WSL/wsl/layers/pcl_loss.py:8-124).
Same call signature and autograd contract (gradient only w.r.t. the probabilities, once-differentiable; like the
reference the backward does not apply the upstream gradient).  The reference moves the probabilities to the CPU and
back on every forward and backward; here everything stays on the device."""
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import ops


class _PCLLoss(Function):
    @staticmethod
    def forward(ctx, pcl_probs, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                img_cls_loss_weights, im_labels):
        dev = pcl_probs.device
        small = [t.to(dev) for t in (labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                                     img_cls_loss_weights, im_labels)]
        probs = pcl_probs.detach()
        ctx.save_for_backward(probs, *small)
        loss, _ = ops.pcl_loss_forward(probs, small[0], small[1], small[3], small[4], small[6], small[7])
        return loss.reshape(())

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        probs, labels, clw, gta, pcl, pcp, pcc, iclw, iml = ctx.saved_tensors
        grad = ops.pcl_loss_backward(probs, labels, clw, gta, pcl, pcp, pcc, iclw, iml)
        return grad, None, None, None, None, None, None, None, None


pcl_loss = _PCLLoss.apply


class PCLLoss(nn.Module):
    def forward(self, pcl_prob, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                img_cls_loss_weights, im_labels_real):
        return pcl_loss(pcl_prob, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                        img_cls_loss_weights, im_labels_real)

    def __repr__(self):
        return self.__class__.__name__ + "()"
