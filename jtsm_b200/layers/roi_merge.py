"""ROIMerge: drop-in for `wsl.layers.ROIMerge` (projects/WSL/wsl/layers/roi_merge.py:10-66), the CMIL head's clique merge.

Same constructor, call signature `(S, J, C, D) -> (MC, MD)`, autograd contract (gradients w.r.t. C and D only) and the
same 8-int state `P` (debug_info, display, cur_iter, max_epoch, size_epoch + three display accumulators) saved with the
module.  The reference op is CPU only and its caller moves S, the R x R IoU matrix J, C and D to the host on every
iteration (fast_rcnn_wsddn.py:666); here everything runs on the device (`jtsm_roi_merge_forward`).  The number of clusters
is data dependent, so returning `[num_id, K]` tensors costs ONE 4-byte device -> host read per call;
`roi_merge_padded` returns the zero-padded `[R, K]` form plus the device-side count with no host read at all."""
import torch
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import ops


class _ROIMerge(Function):
    @staticmethod
    def forward(ctx, S, J, C, D, lam, trim):
        out = ops.roi_merge_forward(S.detach(), J.detach(), C.detach(), D.detach(), lam)
        num_id = int(out["stats"][0].item()) if trim else C.shape[0]
        ctx.save_for_backward(out["I"], out["IC"])
        ctx.rows = (C.shape[0], num_id)
        ctx.mark_non_differentiable(out["I"], out["IC"], out["stats"])
        return out["MC"][:num_id], out["MD"][:num_id], out["I"], out["IC"], out["stats"]

    @staticmethod
    @once_differentiable
    def backward(ctx, GMC, GMD, _gi, _gic, _gs):
        I, IC = ctx.saved_tensors
        r, _ = ctx.rows
        GC, GD = ops.roi_merge_backward(GMC, GMD, I, IC, r)
        return None, None, GC, GD, None, None


def roi_merge_padded(S, J, C, D, lam):
    """-> (MC [R,K], MD [R,K] zero rows past the cluster count, I, IC, stats int32[4]) -- no host read."""
    return _ROIMerge.apply(S, J, C, D, lam, False)


class ROIMerge(nn.Module):
    def __init__(self, debug_info, display, max_epoch, size_epoch):
        super().__init__()
        state = torch.zeros(8, dtype=torch.int)
        state[0], state[1], state[3], state[4] = debug_info, display, max_epoch, size_epoch
        self.P = nn.Parameter(state, requires_grad=False)
        self._iter = None  # host mirror of P[2]; read once (checkpoint loads change P behind our back)

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self._iter = None

    def forward(self, S, J, C, D):
        if self._iter is None:
            self._iter = int(self.P[2].item())
            self._cfg = [int(v) for v in self.P[:5].tolist()]
        _, display, _, max_epoch, size_epoch = self._cfg
        lam = ops.roi_merge_lambda(self._iter, size_epoch, max_epoch)
        MC, MD, _, _, stats = _ROIMerge.apply(S, J, C, D, lam, True)
        self._iter += 1
        with torch.no_grad():  # device-side bookkeeping, no host read
            self.P[2] += 1
            self.P[5:8] += stats[1:4].to(self.P.device)
            if display > 0 and self._iter % display == 0:
                acc = self.P[5:8].tolist()
                print("RoIMerge %d\tlambda: %f\tacc_top_num_id: %d\tacc_max_clique: %d\tacc_min_clique: %d"
                      % (self._iter, lam, acc[0] // display, acc[1] // display, acc[2] // display))
                self.P[5:8] = 0
        return MC, MD

    def __repr__(self):
        p = self.P.tolist()
        names = ("debug_info", "display", "cur_iter", "max_epoch", "size_epoch", "acc_num_top_id", "acc_max_clique",
                 "acc_min_clique")
        return self.__class__.__name__ + "(" + " ".join("%s=%d" % (n, v) for n, v in zip(names, p)) + ")"
