"""ROILabel: drop-in for `wsl.layers.ROILabel` (projects/WSL/wsl/layers/roi_label.py:10-82), the CMIL / UWSOD heads'
capped foreground / background labelling.

Same constructor, call signature `(S, U, L, CW) -> (RL, RW)` (both non-differentiable) and the same 14-float state `P`.
The reference op is CPU only: its callers move the scores, the R x R IoU matrix, the image labels and class weights to the
host every branch (roi_heads_cmil.py:419-431); here it is one single-CTA launch on the device
(`jtsm_roi_label_forward`), no host read.  The order in which RoIs claim the capped slots -- std::random_shuffle seeded
with time(0) in the reference -- is drawn with `torch.randperm` on the device (`generator=` for reproducible runs)."""
import random

import torch
from torch import nn

from .. import ops


class ROILabel(nn.Module):
    def __init__(self, fg_threshold, bg_thresh_hi, bg_thresh_lo, num_pos, num_neg, top_k, debug_info, display,
                 generator=None):
        super().__init__()
        state = torch.zeros(14, dtype=torch.float)
        state[:9] = torch.tensor([fg_threshold, bg_thresh_hi, bg_thresh_lo, num_pos, num_neg, top_k, debug_info,
                                  random.randint(1, 9999), display], dtype=torch.float)
        self.P = nn.Parameter(state, requires_grad=False)
        self.generator = generator
        self._cfg = None

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self._cfg = None

    @torch.no_grad()
    def forward(self, S, U, L, CW, perm=None):
        if self._cfg is None:  # thresholds / caps are fixed after construction: one host read
            self._cfg = self.P.tolist()
            self._iter = int(self._cfg[9])
        fg, hi, lo, num_pos, num_neg, top_k, _, uuid, display = self._cfg[:9]
        dev = S.device
        if perm is None:
            perm = torch.randperm(S.shape[0], device=dev, generator=self.generator)
        RL, RW, stats = ops.roi_label_forward(S, U, L, CW, perm, fg, hi, lo, int(num_pos), int(num_neg), int(top_k))
        self._iter += 1
        self.P[9] += 1
        self.P[10:14] += stats.to(self.P.device)
        if display > 0 and self._iter % int(display) == 0:
            acc = self.P[10:14].tolist()
            print("RoILabel %d\tfg_rois: %d\tbg_rois: %d\tfg_weight: %f\tbg_weight: %f"
                  % (int(uuid), int(acc[0]) // int(display), int(acc[1]) // int(display),
                     acc[2] / max(acc[0], 1.0), acc[3] / max(acc[1], 1.0)))
            self.P[10:14] = 0
        return RL, RW

    def __repr__(self):
        names = ("fg_threshold", "bg_thresh_hi", "bg_thresh_lo", "num_pos", "num_neg", "top_k", "debug_info", "uuid",
                 "display", "cur_iter", "acc_fg_rois", "acc_bg_rois", "acc_fg_weight", "acc_bg_weight")
        return self.__class__.__name__ + "(" + " ".join("%s=%s" % (n, v) for n, v in zip(names, self.P.tolist())) + ")"
