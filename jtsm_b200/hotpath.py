"""One pass of the per-proposal hot path, composed from the ops layer (-> C ABI -> sm_100a kernels).

Mirrors what `OICRROIHeads._forward_box` + autograd execute per step in the reference
(projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:373-468), minus the fc6/fc7 GEMMs:
   box_pooler (+ objectness scaling :391-394) -> WSDDN head math + BCE -> Kr x (pseudo-gt,
   label assignment, weighted CE, softmax) -> backward to the head logits and to the feature map.
Used by bench.py and by the parity tests; stages are enabled as their kernels exist.
"""
import torch

from . import ops, synth


class HotPathStep:
    def __init__(self, cfg, device, dtype=torch.float32):
        self.cfg = cfg
        self.dev = device
        self.dtype = dtype
        self.P = (cfg["P"], cfg["P"])
        self.scale = cfg["scale"]
        self._d2h = 0

    # ------------------------------------------------------------------ inputs
    def pin_host_inputs(self, wl):
        host = dict(
            feat=wl["feat"].to(self.dtype).contiguous().pin_memory(),
            rois=wl["rois"].contiguous().pin_memory(),
            objectness=wl["objectness"].contiguous().pin_memory(),
        )
        # gradient handed over by fc6's backward lives on the device in the real system
        host["_grad_pooled"] = synth.expand_grad_tile(wl["grad_tile"].to(self.dev), wl["rois"].shape[0]).to(self.dtype)
        return host

    def upload(self, host):
        res = {k: (v.to(self.dev, non_blocking=True) if not k.startswith("_") else v) for k, v in host.items()}
        res["row_scale"] = res["objectness"] + 1.0
        return res

    def h2d_bytes(self, host):
        return sum(v.numel() * v.element_size() for k, v in host.items() if not k.startswith("_"))

    def d2h_bytes(self):
        return self._d2h

    def stages(self):
        return ["roi_pool_fwd(x objectness)", "roi_pool_bwd"]

    def dominant_kernel(self):
        return "roi_pool_fwd_kernel"

    def ncu_traffic(self):
        return None

    # ------------------------------------------------------------------ the step
    def run(self, d, time_dominant=False):
        x, rois = d["feat"], d["rois"]
        adt = torch.int16 if x.shape[2] * x.shape[3] < 65535 else torch.int32
        if time_dominant:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        pooled, argmax = ops.roi_pool_forward(x, rois, self.scale, self.P, row_scale=d["row_scale"], argmax_dtype=adt)
        if time_dominant:
            e1.record()
        grad_feat = ops.roi_pool_backward(d["_grad_pooled"], rois, argmax, x.shape, self.P, row_scale=d["row_scale"])
        self.last = dict(pooled=pooled, grad_feat=grad_feat)
        return (e0, e1) if time_dominant else None

    def run_from_host(self, host):
        d = self.upload(host)
        self.run(d)
        out = self.last["grad_feat"].to("cpu")
        self._d2h = out.numel() * out.element_size()
        return out
