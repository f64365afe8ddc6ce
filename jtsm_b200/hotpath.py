"""One pass of the per-proposal hot path, composed from the ops layer (-> C ABI -> sm_100a kernels).

Mirrors what `OICRROIHeads._forward_box` + autograd execute per step in the reference
(projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:373-468), minus the fc6/fc7 GEMMs:
   box_pooler (+ objectness scaling :391-394) -> WSDDN head math + BCE -> Kr x (pseudo-gt,
   label assignment, weighted CE, softmax) -> backward to the head logits and to the feature map.
Used by bench.py and by the parity tests; stages are enabled as their kernels exist.
"""
import torch

from . import heads, ops, synth


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
            mil_cls=wl["mil_cls"].contiguous().pin_memory(),
            mil_det=wl["mil_det"].contiguous().pin_memory(),
        )
        for k, t in enumerate(wl["ref_logits"]):
            host["ref_logits%d" % k] = t.contiguous().pin_memory()
        for k, t in enumerate(wl["ref_deltas"]):
            host["ref_deltas%d" % k] = t.contiguous().pin_memory()
        cfg = self.cfg
        self.counts = [int(b.shape[0]) for b in wl["boxes"]]
        self.labels = wl["labels"]
        oh = torch.zeros(len(self.counts), cfg["K"])
        for i, l in enumerate(self.labels):
            oh[i, l] = 1.0
        host["labels_onehot"] = oh.pin_memory()
        # gradient handed over by fc6's backward lives on the device in the real system
        host["_grad_pooled"] = synth.expand_grad_tile(wl["grad_tile"].to(self.dev), wl["rois"].shape[0]).to(self.dtype)
        return host

    def upload(self, host):
        res = {k: (v.to(self.dev, non_blocking=True) if not k.startswith("_") else v) for k, v in host.items()}
        res["row_scale"] = res["objectness"] + 1.0
        res["proposals"] = res["rois"][:, 1:].contiguous()
        res["ref_logits"] = [res["ref_logits%d" % k] for k in range(self.cfg["Kr"])]
        res["ref_deltas"] = [res["ref_deltas%d" % k] for k in range(self.cfg["Kr"])] if self.cfg["reg"] else []
        return res

    def h2d_bytes(self, host):
        return sum(v.numel() * v.element_size() for k, v in host.items() if not k.startswith("_"))

    def d2h_bytes(self):
        return self._d2h

    def stages(self):
        st = ["roi_pool_fwd(x objectness)", "wsddn_fwd+bce", "wsddn_bwd"]
        st += ["refine_assign", "refine_ce_fwd_bwd_softmax", "apply_deltas"] * self.cfg["Kr"]
        if self.cfg["reg"]:
            st += ["refine_box_reg"] * self.cfg["Kr"]
        return st + ["roi_pool_bwd"]

    def dominant_kernel(self):
        return "roi_pool_fwd_rmq_kernel"

    def ncu_traffic(self, name=None):
        """DRAM bytes per launch of the dominant kernel from the committed ncu capture of this
        configuration (profiles/traffic.json), or None when there is none."""
        import json
        import os
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
        if name is None or not os.path.exists(path):
            return None
        rec = json.load(open(path)).get(name)
        return rec["dram_bytes_per_launch"] if rec else None

    # ------------------------------------------------------------------ the step
    def run(self, d, time_dominant=False):
        x, rois = d["feat"], d["rois"]
        adt = torch.int16 if x.shape[2] * x.shape[3] < 65535 else torch.int32
        if time_dominant:
            # keep the GPU busy while the host enqueues the op, so the events bracket back-to-back device
            # work (memsets, prep kernel, pooling kernel) and not the host's launch latency
            torch.cuda._sleep(2_000_000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        pooled, argmax = ops.roi_pool_forward(x, rois, self.scale, self.P, row_scale=d["row_scale"], argmax_dtype=adt)
        if time_dominant:
            e1.record()
        cfg = self.cfg
        h = heads.head_step(d["mil_cls"], d["mil_det"], d["ref_logits"], d["ref_deltas"], d["proposals"], self.counts,
                            self.labels, cfg["K"], d["labels_onehot"], reg=cfg["reg"])
        grad_feat = ops.roi_pool_backward(d["_grad_pooled"], rois, argmax, x.shape, self.P, row_scale=d["row_scale"])
        self.last = dict(pooled=pooled, grad_feat=grad_feat, heads=h)
        return (e0, e1) if time_dominant else None

    # ------------------------------------------------------------------ CUDA graph of the step
    def capture(self, d):
        """Capture one step over the device-resident inputs `d` into a CUDA graph (all launches go
        through the C ABI on the capturing stream; allocations come from the graph's private pool).
        Afterwards `replay()` re-runs it; the outputs in `self.last` are the graph's static buffers."""
        from . import _lib

        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):  # warm caches (segment tables, function attributes) outside the capture
                self.run(d)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.run(d)
        self.launches_per_step = _lib.launch_count() - before
        self.static_inputs = d
        self.graph_last = self.last  # the graph's static output buffers
        return self.graph

    def replay(self):
        self.graph.replay()

    def results(self, last=None):
        last = last or self.last
        h = last["heads"]
        # what the host consumes per step: losses, pseudo-labels and the logging counters.  The gradient
        # w.r.t. the feature map stays on the device (it feeds the backbone's backward); a checksum of it
        # is read back so the D2H depends on the whole step.
        res = [torch.cat([v.reshape(1) for v in h["losses"].values()]), last["grad_feat"].sum().reshape(1)]
        for b in h["branches"]:
            res += [b["gt_classes"], b["counters"], b["acc"]]
        return res

    def replay_from_host(self, host):
        """H2D of the step's inputs into the graph's static buffers, replay, blocking D2H of the results."""
        d = self.static_inputs
        for k, v in host.items():
            if not k.startswith("_"):
                d[k].copy_(v, non_blocking=True)
        d["row_scale"].copy_(d["objectness"] + 1.0)
        d["proposals"].copy_(d["rois"][:, 1:])
        self.graph.replay()
        out = [t.to("cpu") for t in self.results(self.graph_last)]
        self._d2h = sum(t.numel() * t.element_size() for t in out)
        return out

    def run_from_host(self, host):
        d = self.upload(host)
        self.run(d)
        out = [t.to("cpu") for t in self.results()]  # blocking D2H: losses, pseudo-labels, counters
        self._d2h = sum(t.numel() * t.element_size() for t in out)
        return out
