"""One pass of the per-proposal hot path, composed from the ops layer (-> C ABI -> sm_100a kernels).

Mirrors what `OICRROIHeads._forward_box` + autograd execute per step in the reference
(projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:373-468), minus the fc6/fc7 GEMMs:
   box_pooler (+ objectness scaling :391-394) -> WSDDN head math + BCE -> Kr x (pseudo-gt,
   label assignment, weighted CE, softmax) -> backward to the head logits and to the feature map.
Used by bench.py and by the parity tests; stages are enabled as their kernels exist.
"""
import os

import torch

from . import heads, ops, synth


class HotPathStep:
    def __init__(self, cfg, device, dtype=torch.float32):
        self.cfg = cfg
        self.dev = device
        self.dtype = dtype
        self.P = (cfg["P"], cfg["P"])
        self.scale = cfg["scale"]
        self.pooler = cfg.get("pooler", "ROIPool")
        if self.pooler not in ("ROIPool", "MOIPool"):
            raise RuntimeError("HotPathStep: pooler %r not wired into the step (ROIPool, MOIPool)" % self.pooler)
        self._d2h = 0
        # side streams of the head: one per refinement branch + one for the MIL backward (heads.head_step)
        self.side_streams = [torch.cuda.Stream(device=device) for _ in range(cfg["Kr"] + 1)] \
            if os.environ.get("JTSM_SERIAL_HEADS", "0") != "1" else None

    # ------------------------------------------------------------------ inputs
    def pin_host_inputs(self, wl):
        """Host-side inputs of one step as views of ONE pinned arena (host["__arena"]), so a step's upload is a
        single copy instead of ten small ones."""
        cfg = self.cfg
        self.counts = [int(b.shape[0]) for b in wl["boxes"]]
        self.labels = wl["labels"]
        oh = torch.zeros(len(self.counts), cfg["K"])
        for i, l in enumerate(self.labels):
            oh[i, l] = 1.0
        # image-level class lists (gt_classes_img_int) as device data: offsets [N+1] + ids padded to N * cap, so a
        # captured step serves batches with different labels (the pseudo-gt seeds follow the uploaded lists)
        n_img = len(self.counts)
        self.cls_cap = min(cfg["K"], int(cfg.get("max_img_classes", 16)))
        offs, ids = [0], []
        for l in self.labels:
            l = [int(v) for v in (l.tolist() if torch.is_tensor(l) else l)]
            if len(l) > self.cls_cap:
                raise RuntimeError("HotPathStep: %d image-level classes > max_img_classes=%d" % (len(l), self.cls_cap))
            ids.extend(l)
            offs.append(len(ids))
        ids += [0] * (n_img * self.cls_cap - len(ids))
        src = dict(feat=wl["feat"].to(self.dtype), rois=wl["rois"], objectness=wl["objectness"], mil_cls=wl["mil_cls"],
                   mil_det=wl["mil_det"], labels_onehot=oh, cls_off=torch.tensor(offs, dtype=torch.int32),
                   cls_ids=torch.tensor(ids, dtype=torch.int64))
        if self.pooler == "MOIPool":
            src["oh_labels"] = wl["oh_labels"]
            src["superpixels"] = wl["superpixels"]
        for k, t in enumerate(wl["ref_logits"]):
            src["ref_logits%d" % k] = t
        for k, t in enumerate(wl["ref_deltas"]):
            src["ref_deltas%d" % k] = t
        self._layout, off = [], 0
        for k, t in src.items():
            nbytes = t.numel() * t.element_size()
            self._layout.append((k, off, nbytes, t.dtype, tuple(t.shape)))
            off = (off + nbytes + 255) & ~255
        arena = torch.empty(off, dtype=torch.uint8).pin_memory()
        host = self._views(arena)
        for k, t in src.items():
            host[k].copy_(t)
        host["__arena"] = arena
        # gradient handed over by fc6's backward lives on the device in the real system
        host["_grad_pooled"] = synth.expand_grad_tile(wl["grad_tile"].to(self.dev), wl["rois"].shape[0]).to(self.dtype)
        return host

    def _views(self, arena):
        return {k: arena[off:off + nbytes].view(dt).view(shape) for k, off, nbytes, dt, shape in self._layout}

    def upload(self, host):
        arena = host["__arena"].to(self.dev, non_blocking=True)
        res = self._views(arena)
        res["__arena"] = arena
        res["_grad_pooled"] = host["_grad_pooled"]
        res["row_scale"] = res["objectness"] + 1.0
        res["proposals"] = res["rois"][:, 1:].contiguous()
        res["ref_logits"] = [res["ref_logits%d" % k] for k in range(self.cfg["Kr"])]
        res["ref_deltas"] = [res["ref_deltas%d" % k] for k in range(self.cfg["Kr"])] if self.cfg["reg"] else []
        return res

    def h2d_bytes(self, host):
        return sum(nbytes for _, _, nbytes, _, _ in self._layout)

    def d2h_bytes(self):
        return self._d2h

    def stages(self):
        pool = "roi_pool_fwd(x objectness)" if self.pooler == "ROIPool" else \
            "moi_pool_fwd(cell ids, row bits, visit lists, pool x mask_scale x objectness)"
        st = [pool, "wsddn_fwd+bce", "wsddn_bwd"]
        st += ["refine_assign", "refine_ce_fwd_bwd_softmax", "apply_deltas"] * self.cfg["Kr"]
        if self.cfg["reg"]:
            st += ["refine_box_reg"] * self.cfg["Kr"]
        return st + ["roi_pool_bwd" if self.pooler == "ROIPool" else "moi_pool_bwd(x scales)"]

    def schedule(self):
        """How the stages are ordered on the device."""
        if self.side_streams is None:
            return "one stream, stages in order"
        return ("pool fwd -> MIL fwd -> {branch k: softmax(prev logits), assign, CE[, box reg]} x Kr and MIL bwd "
                "concurrently on side streams -> pool bwd")

    def dominant_kernel(self):
        return "roi_pool_fwd_rmq_kernel" if self.pooler == "ROIPool" else "moi_pool_fwd_list_kernel"

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
    def pool_forward(self, d):
        """-> (pooled, argmax, row scale the backward multiplies by)."""
        x = d["feat"]
        if self.pooler == "MOIPool":  # JTSM: roi_heads_jtsm.py:607-633
            out, arg, scales = ops.moi_pool_forward(x, d["rois"], self.scale, self.P[0], self.P[1], d["oh_labels"],
                                                    d["superpixels"], row_scale=d["row_scale"], mask_scale=True,
                                                    return_scales=True)
            return out, arg, scales[2]
        adt = torch.int16 if x.shape[2] * x.shape[3] < 65535 else torch.int32
        out, arg = ops.roi_pool_forward(x, d["rois"], self.scale, self.P, row_scale=d["row_scale"], argmax_dtype=adt)
        return out, arg, d["row_scale"]

    def capture_dominant(self, d):
        """The dominant op (RoIPool forward: two memsets, prep kernel, pooling kernel) alone in a CUDA graph, so it
        can be timed with events without host launch gaps, exactly as it runs inside the step's graph."""
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self.pool_forward(d)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.dom_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.dom_graph):
            self._dom_out = self.pool_forward(d)

    def time_dominant(self):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        self.dom_graph.replay()
        e1.record()
        return e0, e1

    def run_forward(self, d):
        """pool fwd + the whole head (everything up to the point where fc6's backward would hand back its gradient)."""
        pooled, argmax, bwd_scale = self.pool_forward(d)
        cfg = self.cfg
        h = heads.head_step(d["mil_cls"], d["mil_det"], d["ref_logits"], d["ref_deltas"], d["proposals"], self.counts,
                            (d["cls_off"], d["cls_ids"], self.cls_cap, len(self.counts) * self.cls_cap), cfg["K"],
                            d["labels_onehot"], reg=cfg["reg"], streams=self.side_streams)
        return dict(pooled=pooled, argmax=argmax, bwd_scale=bwd_scale, heads=h)

    def run_backward(self, d, st):
        grad_feat = ops.roi_pool_backward(d["_grad_pooled"], d["rois"], st["argmax"], d["feat"].shape, self.P,
                                          row_scale=st["bwd_scale"])
        self.last = dict(pooled=st["pooled"], grad_feat=grad_feat, heads=st["heads"])

    def run(self, d):
        self.run_backward(d, self.run_forward(d))

    # ------------------------------------------------------------------ CUDA graph of the step
    def derive_inputs(self, d):
        """Inputs the step derives from the uploaded ones (row scale = objectness + 1, boxes = RoI columns 1..4)."""
        d["row_scale"].copy_(d["objectness"] + 1.0)
        d["proposals"].copy_(d["rois"][:, 1:])

    def capture(self, d, with_io=False):
        """Capture one step over the device-resident inputs `d` into a CUDA graph (all launches go
        through the C ABI on the capturing stream; allocations come from the graph's private pool).
        Afterwards `replay()` re-runs it; the outputs in `self.last` are the graph's static buffers.
        with_io: the graph also derives the dependent inputs first and packs the host-bound results last
        (`self.graph_results`), so a streamed step is one copy in, one replay, a few copies out."""
        from . import _lib

        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):  # warm caches (segment tables, function attributes) outside the capture
                if with_io:
                    self.derive_inputs(d)
                self.run(d)
                if with_io:
                    self.results()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            if with_io:
                self.derive_inputs(d)
            self.run(d)
            self.launches_per_step = _lib.launch_count() - before
            self.graph_results = self.results() if with_io else None
            # all host-bound results in ONE byte buffer: one D2H copy per step instead of one per tensor
            self.graph_packed = self._pack(self.graph_results) if with_io else None
        self.static_inputs = d
        self.graph_last = self.last  # the graph's static output buffers
        return self.graph

    def capture_split(self, d):
        """The streamed step as TWO graphs over the same buffers: A = derive inputs + pool forward + heads, B = pool
        backward + result packing.  The streaming loop records an event between them so that uploads can be held back
        until the pooling forward -- the one kernel a concurrent H2D copy slows down (221 -> 260 us at c2: the DMA
        writes disturb the L2 write-combining its output relies on) -- has finished."""
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self.derive_inputs(d)
                self.run(d)
                self.results()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(ga):
            self.derive_inputs(d)
            st = self.run_forward(d)
        with torch.cuda.graph(gb, pool=ga.pool()):
            self.run_backward(d, st)
            self.graph_results = self.results()
            self.graph_packed = self._pack(self.graph_results)
        self.graph_last = self.last
        return ga, gb

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

    def _pack(self, res):
        self._res_layout, off = [], 0
        for t in res:
            nbytes = t.numel() * t.element_size()
            self._res_layout.append((off, nbytes, t.dtype, tuple(t.shape)))
            off += (nbytes + 15) & ~15
        parts = []
        for t, (o, nbytes, _, _) in zip(res, self._res_layout):
            parts.append(t.contiguous().reshape(-1).view(torch.uint8))
            pad = ((nbytes + 15) & ~15) - nbytes
            if pad:
                parts.append(torch.empty(pad, dtype=torch.uint8, device=t.device))  # alignment only
        return torch.cat(parts)

    def _unpack(self, buf):
        return [buf[o:o + n].view(dt).view(shape) for o, n, dt, shape in self._res_layout]

    def replay_from_host(self, host):
        """H2D of the step's inputs into the graph's static buffers, replay, blocking D2H of the results."""
        d = self.static_inputs
        d["__arena"].copy_(host["__arena"], non_blocking=True)
        self.derive_inputs(d)
        self.graph.replay()
        out = [t.to("cpu") for t in self.results(self.graph_last)]
        self._d2h = sum(t.numel() * t.element_size() for t in out)
        return out

    # ------------------------------------------------------------------ pipelined stream of steps from host memory
    def prepare_stream(self, host, depth=None, overlap=None):
        """`depth` captured copies of the step (sets of static device buffers) + pinned result buffers, so that
        step i + 1's inputs upload and step i - 1's results download while step i computes.
        overlap: "any" (default; JTSM_STREAM_OVERLAP) = the upload of step i + 1 starts as soon as the host has queued
        it (depth: JTSM_STREAM_DEPTH or 2; 3 and 4 measured the same); "after_forward" = three buffer sets, each step
        captured as two graphs and every upload cut in two halves that only run between a pooling forward's end and
        the next one's start.  A concurrent H2D copy slows the pooling forward by 17 % when the two run back to back
        (tools/probe_dma.py), but holding the uploads back measured the same end-to-end rate (3.36 M RoIs/s either
        way at c2), so the simpler schedule stays the default."""
        self.overlap = overlap or os.environ.get("JTSM_STREAM_OVERLAP", "any")
        if self.overlap == "after_forward":
            depth = 3
        else:
            depth = int(depth or os.environ.get("JTSM_STREAM_DEPTH", "2"))
        self.slots = []
        for _ in range(max(2, depth)):
            d = self.upload(host)
            torch.cuda.synchronize()
            if self.overlap == "after_forward":
                ga, gb = self.capture_split(d)
                graph = None
            else:
                graph, ga, gb = self.capture(d, with_io=True), None, None
            res = self.graph_packed
            out_buf = torch.empty(res.shape, dtype=torch.uint8).pin_memory()
            self.slots.append(dict(graph=graph, graph_a=ga, graph_b=gb, inputs=d, last=self.graph_last, res=res,
                                   out_buf=out_buf, out=self._unpack(out_buf),
                                   ev_in=torch.cuda.Event(), ev_done=torch.cuda.Event(), ev_out=torch.cuda.Event(),
                                   ev_fwd=torch.cuda.Event()))
        self.copy_stream, self.d2h_stream = torch.cuda.Stream(), torch.cuda.Stream()
        self._d2h = self.slots[0]["out_buf"].numel()
        n = self.slots[0]["inputs"]["__arena"].numel()
        self._half = (n // 2 + 255) & ~255

    def run_stream(self, host_batches, consume=None, before_step=None):
        """Run one step per element of `host_batches` (dicts of pinned host tensors as from pin_host_inputs).
        Per step: H2D of its inputs on a copy stream, graph replay on the current stream, D2H of its results
        (losses, pseudo-labels, counters) into pinned buffers on a third stream; `consume(i, results)` is called on
        the host once step i's results have landed (len(slots) - 1 steps behind the submission)."""
        if self.overlap == "after_forward":
            return self._run_stream_after_forward(list(host_batches), consume, before_step)
        comp = torch.cuda.current_stream()
        self.copy_stream.wait_stream(comp)
        n = 0
        ns = len(self.slots)
        lag = ns - 1
        for i, host in enumerate(host_batches):
            sl = self.slots[i % ns]
            with torch.cuda.stream(self.copy_stream):
                if i >= ns:
                    self.copy_stream.wait_event(sl["ev_done"])  # the slot's previous step no longer reads its inputs
                sl["inputs"]["__arena"].copy_(host["__arena"], non_blocking=True)  # one copy: all inputs of the step
                sl["ev_in"].record(self.copy_stream)
            comp.wait_event(sl["ev_in"])
            if i >= ns:
                comp.wait_event(sl["ev_out"])  # the slot's previous results have left the device
            if before_step is not None:
                before_step()
            sl["graph"].replay()  # derive inputs, the step, pack the results: all inside the slot's graph
            sl["ev_done"].record(comp)
            self._download(sl)
            if i >= lag:
                prev = self.slots[(i - lag) % ns]
                prev["ev_out"].synchronize()
                if consume is not None:
                    consume(i - lag, prev["out"])
            n = i + 1
        self._drain(n, lag, consume)
        return n

    def _download(self, sl):
        with torch.cuda.stream(self.d2h_stream):
            self.d2h_stream.wait_event(sl["ev_done"])
            sl["out_buf"].copy_(sl["res"], non_blocking=True)  # one copy: every host-bound result of the step
            sl["ev_out"].record(self.d2h_stream)

    def _drain(self, n, lag, consume):
        ns = len(self.slots)
        for j in range(max(0, n - lag), n):  # the last results, in order
            sl = self.slots[j % ns]
            sl["ev_out"].synchronize()
            if consume is not None:
                consume(j, sl["out"])
        if n:
            torch.cuda.current_stream().wait_event(self.slots[(n - 1) % ns]["ev_out"])  # covers the last D2H

    def _run_stream_after_forward(self, hosts, consume, before_step):
        """Uploads only between the end of a pooling forward and the start of the next one.  Step j's arena goes up in
        two halves: the first after the forward of step j - 2, the second after the forward of step j - 1 (each half
        takes ~0.15 ms at c2 against the ~0.3 ms of heads + pooling backward it hides behind)."""
        comp = torch.cuda.current_stream()
        cs = self.copy_stream
        cs.wait_stream(comp)
        ns, lag, half = 3, 2, self._half
        total = len(hosts)

        def up(j, part):
            sl = self.slots[j % ns]
            dst, src = sl["inputs"]["__arena"], hosts[j]["__arena"]
            if part == 0:
                dst[:half].copy_(src[:half], non_blocking=True)
            else:
                dst[half:].copy_(src[half:], non_blocking=True)
                sl["ev_in"].record(cs)

        with torch.cuda.stream(cs):  # start-up: step 0 in full, the first half of step 1
            if total > 0:
                up(0, 0)
                up(0, 1)
            if total > 1:
                up(1, 0)
        n = 0
        for i in range(total):
            sl = self.slots[i % ns]
            comp.wait_event(sl["ev_in"])
            if i >= ns:
                comp.wait_event(sl["ev_out"])  # the slot's previous results have left the device
            if before_step is not None:
                before_step()
            sl["graph_a"].replay()  # derive inputs, pooling forward, heads
            sl["ev_fwd"].record(comp)
            with torch.cuda.stream(cs):
                # behind this step's forward: the second half of step i + 1 and the first half of step i + 2 (its buffer
                # set was last read by step i - 1, which ended before this step began)
                cs.wait_event(sl["ev_fwd"])
                if i + 1 < total:
                    up(i + 1, 1)
                if i + 2 < total:
                    up(i + 2, 0)
            sl["graph_b"].replay()  # pooling backward, result packing
            sl["ev_done"].record(comp)
            self._download(sl)
            if i >= lag:
                prev = self.slots[(i - lag) % ns]
                prev["ev_out"].synchronize()
                if consume is not None:
                    consume(i - lag, prev["out"])
            n = i + 1
        self._drain(n, lag, consume)
        return n

    def run_from_host(self, host):
        d = self.upload(host)
        self.run(d)
        out = [t.to("cpu") for t in self.results()]  # blocking D2H: losses, pseudo-labels, counters
        self._d2h = sum(t.numel() * t.element_size() for t in out)
        return out
