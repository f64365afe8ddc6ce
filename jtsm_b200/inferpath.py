"""One pass of the multi-scale + flip WSL INFERENCE path (BASELINE.json configs[3]), composed from the ops layer.

Mirrors GeneralizedRCNNWithTTAAVG._inference_one_image (projects/WSL/wsl/modeling/test_time_augmentation_avg.py:
264-404) minus backbone and fc6 / fc7: for every augmentation (5 scales x flip)
    box_pooler x (objectness + 1)                 roi_heads_oicr.py:391-394
    [given the Kr refinement heads' logits / deltas]
    predict_probs_K / predict_boxes_K             fast_rcnn_oicr.py:712-783   (mean softmax, decoded mean deltas)
then
    _get_augmented_boxes                          :364-390  inverse transforms + mean over augmentations (one launch)
    _merge_detections -> fast_rcnn_inference_single_image   :392-404  threshold, per-class NMS, top-100 (one call)
Everything stays on the device, nothing synchronises: the whole chain is captured into one CUDA graph."""
import torch

from . import inference, ops, tta


class InferenceStep:
    def __init__(self, cfg, device, dtype=torch.float32):
        self.cfg, self.dev, self.dtype = cfg, device, dtype
        self.P = (cfg["P"], cfg["P"])
        hw = (cfg["img_h"], cfg["img_w"])
        self.tfms = tta.tta_transforms(hw, hw, cfg["min_sizes"], cfg["max_size"], cfg["flip"])
        self.inv_stages = [t.inverse().stages() for t in self.tfms]
        self.orig_hw = hw

    # ------------------------------------------------------------------ inputs
    def host_inputs(self, wl):
        """Pinned host copies of what one step uploads: per augmentation the conv5 map, RoIs, head logits / deltas."""
        host = dict(objectness=wl["objectness"].pin_memory(), augs=[])
        for a in wl["augs"]:
            host["augs"].append(dict(feat=a["feat"].to(self.dtype).pin_memory(), rois=a["rois"].pin_memory(),
                                     ref_logits=[t.pin_memory() for t in a["ref_logits"]],
                                     ref_deltas=[t.pin_memory() for t in a["ref_deltas"]]))
        return host

    def h2d_bytes(self, host):
        n = host["objectness"].numel() * 4
        for a in host["augs"]:
            n += a["feat"].numel() * a["feat"].element_size() + a["rois"].numel() * 4
            n += sum(t.numel() * 4 for t in a["ref_logits"]) + sum(t.numel() * 4 for t in a["ref_deltas"])
        return n

    def upload(self, host, into=None):
        """Host -> device (non-blocking); `into`: an existing device set to overwrite (static graph inputs)."""
        if into is None:
            d = dict(objectness=host["objectness"].to(self.dev, non_blocking=True), augs=[])
            for a in host["augs"]:
                d["augs"].append({k: ([t.to(self.dev, non_blocking=True) for t in v] if isinstance(v, list)
                                      else v.to(self.dev, non_blocking=True)) for k, v in a.items()})
            d["row_scale"] = d["objectness"] + 1.0
            return d
        into["objectness"].copy_(host["objectness"], non_blocking=True)
        for da, ha in zip(into["augs"], host["augs"]):
            for k, v in ha.items():
                if isinstance(v, list):
                    for dt, ht in zip(da[k], v):
                        dt.copy_(ht, non_blocking=True)
                else:
                    da[k].copy_(v, non_blocking=True)
        return into

    # ------------------------------------------------------------------ the step
    def stages(self):
        a = self.cfg["A"]
        return ["roi_pool_fwd(x objectness)"] * a + ["softmax_rows x Kr (mean)", "apply_deltas(mean deltas)"] * a + \
               ["tta_merge", "fast_rcnn_inference(valid rows, threshold + compaction, per-class NMS, top-k)"]

    def run(self, d):
        cfg = self.cfg
        d["row_scale"].copy_(d["objectness"] + 1.0)
        all_boxes, all_scores, pooled = [], [], []
        adt = torch.int16
        for a in d["augs"]:
            hw = a["feat"].shape[2] * a["feat"].shape[3]
            out, _ = ops.roi_pool_forward(a["feat"], a["rois"], cfg["scale"], self.P, row_scale=d["row_scale"],
                                          argmax_dtype=adt if hw < 65535 else torch.int32)
            pooled.append(out)  # -> fc6 / fc7 (torch GEMMs, outside the path; the heads' logits are given)
            all_scores.append(inference.predict_probs_K(a["ref_logits"]))
            all_boxes.append(inference.predict_boxes_K(a["ref_deltas"], a["rois"][:, 1:]))
        mb, ms = ops.tta_merge(all_boxes, all_scores, self.inv_stages)
        det = inference.fast_rcnn_inference_device(mb, ms, self.orig_hw, cfg["score_thresh"], cfg["nms_thresh"],
                                                   cfg["topk"])
        self.last = dict(det=det, mean_boxes=mb, mean_scores=ms, pooled=pooled)
        return det

    def results(self, det=None):
        """What the host consumes per image: the (padded) detections, one packed buffer."""
        det = det or self.last["det"]
        return torch.cat([det["pred_boxes"].reshape(-1), det["scores"], det["pred_classes"].to(torch.float32),
                          det["pred_inds"].to(torch.float32), det["num"].to(torch.float32)])

    def capture(self, d):
        from . import _lib

        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self.run(d)
                self.results()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.run(d)
            self.launches_per_step = _lib.launch_count() - before
            self.graph_packed = self.results()
        self.static_inputs = d
        return self.graph
