"""CPU leg of bench.py for the c4 INFERENCE chain (TEST / BASELINE INFRASTRUCTURE -- never imported by jtsm_b200/).

What the reference executes with MODEL.DEVICE=cpu for this path, per image
(test_time_augmentation_avg.py:264-404): per augmentation torchvision CPU roi_pool x (objectness + 1)
(wsl/modeling/poolers.py:183-186, roi_heads_oicr.py:391-394), the mean of the Kr softmaxes and the decoded mean
deltas (fast_rcnn_oicr.py:712-783), then _get_augmented_boxes (:364-390; oracle/ref_tta.py, pinned to the reference's
own method) and fast_rcnn_inference_single_image (fast_rcnn_oicr.py:95-163; oracle/ref_heads.py, pinned likewise).
The pooling runs on a BOUNDED sample of the proposals (its cost is linear in them); everything after it runs on all."""
import time

import torch

from . import ref_heads, ref_tta


def heads_and_detections(wl, sequential_mean=True):
    """The part after pooling, on the full proposal set -> (detections dict, mean boxes, mean scores).  The checker of
    tests/test_gpu_infer_step.py uses the same function."""
    cfg = wl["cfg"]
    hw = (cfg["img_h"], cfg["img_w"])
    tf = ref_tta.tta_transforms(hw, hw, cfg["min_sizes"], cfg["max_size"], cfg["flip"])
    all_boxes, all_scores = [], []
    for a in wl["augs"]:
        probs = sum(torch.softmax(x, dim=-1) for x in a["ref_logits"]) / len(a["ref_logits"])
        deltas = torch.zeros_like(a["ref_deltas"][0])
        for dlt in a["ref_deltas"]:
            deltas += dlt
        deltas = deltas / len(a["ref_deltas"])
        all_scores.append(probs.unsqueeze(0))
        all_boxes.append(ref_heads.apply_deltas(deltas, a["boxes"]).unsqueeze(0))
    # sequential_mean: the machine-independent definition the device kernel follows (checker); False: torch.mean as
    # the reference calls it (timed leg)
    mb, ms = ref_tta.get_augmented_boxes(all_boxes, all_scores, tf,
                                         mean=ref_tta.mean_sequential if sequential_mean else None)
    det = ref_heads.fast_rcnn_inference_single_image(mb, ms, hw, cfg["score_thresh"], cfg["nms_thresh"], cfg["topk"])
    return det, mb, ms


def auto_sample_rois(wl, budget_s=6.0):
    cfg = wl["cfg"]
    per_roi_aug = 0.35e-3 * cfg["C"] / 512.0  # torchvision CPU roi_pool forward, ~16 threads, measured on the GPU box
    n = int(budget_s / (per_roi_aug * cfg["A"]))
    return max(64, min(n, cfg["R"]))


def time_step(wl, sample_rois, threads):
    import torchvision

    cfg = wl["cfg"]
    r = min(sample_rois, cfg["R"])
    P = (cfg["P"], cfg["P"])
    scale = (wl["objectness"][:r] + 1.0).view(-1, 1, 1, 1)
    torch.set_num_threads(threads)
    t0 = time.perf_counter()
    with torch.no_grad():
        for a in wl["augs"]:
            torchvision.ops.roi_pool(a["feat"].float(), a["rois"][:r].contiguous(), P, cfg["scale"]) * scale
        t_pool = time.perf_counter() - t0
        heads_and_detections(wl, sequential_mean=False)
    t_rest = time.perf_counter() - t0 - t_pool
    # pooling is linear in the proposals: scale the sampled time to all R, add the rest as measured
    t_full = t_pool * cfg["R"] / r + t_rest
    desc = ("%d augmentations x torchvision CPU roi_pool on the first %d of %d proposals (scaled linearly to %d) + the "
            "reference's head / TTA-merge / inference composition on all of them, %d threads" %
            (cfg["A"], r, cfg["R"], cfg["R"], threads))
    return t_full, cfg["A"] * cfg["R"], desc, dict(pool_sample_s=t_pool, rest_s=t_rest)
