"""Seeded synthetic inputs for the BASELINE.json configurations.

Everything is generated on the host with a CPU torch.Generator so that the CPU oracle
and the GPU see identical bits.  seed = 20210600 + 100 * config + sweep_index.
"""
import math

import torch

CONFIGS = {
    # name: dict(N, C, stride, img_h, img_w, R (per image), P, K, Kr, reg)
    "c1": dict(N=1, C=512, stride=16, img_h=600, img_w=800, R=2000, P=7, K=20, Kr=0, reg=False),
    "c2": dict(N=1, C=512, stride=8, img_h=600, img_w=800, R=2000, P=7, K=20, Kr=3, reg=False),
    "c3": dict(N=2, C=2048, stride=8, img_h=600, img_w=800, R=4000, P=7, K=80, Kr=4, reg=True),
    # config 2's training step with 4000 proposals (round 1 called this "c4"; BASELINE.json's c4 is INFERENCE: infer_case)
    "c2_r4000": dict(N=1, C=512, stride=8, img_h=600, img_w=800, R=4000, P=7, K=20, Kr=3, reg=False),
    # c3 with JTSM's own pooler (POOLER_TYPE "MOIPool", jtsm_WSR_50_DC5_1x.yaml:17): superpixel-masked pooling with
    # the heads' mask_scale x (objectness + 1) epilogue (roi_heads_jtsm.py:607-633); ~1000 superpixels per image
    "c3m": dict(N=2, C=2048, stride=8, img_h=600, img_w=800, R=4000, P=7, K=80, Kr=4, reg=True, pooler="MOIPool",
                sp_block=22),
}


def config_index(name):
    """c2 -> 2, c3m -> 3, c2_r4000 -> 4 (its round-1 seed) (seed / label rules follow the base configuration)."""
    if name == "c2_r4000":
        return 4
    return int("".join(ch for ch in name[1:] if ch.isdigit()))


# BASELINE.json configs[3]: multi-scale (5 scales + flip) WSL inference, 4000 proposals, OICR scoring + per-class NMS.
# TEST.AUG of the WSL VGG configs: MIN_SIZES (480, 576, 688, 864, 1200), MAX_SIZE 4000, FLIP True
# (projects/WSL/configs/PascalVOC-Detection/oicr_V_16_DC5_1x.yaml); thresholds :22-23; top 100 detections.
INFER_CONFIGS = {
    "c4": dict(N=1, C=512, stride=8, img_h=600, img_w=800, R=4000, P=7, K=20, Kr=3, reg=False,
               min_sizes=(480, 576, 688, 864, 1200), max_size=4000, flip=True, score_thresh=1e-5, nms_thresh=0.3, topk=100),
}


def infer_case(name, sweep_index=0, **over):
    """Everything one c4 inference step consumes, generated on the host: per augmentation the conv5 map of the resized
    (and flipped) image, the proposals mapped into that image (transform_proposals, test_time_augmentation_avg.py:27-70),
    objectness, and the Kr refinement heads' logits (+ deltas with box regression) that fc6 / fc7 would produce."""
    from . import tta_shapes

    cfg = dict(INFER_CONFIGS[name])
    cfg.update(over)
    g = generator(seed_for(config_index(name), sweep_index))
    hw = (cfg["img_h"], cfg["img_w"])
    boxes = proposals(g, cfg["R"], *hw)
    obj = torch.rand(cfg["R"], generator=g)
    augs = []
    for (nh, nw, flip) in tta_shapes.augmented_shapes(hw, cfg["min_sizes"], cfg["max_size"], cfg["flip"]):
        fh, fw = int(math.ceil(nh / cfg["stride"])), int(math.ceil(nw / cfg["stride"]))
        sx, sy = float(torch.tensor(nw * 1.0 / hw[1], dtype=torch.float32)), float(torch.tensor(nh * 1.0 / hw[0], dtype=torch.float32))
        b = boxes * torch.tensor([sx, sy, sx, sy])
        if flip:
            b = torch.stack([nw - b[:, 2], b[:, 1], nw - b[:, 0], b[:, 3]], dim=1)
        b[:, 0::2] = b[:, 0::2].clamp(0, nw)
        b[:, 1::2] = b[:, 1::2].clamp(0, nh)
        augs.append(dict(hw=(nh, nw), flip=flip, feat=feature_map(g, 1, cfg["C"], fh, fw), boxes=b,
                         rois=rois_from_boxes([b]),
                         ref_logits=[torch.randn(cfg["R"], cfg["K"] + 1, generator=g) * 2 for _ in range(cfg["Kr"])],
                         ref_deltas=[(torch.randn(cfg["R"], 4 * cfg["K"], generator=g) * 0.05 if cfg["reg"] else
                                      torch.zeros(cfg["R"], 4 * cfg["K"])) for _ in range(cfg["Kr"])]))
    cfg.update(scale=1.0 / cfg["stride"], A=len(augs))
    return dict(name=name, cfg=cfg, boxes=boxes, objectness=obj, augs=augs)


def seed_for(config_index, sweep_index=0):
    return 20210600 + 100 * int(config_index) + int(sweep_index)


def generator(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def feature_map(g, n, c, h, w):
    """relu(randn): non-negative like a post-ReLU conv5 map (ROILoopPool assumes >= 0)."""
    return torch.relu(torch.randn(n, c, h, w, generator=g))


def proposals(g, num, img_h, img_w, min_size=20.0):
    """MCG-like boxes [num,4] xyxy fp32: centre ~ U(image), sqrt(area) ~ logU(min_size,
    min(H,W)), aspect ~ logU(1/3, 3), clipped to the image."""
    cx = torch.rand(num, generator=g) * img_w
    cy = torch.rand(num, generator=g) * img_h
    lo, hi = math.log(min_size), math.log(float(min(img_h, img_w)))
    s = torch.exp(lo + torch.rand(num, generator=g) * (hi - lo))
    a = torch.exp((torch.rand(num, generator=g) * 2 - 1) * math.log(3.0))
    w = s * torch.sqrt(a)
    h = s / torch.sqrt(a)
    x1 = (cx - w / 2).clamp(0, img_w)
    y1 = (cy - h / 2).clamp(0, img_h)
    x2 = (cx + w / 2).clamp(0, img_w)
    y2 = (cy + h / 2).clamp(0, img_h)
    # keep boxes non-degenerate (>= 1 px) like the reference's proposal filtering
    x2 = torch.maximum(x2, x1 + 1.0).clamp(max=img_w)
    y2 = torch.maximum(y2, y1 + 1.0).clamp(max=img_h)
    x1 = torch.minimum(x1, x2 - 1.0)
    y1 = torch.minimum(y1, y2 - 1.0)
    return torch.stack([x1, y1, x2, y2], dim=1).float()


def random_boxes(g, num, max_coord=100):
    """Same construction as detectron2/utils/testing.py:29-40 `random_boxes`."""
    boxes = torch.rand(num, 4, generator=g) * (max_coord * 0.5)
    boxes.clamp_(min=1.0)
    boxes[:, 2:] += boxes[:, :2]
    return boxes


def rois_from_boxes(boxes_per_image):
    parts = []
    for i, b in enumerate(boxes_per_image):
        parts.append(torch.cat([torch.full((b.shape[0], 1), float(i)), b], dim=1))
    return torch.cat(parts, dim=0)


def image_labels(g, n, k, max_classes=3):
    """Per image: sorted distinct class ids (1..max_classes of them)."""
    out = []
    for _ in range(n):
        m = int(torch.randint(1, max_classes + 1, (1,), generator=g))
        out.append(torch.randperm(k, generator=g)[:m].sort().values.to(torch.int64))
    return out


def pool_case(name, sweep_index=0, **over):
    """Feature map + rois (+ objectness) for a named config."""
    cfg = dict(CONFIGS[name])
    cfg.update(over)
    cfg.setdefault("pooler", "ROIPool")
    g = generator(seed_for(config_index(name), sweep_index))
    h = int(math.ceil(cfg["img_h"] / cfg["stride"]))
    w = int(math.ceil(cfg["img_w"] / cfg["stride"]))
    feat = feature_map(g, cfg["N"], cfg["C"], h, w)
    boxes = [proposals(g, cfg["R"], cfg["img_h"], cfg["img_w"]) for _ in range(cfg["N"])]
    rois = rois_from_boxes(boxes)
    obj = torch.rand(rois.shape[0], generator=g)
    cfg.update(H=h, W=w, scale=1.0 / cfg["stride"])
    return cfg, feat, boxes, rois, obj


def step_case(name, sweep_index=0, **over):
    """Everything one hot-path step consumes, generated on the host:
    feature map, proposals, objectness, image labels, MIL logits (C, D), per-branch refinement
    logits (+ deltas when box regression is on) and the gradient that fc6's backward would
    hand to the pooling op."""
    cfg, feat, boxes, rois, obj = pool_case(name, sweep_index, **over)
    g = generator(seed_for(config_index(name), sweep_index) + 7)
    r_total, k = rois.shape[0], cfg["K"]
    labels = image_labels(g, cfg["N"], k, max_classes=3 if k <= 20 else 7)
    sigma = 5.0
    mil_cls = torch.randn(r_total, k, generator=g) * sigma
    mil_det = torch.randn(r_total, k, generator=g) * sigma
    ref_logits = [torch.randn(r_total, k + 1, generator=g) for _ in range(cfg["Kr"])]
    ref_deltas = [torch.randn(r_total, 4 * k, generator=g) * 0.1 for _ in range(cfg["Kr"])] if cfg["reg"] else []
    n_pooled = r_total * cfg["C"] * cfg["P"] * cfg["P"]
    # one [C,P,P] gradient tile per RoI-slot mod 64 keeps host memory bounded at config 3
    tile = torch.randn(min(r_total, 64), cfg["C"], cfg["P"], cfg["P"], generator=g)
    out = dict(name=name, cfg=cfg, feat=feat, boxes=boxes, rois=rois, objectness=obj, labels=labels,
               mil_cls=mil_cls, mil_det=mil_det, ref_logits=ref_logits, ref_deltas=ref_deltas,
               grad_tile=tile, n_pooled=n_pooled)
    if cfg["pooler"] == "MOIPool":
        sp, oh, cnt, s = superpixel_case(generator(seed_for(config_index(name), sweep_index) + 13), boxes,
                                         cfg["img_h"], cfg["img_w"], cfg.get("sp_block", 22))
        out.update(superpixels=sp, oh_labels=oh, cnt_sp=cnt, num_superpixels=s)
        cfg["S"] = s
    return out


def expand_grad_tile(tile, num_rois):
    """grad_pooled[r] = tile[r % tile.shape[0]] (host or device)."""
    reps = (num_rois + tile.shape[0] - 1) // tile.shape[0]
    return tile.repeat(reps, 1, 1, 1)[:num_rois].contiguous()


def superpixels(g, n, hs, ws, block=24):
    """Blocky superpixel label images [n, hs, ws] int32 (ids 0..S-1, jittered block grid) -> (labels, S)."""
    rows, cols = (hs + block - 1) // block, (ws + block - 1) // block
    s = rows * cols
    out = []
    for _ in range(n):
        perm = torch.randperm(s, generator=g).to(torch.int32).view(rows, cols)
        jy = torch.randint(0, block // 2 + 1, (1,), generator=g).item()
        jx = torch.randint(0, block // 2 + 1, (1,), generator=g).item()
        yy = ((torch.arange(hs) + jy) // block).clamp(max=rows - 1)
        xx = ((torch.arange(ws) + jx) // block).clamp(max=cols - 1)
        out.append(perm[yy][:, xx])
    return torch.stack(out).contiguous(), s


def oh_labels(g, num_rois, s, p=0.35):
    """oh_labels [R, S] int32: which superpixels are switched on for each RoI."""
    return (torch.rand(num_rois, s, generator=g) < p).to(torch.int32)


def pcl_case(g, r, k, p):
    """Synthetic PCL-loss inputs with the shapes / dtypes wsl/layers/pcl_loss.py passes (all fp32): probs [R, K+1]
    (column 0 = background), labels / cls_loss_weights / gt_assignment [1, R], pc_* / img_cls_loss_weights [1, P],
    im_labels [1, K+1].  Proposal clusters: P clusters with distinct-ish classes; labels[i] = class of its cluster or 0."""
    probs = torch.softmax(torch.randn(r, k + 1, generator=g) * 2, dim=1)
    probs[::17, 0] = 1e-8  # exercises the eps clamps
    pc_labels = torch.randint(1, k + 1, (1, p), generator=g).float()
    gt_assignment = torch.randint(0, p, (1, r), generator=g).float()
    is_bg = torch.rand(1, r, generator=g) < 0.6
    labels = torch.where(is_bg, torch.zeros(1, r), pc_labels[0, gt_assignment[0].long()][None])
    cls_loss_weights = torch.rand(1, r, generator=g)
    pc_probs = torch.rand(1, p, generator=g) * 0.9 + 1e-7
    pc_probs[0, ::5] = 1e-9
    pc_count = torch.randint(1, 40, (1, p), generator=g).float()
    img_cls_loss_weights = torch.rand(1, p, generator=g)
    im_labels = torch.zeros(1, k + 1)
    im_labels[0, 0] = 1.0
    im_labels[0, pc_labels[0, : max(1, p - 1)].long()] = 1.0  # the last cluster's class may be switched off
    return dict(probs=probs, labels=labels, cls_loss_weights=cls_loss_weights, gt_assignment=gt_assignment,
                pc_labels=pc_labels, pc_probs=pc_probs, pc_count=pc_count, img_cls_loss_weights=img_cls_loss_weights,
                im_labels=im_labels)


def superpixel_case(g, boxes_per_image, hs, ws, block=22):
    """Superpixels + box-consistent oh_labels: blocky label images as `superpixels` and
    oh_labels[r, s] = 1 iff at least half of segment s lies inside box r (segments are rectangles, so the overlap is
    analytic).  -> (superpixels int32 [N, hs, ws], oh_labels int32 [R, S], cnt_sp int64 [N, S], S)."""
    n = len(boxes_per_image)
    rows, cols = (hs + block - 1) // block, (ws + block - 1) // block
    s = rows * cols
    sps, ohs, cnts = [], [], []
    for i in range(n):
        perm = torch.randperm(s, generator=g).to(torch.int32).view(rows, cols)
        jy = torch.randint(0, block // 2 + 1, (1,), generator=g).item()
        jx = torch.randint(0, block // 2 + 1, (1,), generator=g).item()
        yy = ((torch.arange(hs) + jy) // block).clamp(max=rows - 1)
        xx = ((torch.arange(ws) + jx) // block).clamp(max=cols - 1)
        sp = perm[yy][:, xx]
        sps.append(sp)
        ys = torch.tensor([(yy == r_).nonzero()[0, 0].item() if (yy == r_).any() else hs for r_ in range(rows)]).float()
        ye = torch.tensor([(yy == r_).nonzero()[-1, 0].item() + 1 if (yy == r_).any() else hs for r_ in range(rows)]).float()
        xs = torch.tensor([(xx == c_).nonzero()[0, 0].item() if (xx == c_).any() else ws for c_ in range(cols)]).float()
        xe = torch.tensor([(xx == c_).nonzero()[-1, 0].item() + 1 if (xx == c_).any() else ws for c_ in range(cols)]).float()
        b = boxes_per_image[i]
        ovy = (torch.minimum(ye[None], b[:, 3:4]) - torch.maximum(ys[None], b[:, 1:2])).clamp(min=0)  # [R, rows]
        ovx = (torch.minimum(xe[None], b[:, 2:3]) - torch.maximum(xs[None], b[:, 0:1])).clamp(min=0)  # [R, cols]
        area = (ye - ys)[:, None] * (xe - xs)[None]  # [rows, cols]
        inside = ovy[:, :, None] * ovx[:, None, :] >= 0.5 * area[None].clamp(min=1.0)
        oh = torch.zeros(b.shape[0], s, dtype=torch.int32)
        oh[:, perm.reshape(-1).long()] = inside.reshape(b.shape[0], -1).to(torch.int32)
        ohs.append(oh)
        cnts.append(torch.bincount(sp.reshape(-1).long(), minlength=s))
    return torch.stack(sps).contiguous(), torch.cat(ohs).contiguous(), torch.stack(cnts), s
