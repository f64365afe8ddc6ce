"""CPU restatement (eager PyTorch, fp32) of the reference's MIL / refinement head math.

TEST / BASELINE INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's CPU legs.
The reference has no native op for these stages -- it runs exactly this kind of eager composition
(also with MODEL.DEVICE=cpu), so this module is both the parity oracle and the timed "reference CPU
path" for the head stages.  Each function cites the reference lines it follows (paths relative to the
reference checkout).  Pinned against the reference's own code executed in place by
tests/test_ref_parity.py (oracle/ref_loader.py) and by the fixtures in tests/golden/ generated from it.
"""
import math

import torch
import torch.nn.functional as F


# ---------------------------------------------------------------- WSDDN (fast_rcnn_wsddn.py)
def wsddn_scores(C, D, num_preds_per_image):
    """softmax(C, dim=1) * softmax(D, dim=0) per image.  fast_rcnn_wsddn.py:571-584."""
    if len(num_preds_per_image) == 1:
        return F.softmax(C, dim=1) * F.softmax(D, dim=0)
    return torch.cat([F.softmax(c, dim=1) * F.softmax(d, dim=0)
                      for c, d in zip(C.split(num_preds_per_image, dim=0), D.split(num_preds_per_image, dim=0))], dim=0)


def wsddn_img_scores(scores, num_preds_per_image):
    """Proposal-sum image scores, clamped.  fast_rcnn_wsddn.py:364-379."""
    img = torch.cat([torch.sum(x, dim=0, keepdim=True) for x in scores.split(num_preds_per_image, dim=0)], dim=0)
    return torch.clamp(img, min=1e-6, max=1.0 - 1e-6)


def wsddn_bce(img_scores, gt_classes_img_oh, mean_loss=True):
    """fast_rcnn_wsddn.py:346-362."""
    if not mean_loss:
        return F.binary_cross_entropy(img_scores, gt_classes_img_oh, reduction="sum") / gt_classes_img_oh.size(0)
    return F.binary_cross_entropy(img_scores, gt_classes_img_oh, reduction="mean")


# ---------------------------------------------------------------- boxes (detectron2/structures/boxes.py)
def box_area(b):
    return (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])  # boxes.py:184-193


def pairwise_iou(b1, b2):
    """boxes.py:345-391 on raw tensors."""
    area1, area2 = box_area(b1), box_area(b2)
    wh = torch.min(b1[:, None, 2:], b2[:, 2:]) - torch.max(b1[:, None, :2], b2[:, :2])
    wh.clamp_(min=0)
    inter = wh.prod(dim=2)
    return torch.where(inter > 0, inter / (area1[:, None] + area2 - inter), torch.zeros(1, dtype=inter.dtype, device=inter.device))


def matcher(quality, thresholds=(0.5,), labels=(0, 1)):
    """Matcher.__call__ without low-quality matches.  detectron2/modeling/matcher.py:61-103."""
    if quality.numel() == 0:
        n = quality.size(1)
        return (torch.zeros(n, dtype=torch.int64, device=quality.device),
                torch.full((n,), labels[0], dtype=torch.int8, device=quality.device))
    vals, matches = quality.max(dim=0)
    out = torch.full(matches.size(), 1, dtype=torch.int8, device=quality.device)
    th = [-float("inf")] + list(thresholds) + [float("inf")]
    for l, low, high in zip(labels, th[:-1], th[1:]):
        out[(vals >= low) & (vals < high)] = l
    return matches, out


# ---------------------------------------------------------------- Box2BoxTransform (box_regression.py)
SCALE_CLAMP = math.log(1000.0 / 16)


def get_deltas(src, tgt, weights=(10.0, 10.0, 5.0, 5.0)):
    """box_regression.py:38-71."""
    sw, sh = src[:, 2] - src[:, 0], src[:, 3] - src[:, 1]
    sx, sy = src[:, 0] + 0.5 * sw, src[:, 1] + 0.5 * sh
    tw, th = tgt[:, 2] - tgt[:, 0], tgt[:, 3] - tgt[:, 1]
    tx, ty = tgt[:, 0] + 0.5 * tw, tgt[:, 1] + 0.5 * th
    wx, wy, ww, wh = weights
    return torch.stack((wx * (tx - sx) / sw, wy * (ty - sy) / sh, ww * torch.log(tw / sw), wh * torch.log(th / sh)), dim=1)


def apply_deltas(deltas, boxes, weights=(10.0, 10.0, 5.0, 5.0), scale_clamp=SCALE_CLAMP):
    """box_regression.py:73-109."""
    deltas = deltas.float()
    boxes = boxes.to(deltas.dtype)
    w, h = boxes[:, 2] - boxes[:, 0], boxes[:, 3] - boxes[:, 1]
    cx, cy = boxes[:, 0] + 0.5 * w, boxes[:, 1] + 0.5 * h
    wx, wy, ww, wh = weights
    dx, dy = deltas[:, 0::4] / wx, deltas[:, 1::4] / wy
    dw = torch.clamp(deltas[:, 2::4] / ww, max=scale_clamp)
    dh = torch.clamp(deltas[:, 3::4] / wh, max=scale_clamp)
    pcx, pcy = dx * w[:, None] + cx[:, None], dy * h[:, None] + cy[:, None]
    pw, ph = torch.exp(dw) * w[:, None], torch.exp(dh) * h[:, None]
    out = torch.zeros_like(deltas)
    out[:, 0::4] = pcx - 0.5 * pw
    out[:, 1::4] = pcy - 0.5 * ph
    out[:, 2::4] = pcx + 0.5 * pw
    out[:, 3::4] = pcy + 0.5 * ph
    return out


# ---------------------------------------------------------------- pseudo ground truth (roi_heads_oicr.py)
def get_pgt_top1(prev_boxes, prev_scores, gt_classes_img_int, pred_class_img_logits, num_classes):
    """get_pgt_top_k with top_k=1, thres=0.  roi_heads_oicr.py:675-811.
    prev_boxes: list of [R_i,4] (proposal boxes) or [R_i, 4K]; prev_scores: list of [R_i, K or K+1].
    -> per image (pgt_scores [G], pgt_boxes [G,4], pgt_classes [G] int64, pgt_weights [G])."""
    out = []
    for i, (pb, ps, gt_int) in enumerate(zip(prev_boxes, prev_scores, gt_classes_img_int)):
        n = pb.size(0)
        if pb.numel() == n * 4:
            pb = pb.unsqueeze(1).expand(n, num_classes, 4)  # :690-704
        pb = pb.reshape(-1, num_classes, 4)
        ps = torch.index_select(ps, 1, gt_int)  # :713-716
        pb = torch.index_select(pb, 1, gt_int)
        k = min(n, 1)
        sc, idx = torch.topk(ps, k, dim=0)  # :729-732
        idx = idx.unsqueeze(2).expand(k, gt_int.numel(), 4)
        bx = torch.gather(pb, 0, idx)
        cl = gt_int.unsqueeze(0).expand(k, gt_int.numel())
        wt = torch.index_select(pred_class_img_logits[i:i + 1], 1, gt_int).expand(k, gt_int.numel())  # :748-754
        out.append((sc.reshape(-1), bx.reshape(-1, 4), cl.reshape(-1), wt.reshape(-1)))
    return out


def get_pgt_top_k(prev_boxes, prev_scores, gt_classes_img_int, pred_class_img_logits, num_classes, top_k=1, thres=0.0):
    """General get_pgt_top_k (roi_heads_oicr.py:675-811): top_k >= 1 -> min(R_i, int(top_k)); 0 < top_k < 1 ->
    max(int(R_i * top_k), 1); seeds flattened k-major ([top_k, G] -> reshape(-1)); thres > 0 keeps row 0 and the
    rows with score >= thres (:756-785).  -> per image (scores, boxes [.,4], classes int64, weights)."""
    out = []
    for i, (pb, ps, gt_int) in enumerate(zip(prev_boxes, prev_scores, gt_classes_img_int)):
        n = pb.size(0)
        if pb.numel() == n * 4:
            pb = pb.unsqueeze(1).expand(n, num_classes, 4)
        pb = pb.reshape(-1, num_classes, 4)
        ps = torch.index_select(ps, 1, gt_int)
        pb = torch.index_select(pb, 1, gt_int)
        if top_k >= 1:
            k = min(n, int(top_k))
        elif top_k > 0:
            k = max(int(n * top_k), 1)
        else:
            k = min(n, 1)
        g = gt_int.numel()
        sc, idx = torch.topk(ps, k, dim=0)
        bx = torch.gather(pb, 0, idx.unsqueeze(2).expand(k, g, 4))
        cl = gt_int.unsqueeze(0).expand(k, g)
        wt = torch.index_select(pred_class_img_logits[i:i + 1], 1, gt_int).expand(k, g)
        if thres > 0:
            mask = sc.ge(thres)
            mask = torch.cat([torch.full_like(mask[0:1, :], True), mask[1:, :]], dim=0)
            sc, cl, wt = sc[mask], cl[mask], wt[mask]
            bx = bx[mask.unsqueeze(2).expand(k, g, 4)]
        out.append((sc.reshape(-1), bx.reshape(-1, 4), cl.reshape(-1), wt.reshape(-1)))
    return out


def get_pgt_mist(prev_boxes, prev_scores, gt_classes_img_int, pred_class_img_logits, num_classes, top_pro=0.15):
    """roi_heads_oicr.py:549-591: top `top_pro` fraction per image class, class-agnostic NMS at 0.2 over all of an
    image's candidates, keep order = descending score.  As in the reference (:579-581 zips pgt_scores twice) the
    returned gt_weights ARE the seed scores."""
    out = []
    for sc, bx, cl, _ in get_pgt_top_k(prev_boxes, prev_scores, gt_classes_img_int, pred_class_img_logits, num_classes,
                                       top_k=top_pro, thres=0.0):
        keep = batched_nms(bx, sc, torch.zeros_like(cl), 0.2)
        out.append((sc[keep], bx[keep], cl[keep], sc[keep]))
    return out


def label_proposals(pgt, proposal_boxes, num_classes, thresholds=(0.5,), labels=(0, 1)):
    """label_and_sample_proposals + _sample_proposals (no sub-sampling) for one image.
    roi_heads.py:222-254, 263-370.  -> dict(matched_idx, gt_classes, gt_boxes, gt_scores, gt_weights)."""
    sc, bx, cl, wt = pgt
    has_gt = cl.numel() > 0
    q = pairwise_iou(bx, proposal_boxes)
    midx, mlab = matcher(q, thresholds, labels)
    if has_gt:
        gc = cl[midx].clone()
        gc[mlab == 0] = num_classes
        gc[mlab == -1] = -1
        return dict(matched_idx=midx, gt_classes=gc, gt_boxes=bx[midx], gt_scores=sc[midx], gt_weights=wt[midx],
                    matched_iou=q.max(dim=0)[0])
    r = proposal_boxes.size(0)
    return dict(matched_idx=midx, gt_classes=torch.zeros_like(midx) + num_classes, gt_boxes=proposal_boxes,
                gt_scores=proposal_boxes.new_zeros(r), gt_weights=proposal_boxes.new_zeros(r),
                matched_iou=proposal_boxes.new_zeros(r))


# ---------------------------------------------------------------- OICR losses (fast_rcnn_oicr.py)
def oicr_weights(gt_classes, gt_weights):
    """fast_rcnn_oicr.py:243-247."""
    w = gt_weights.clone()
    w[gt_classes == -1] = 0.0
    valid = torch.zeros_like(w)
    valid[w > 1e-12] = 1.0
    return w, valid


def oicr_ce_loss(logits, gt_classes, gt_weights):
    """fast_rcnn_oicr.py:282-298."""
    w, valid = oicr_weights(gt_classes, gt_weights)
    loss = F.cross_entropy(logits, gt_classes, reduction="none", ignore_index=-1)
    return (loss * w).sum() / valid.sum()


def oicr_accuracy(logits, gt_classes):
    """_log_accuracy counters.  fast_rcnn_oicr.py:252-280."""
    pred = logits.argmax(dim=1)
    bg = logits.shape[1] - 1
    fg = (gt_classes >= 0) & (gt_classes < bg)
    return [int((pred == gt_classes).sum()), int(fg.sum()), int((pred[fg] == gt_classes[fg]).sum()),
            int((pred[fg] == bg).sum())]


def smooth_l1(x, y, beta):
    # fvcore.nn.smooth_l1_loss (third-party, absent from the checkout; published definition)
    if beta < 1e-5:
        return torch.abs(x - y)
    n = torch.abs(x - y)
    return torch.where(n < beta, 0.5 * n ** 2 / beta, n - 0.5 * beta)


def oicr_box_reg_loss(deltas, proposals, gt_boxes, gt_classes, gt_weights, num_classes, weights=(10.0, 10.0, 5.0, 5.0),
                      beta=0.0, weighted=True):
    """fast_rcnn_oicr.py:300-380 ("smooth_l1" / "smooth_l1_weighted")."""
    w, _ = oicr_weights(gt_classes, gt_weights)
    fg = torch.nonzero((gt_classes >= 0) & (gt_classes < num_classes), as_tuple=True)[0]
    if deltas.size(1) == 4:
        cols = torch.arange(4, device=deltas.device)
    else:
        cols = 4 * gt_classes[fg][:, None] + torch.arange(4, device=deltas.device)
    tgt = get_deltas(proposals, gt_boxes, weights)
    l = smooth_l1(deltas[fg[:, None], cols], tgt[fg], beta)
    if weighted:
        l = l * w[fg, None]
    return l.sum() / gt_classes.numel()


# ---------------------------------------------------------------- the whole head step
def head_step(mil_cls, mil_det, ref_logits, ref_deltas, proposal_boxes, labels, num_classes, mean_loss=True,
              reg=False, reg_weighted=True, thresholds=(0.5,), iou_labels=(0, 1), need_grad=True, refine_mist=False,
              mist_top_pro=0.15):
    """One training pass of the head math of OICRROIHeads._forward_box (roi_heads_oicr.py:404-449):
    WSDDN scores + BCE, then per refinement branch: pseudo-gt -> labels -> weighted CE (+ box reg)
    -> softmax / decoded boxes for the next branch.  Returns losses, per-branch labels, and
    (need_grad) gradients w.r.t. every logit tensor.  `proposal_boxes`: list of [R_i,4]; `labels`:
    list of int64 class-id tensors (gt_classes_img_int)."""
    num_preds = [int(b.shape[0]) for b in proposal_boxes]
    n = len(num_preds)
    C = mil_cls.clone().requires_grad_(need_grad)
    D = mil_det.clone().requires_grad_(need_grad)
    RL = [x.clone().requires_grad_(need_grad) for x in ref_logits]
    RD = [x.clone().requires_grad_(need_grad) for x in ref_deltas]
    oh = torch.zeros(n, num_classes, device=mil_cls.device)
    for i, l in enumerate(labels):
        oh[i, l] = 1.0
    scores = wsddn_scores(C, D, num_preds)
    img = wsddn_img_scores(scores, num_preds)
    losses = {"loss_cls": wsddn_bce(img, oh, mean_loss)}
    img_det = img.clone().detach()
    prev_scores = list(scores.detach().split(num_preds, dim=0))
    prev_boxes = list(proposal_boxes)
    allp = torch.cat(proposal_boxes, dim=0)
    branches = []
    for k in range(len(RL)):
        if refine_mist:  # roi_heads_oicr.py:418-424 (term weight 3 on the first branch)
            pgt = get_pgt_mist(prev_boxes, prev_scores, labels, img_det, num_classes, mist_top_pro)
        else:
            pgt = get_pgt_top1(prev_boxes, prev_scores, labels, img_det, num_classes)
        tw = 3.0 if refine_mist and k == 0 else 1.0
        lab = [label_proposals(p, b, num_classes, thresholds, iou_labels) for p, b in zip(pgt, proposal_boxes)]
        gt_classes = torch.cat([l["gt_classes"] for l in lab])
        gt_weights = torch.cat([l["gt_weights"] for l in lab])
        gt_boxes = torch.cat([l["gt_boxes"] for l in lab])
        losses["loss_cls_r%d" % k] = oicr_ce_loss(RL[k], gt_classes, gt_weights) * tw
        if reg:
            losses["loss_box_reg_r%d" % k] = oicr_box_reg_loss(RD[k], allp, gt_boxes, gt_classes, gt_weights,
                                                               num_classes, weighted=reg_weighted) * tw
        probs = F.softmax(RL[k].detach(), dim=-1)
        dl = RD[k].detach() if reg else torch.zeros(allp.shape[0], 4 * num_classes, device=allp.device)
        boxes = apply_deltas(dl, allp)
        prev_scores = list(probs.split(num_preds, dim=0))
        prev_boxes = list(boxes.split(num_preds, dim=0))
        branches.append(dict(pgt=pgt, gt_classes=gt_classes, gt_weights=gt_weights, gt_boxes=gt_boxes,
                             matched_idx=torch.cat([l["matched_idx"] for l in lab]),
                             matched_iou=torch.cat([l["matched_iou"] for l in lab]), probs=probs, boxes=boxes,
                             acc=oicr_accuracy(RL[k].detach(), gt_classes)))
    out = dict(losses={k: v.detach() for k, v in losses.items()}, scores=scores.detach(), img_scores=img_det,
               branches=branches)
    if need_grad:
        total = sum(losses.values())
        total.backward()
        out["grad_cls"], out["grad_det"] = C.grad, D.grad
        out["grad_ref_logits"] = [x.grad for x in RL]
        out["grad_ref_deltas"] = [x.grad if x.grad is not None else torch.zeros_like(x) for x in RD]
    return out


# ---------------------------------------------------------------- inference (fast_rcnn_oicr.py:95-163, nms.py)
def batched_nms(boxes, scores, idxs, thr):
    """Per-class greedy NMS on the original coordinates (detectron2/layers/nms.py:24-30, the >= 40000
    branch; torchvision's _batched_nms_vanilla is the same computation), keep sorted by descending
    score with ties broken by lower index.  NMS arithmetic: oracle C (torchvision CPU kernel order)."""
    import numpy as np

    from . import capi

    b, s, c = boxes.numpy(), scores.numpy(), idxs.numpy()
    keep = []
    for cl in np.unique(c):
        m = np.nonzero(c == cl)[0]
        k = capi.nms(b[m], s[m], thr)
        keep.append(m[k])
    keep = np.concatenate(keep) if keep else np.zeros(0, np.int64)
    order = np.lexsort((keep, -s[keep]))
    return torch.from_numpy(keep[order].astype(np.int64))


def fast_rcnn_inference_single_image(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image):
    """fast_rcnn_oicr.py:95-163 on CPU tensors -> dict(pred_boxes, scores, pred_classes, pred_inds, kept_rois).
    NMS = `batched_nms` above (per class on the original coordinates)."""
    pred_inds = torch.arange(scores.size(0), dtype=torch.long).unsqueeze(1).repeat(1, scores.size(1))
    valid = torch.isfinite(boxes).all(dim=1) & torch.isfinite(scores).all(dim=1)
    if not valid.all():
        boxes, scores, pred_inds = boxes[valid], scores[valid], pred_inds[valid]
    scores = scores[:, :-1]
    nb = boxes.shape[1] // 4
    h, w = image_shape
    b = boxes.reshape(-1, 4).clone()
    b[:, 0].clamp_(min=0, max=w)
    b[:, 1].clamp_(min=0, max=h)
    b[:, 2].clamp_(min=0, max=w)
    b[:, 3].clamp_(min=0, max=h)
    b = b.view(-1, nb, 4)
    pred_inds = pred_inds[:, :-1]
    mask = scores > score_thresh
    fi = mask.nonzero()
    b = b[fi[:, 0], 0] if nb == 1 else b[mask]
    s = scores[mask]
    pred_inds = pred_inds[mask]
    keep = batched_nms(b, s, fi[:, 1], nms_thresh)
    if topk_per_image >= 0:
        keep = keep[:topk_per_image]
    return dict(pred_boxes=b[keep], scores=s[keep], pred_classes=fi[keep][:, 1], pred_inds=pred_inds[keep],
                kept_rois=fi[keep][:, 0])


def pairwise_iou_sp(oh_labels1, oh_labels2, cnt_sp):
    """pairwise_iou_sp, roi_heads_jtsm.py:49-73, without the [n1, n2, S] temporaries: the three sums are exact
    integer sums, so any order gives the reference's values; the division is the same fp32 division."""
    c = cnt_sp.reshape(1, -1).to(torch.int64)
    a, b = oh_labels1.to(torch.int64), oh_labels2.to(torch.int64)
    inter = (a * c) @ b.t()
    area1 = (a * c).sum(1, keepdim=True)
    area2 = (b * c).sum(1).unsqueeze(0)
    return torch.div(inter.to(torch.float32), (area1 + area2 - inter).to(torch.float32))
