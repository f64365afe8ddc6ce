"""Head math of the WSL ROI heads on top of the fused kernels (ops.py -> C ABI).

Functional mirror of what `OICRROIHeads._forward_box` does after fc6/fc7
(projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:404-449): WSDDN scores + BCE, then per refinement
branch pseudo-gt -> label assignment -> weighted CE (+ box regression) -> softmax / decoded boxes.
Everything stays on the device; no host synchronisation inside (the reference does 3-7 `.item()` /
`nonzero` syncs per branch and image)."""
import torch

from . import ops


def head_step(mil_cls, mil_det, ref_logits, ref_deltas, proposals, num_preds_per_image, gt_classes_img_int,
              num_classes, labels_onehot, mean_loss=True, reg=False, reg_weighted=True, iou_thresholds=(0.5,),
              iou_labels=(0, 1), need_grad=True, term_weights=None, smooth_l1_beta=0.0,
              bbox_reg_weights=(10.0, 10.0, 5.0, 5.0), refine_mist=False, mist_top_pro=0.15, streams=None,
              compact_seeds=False):
    """proposals: [R,4] (all images concatenated).  Returns dict(losses, scores, img_scores, branches[...],
    grad_cls, grad_det, grad_ref_logits, grad_ref_deltas).  Launches: 1 (MIL fwd+loss) + 1 (MIL bwd)
    + per branch 2 (assign, CE fwd/bwd/softmax) [+ 2 with box regression].
    refine_mist: WSL.REFINE_MIST targets (roi_heads_oicr.py:418-424) -- seeds from get_pgt_mist, branch 0 weighted x3
    unless term_weights is given; the NMS survivor counts stay on the device (padded seed slots + jtsm_refine_match_counted),
    so this variant is graph-capturable too when the image labels are device tensors.  compact_seeds=True trims the
    returned pgt_* to the seeds proper (one host read; tests / logging).
    streams: optional list of >= len(ref_logits) + 1 side streams.  The branches only meet through the previous
    branch's softmax scores and decoded boxes, and neither depends on the labels, so with side streams every branch
    (its own softmax of the previous logits, assign, CE, box regression) and the MIL backward run concurrently
    after the MIL forward instead of as one chain of small latency-bound launches; same kernels, same results."""
    fwd = ops.wsddn_forward(mil_cls, mil_det, num_preds_per_image, labels_onehot, mean_loss)
    out = dict(losses={"loss_cls": fwd["loss"]}, scores=fwd["scores"], img_scores=fwd["img_scores"], branches=[])
    if streams is not None and not refine_mist and len(ref_logits) > 0 and len(streams) > len(ref_logits):
        return _head_step_concurrent(fwd, out, ref_logits, ref_deltas, proposals, num_preds_per_image,
                                     gt_classes_img_int, num_classes, reg, reg_weighted, iou_thresholds, iou_labels,
                                     need_grad, term_weights, smooth_l1_beta, bbox_reg_weights, streams)
    prev_scores, prev_boxes = fwd["scores"], proposals
    grads_l, grads_d = [], []
    boxes_noreg = None
    for k, logits in enumerate(ref_logits):
        tw = (3.0 if refine_mist and k == 0 else 1.0) if term_weights is None else float(term_weights[k])
        if refine_mist:
            a = refine_labels_mist(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int,
                                   fwd["img_scores"], num_classes, mist_top_pro, iou_thresholds, iou_labels, want_iou=True,
                                   compact=compact_seeds)
        else:
            a = ops.refine_assign(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int,
                                  fwd["img_scores"], num_classes, iou_thresholds, iou_labels, want_iou=True)
        ce = ops.refine_ce(logits, a["gt_classes"], a["loss_weights"], a["valid_count"], loss_scale=tw,
                           want_grad=need_grad)
        out["losses"]["loss_cls_r%d" % k] = ce["loss"]
        grads_l.append(ce["grad_logits"])
        if reg:
            br = ops.refine_box_reg(ref_deltas[k], proposals, a["gt_boxes"], a["gt_classes"],
                                    a["loss_weights"] if reg_weighted else None, num_classes, bbox_reg_weights,
                                    smooth_l1_beta, loss_scale=tw, want_grad=need_grad)
            out["losses"]["loss_box_reg_r%d" % k] = br["loss"]
            grads_d.append(br["grad_deltas"])
            boxes = ops.apply_deltas(ref_deltas[k], proposals, bbox_reg_weights)
        else:
            # zero deltas: predict_boxes still round-trips through (ctr, size) (fast_rcnn_oicr.py:684-709)
            if boxes_noreg is None:
                boxes_noreg = ops.apply_deltas(torch.zeros(proposals.shape[0], 4, device=proposals.device),
                                               proposals, bbox_reg_weights)
            boxes = boxes_noreg
        a.update(probs=ce["probs"], boxes=boxes, acc=ce["acc"])
        out["branches"].append(a)
        prev_scores, prev_boxes = ce["probs"], boxes
    if need_grad:
        out["grad_cls"], out["grad_det"] = ops.wsddn_backward(fwd)
        out["grad_ref_logits"], out["grad_ref_deltas"] = grads_l, grads_d
    return out


def _head_step_concurrent(fwd, out, ref_logits, ref_deltas, proposals, num_preds_per_image, gt_classes_img_int,
                          num_classes, reg, reg_weighted, iou_thresholds, iou_labels, need_grad, term_weights,
                          smooth_l1_beta, bbox_reg_weights, streams):
    main = torch.cuda.current_stream()
    kr = len(ref_logits)
    boxes_noreg = None
    if not reg:  # zero deltas: predict_boxes still round-trips through (ctr, size) (fast_rcnn_oicr.py:684-709)
        boxes_noreg = ops.apply_deltas(torch.zeros(proposals.shape[0], 4, device=proposals.device), proposals,
                                       bbox_reg_weights)
    start = torch.cuda.Event()
    start.record(main)
    boxes, box_ready = [None] * kr, [None] * kr
    branches, losses, grads_l, grads_d = [None] * kr, [dict() for _ in range(kr)], [None] * kr, [None] * kr
    for k, logits in enumerate(ref_logits):
        st = streams[k]
        st.wait_event(start)
        tw = 1.0 if term_weights is None else float(term_weights[k])
        with torch.cuda.stream(st):
            if reg:
                boxes[k] = ops.apply_deltas(ref_deltas[k], proposals, bbox_reg_weights)
                box_ready[k] = torch.cuda.Event()
                box_ready[k].record(st)
            else:
                boxes[k] = boxes_noreg
            if k == 0:
                prev_scores, prev_boxes = fwd["scores"], proposals
            else:
                prev_scores = ops.softmax_rows(ref_logits[k - 1])  # = the previous branch's probs, bit for bit
                prev_boxes = boxes[k - 1]
                if reg:
                    st.wait_event(box_ready[k - 1])
            a = ops.refine_assign(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int,
                                  fwd["img_scores"], num_classes, iou_thresholds, iou_labels, want_iou=True)
            ce = ops.refine_ce(logits, a["gt_classes"], a["loss_weights"], a["valid_count"], loss_scale=tw,
                               want_grad=need_grad)
            losses[k]["loss_cls_r%d" % k] = ce["loss"]
            grads_l[k] = ce["grad_logits"]
            if reg:
                br = ops.refine_box_reg(ref_deltas[k], proposals, a["gt_boxes"], a["gt_classes"],
                                        a["loss_weights"] if reg_weighted else None, num_classes, bbox_reg_weights,
                                        smooth_l1_beta, loss_scale=tw, want_grad=need_grad)
                losses[k]["loss_box_reg_r%d" % k] = br["loss"]
                grads_d[k] = br["grad_deltas"]
            a.update(probs=ce["probs"], boxes=boxes[k], acc=ce["acc"])
            branches[k] = a
            for t in list(a.values()) + [ce["loss"], ce["grad_logits"], prev_scores] + \
                    ([br["loss"], br["grad_deltas"]] if reg else []):
                if torch.is_tensor(t):
                    t.record_stream(main)
    if need_grad:
        st = streams[kr]
        st.wait_event(start)
        with torch.cuda.stream(st):
            out["grad_cls"], out["grad_det"] = ops.wsddn_backward(fwd)
            out["grad_cls"].record_stream(main)
            out["grad_det"].record_stream(main)
    for st in streams[:kr + 1]:
        main.wait_stream(st)
    for k in range(kr):
        out["losses"].update(losses[k])
    out["branches"] = branches
    if need_grad:
        out["grad_ref_logits"], out["grad_ref_deltas"] = grads_l, [g for g in grads_d if g is not None]
    return out


# --------------------------------------------------------------------------- pseudo-gt variants beyond top-1
def get_pgt_top_k(prev_scores, prev_boxes, num_preds_per_image, gt_classes_img_int, img_scores, num_classes, top_k=1,
                  thres=0.0, device_counts=False):
    """get_pgt_top_k for any top_k / thres (roi_heads_oicr.py:675-811) on device tensors.
    prev_scores [R, K or K+1], prev_boxes [R,4] or [R,4K] (all images concatenated), img_scores [N,K].
    -> (scores [G], boxes [G,4], classes [G] int64, weights [G], seeds_per_image list[int]); the top-1, thres = 0
    case is fused into ops.refine_assign -- this is the general path (torch.topk + gathers).
    Without `thres` the seed counts follow from shapes.  With thres > 0 they are data dependent:
      device_counts=False  the seed lists are trimmed (boolean indexing: one host sync per image, as the reference);
      device_counts=True   every image keeps its full k x g slot with the surviving seeds moved to the front in the
                           reference's order (stable partition) and a sixth result, int32 [N] on the device, says how
                           many of each slot are seeds -- no host read, graph-capturable (ops.refine_match seed_counts=)."""
    outs, counts, dev_counts = [], [], []
    sc_split = prev_scores.split(num_preds_per_image, dim=0)
    bx_split = prev_boxes.split(num_preds_per_image, dim=0)
    for i, (ps, pb, gt_int) in enumerate(zip(sc_split, bx_split, gt_classes_img_int)):
        n = ps.shape[0]
        gt_int = gt_int.to(ps.device)
        g = gt_int.numel()
        if top_k >= 1:
            k = min(n, int(top_k))
        elif top_k > 0:
            k = max(int(n * top_k), 1)
        else:
            k = min(n, 1)
        if pb.numel() == n * 4:
            pb = pb.view(n, 1, 4).expand(n, num_classes, 4)
        pb = pb.reshape(n, num_classes, 4).index_select(1, gt_int)
        sc, idx = torch.topk(ps.index_select(1, gt_int), k, dim=0)
        bx = torch.gather(pb, 0, idx.unsqueeze(2).expand(k, g, 4))
        cl = gt_int.unsqueeze(0).expand(k, g)
        wt = img_scores[i:i + 1].index_select(1, gt_int).expand(k, g)
        if thres > 0 and device_counts:
            mask = sc.ge(thres)
            mask[0:1, :] = True
            flat = mask.reshape(-1)
            order = torch.argsort((~flat).to(torch.uint8), stable=True)  # survivors first, row-major order kept
            sc, cl, wt = sc.reshape(-1)[order], cl.reshape(-1)[order], wt.reshape(-1)[order]
            bx = bx.reshape(-1, 4)[order]
            dev_counts.append(flat.sum(dtype=torch.int32).reshape(1))
        elif thres > 0:
            mask = sc.ge(thres)
            mask[0:1, :] = True
            sc, cl, wt, bx = sc[mask], cl[mask], wt[mask], bx[mask.unsqueeze(2).expand(k, g, 4)]
        elif device_counts:
            dev_counts.append(torch.full((1,), k * g, dtype=torch.int32, device=ps.device))
        outs.append((sc.reshape(-1), bx.reshape(-1, 4), cl.reshape(-1), wt.reshape(-1)))
        counts.append(int(outs[-1][0].shape[0]))
    cat = [torch.cat([o[j] for o in outs]) for j in range(4)]
    if device_counts:
        return cat[0], cat[1], cat[2], cat[3], counts, torch.cat(dev_counts)
    return cat[0], cat[1], cat[2], cat[3], counts


def get_pgt_mist(prev_scores, prev_boxes, num_preds_per_image, gt_classes_img_int, img_scores, num_classes,
                 top_pro=0.15, nms_thresh=0.2, device_counts=False):
    """get_pgt_mist (roi_heads_oicr.py:549-591): top `top_pro` per image class, then class-agnostic NMS per image
    with the sm_100a NMS kernel.  gt_weights = seed scores, as the reference returns them (:579-581).
    device_counts=False: trimmed seed lists (one host read of the keep count per image, as the reference's indexing by
    `keep`).  device_counts=True: the survivors sit at the front of each image's full slot, in NMS (descending score)
    order, with their number in an int32 [N] device tensor (sixth result) -- no host read."""
    sc, bx, cl, _, counts = get_pgt_top_k(prev_scores, prev_boxes, num_preds_per_image, gt_classes_img_int, img_scores,
                                          num_classes, top_k=top_pro)
    keeps, kept, nums = [], [], []
    off = 0
    for c in counts:
        keep, num = ops.batched_nms_raw(bx[off:off + c], sc[off:off + c], None, nms_thresh)
        if device_counts:
            keeps.append(keep.clamp_(min=0) + off)  # padding (-1) -> a valid row; never read past seed_counts
            nums.append(num.to(torch.int32))
            kept.append(c)
        else:
            k = int(num.item())
            keeps.append(keep[:k] + off)
            kept.append(k)
        off += c
    keep = torch.cat(keeps) if keeps else torch.zeros(0, dtype=torch.int64, device=sc.device)
    if device_counts:
        return sc[keep], bx[keep], cl[keep], sc[keep], kept, torch.cat(nums)
    return sc[keep], bx[keep], cl[keep], sc[keep], kept


def refine_labels_mist(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int, img_scores,
                       num_classes, top_pro=0.15, iou_thresholds=(0.5,), iou_labels=(0, 1), want_iou=False,
                       compact=False):
    """One refinement branch's targets with WSL.REFINE_MIST (roi_heads_oicr.py:418-427): mist seeds ->
    label_and_sample_proposals.  Returns the ops.refine_match dict plus the seeds.
    Nothing is read back on the host: the seeds stay in padded per-image slots (`seed_slots` host ints) and the matching
    kernel takes each image's seed count from `seed_counts` (device).  compact=True additionally trims pgt_* to the
    seeds proper and fills `seeds_per_image` (host ints; one device -> host copy, for tests and logging)."""
    sc, bx, cl, wt, slots, cnt = get_pgt_mist(prev_scores, prev_boxes, num_preds_per_image, gt_classes_img_int,
                                              img_scores, num_classes, top_pro, device_counts=True)
    out = ops.refine_match(bx, cl, sc, wt, slots, proposals, num_preds_per_image, num_classes, iou_thresholds,
                           iou_labels, want_iou, seed_counts=cnt)
    out.update(pgt_scores=sc, pgt_boxes=bx, pgt_classes=cl, pgt_weights=wt, seed_slots=slots, seed_counts=cnt)
    if compact:
        kept = [int(v) for v in cnt.tolist()]
        sel, off = [], 0
        for c, k in zip(slots, kept):
            sel.append(torch.arange(off, off + k, device=sc.device))
            off += c
        sel = torch.cat(sel) if sel else torch.zeros(0, dtype=torch.int64, device=sc.device)
        out.update(pgt_scores=sc[sel], pgt_boxes=bx[sel], pgt_classes=cl[sel], pgt_weights=wt[sel],
                   seeds_per_image=kept)
    return out


def refine_labels_top_k(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int, img_scores,
                        num_classes, top_k=1, thres=0.0, iou_thresholds=(0.5,), iou_labels=(0, 1), want_iou=False):
    """A refinement branch's targets for any top_k / thres of get_pgt_top_k (roi_heads_oicr.py:675-811) followed by
    label_and_sample_proposals, without a host read (thresholded seed counts stay on the device)."""
    sc, bx, cl, wt, slots, cnt = get_pgt_top_k(prev_scores, prev_boxes, num_preds_per_image, gt_classes_img_int,
                                               img_scores, num_classes, top_k, thres, device_counts=True)
    out = ops.refine_match(bx, cl, sc, wt, slots, proposals, num_preds_per_image, num_classes, iou_thresholds,
                           iou_labels, want_iou, seed_counts=cnt)
    out.update(pgt_scores=sc, pgt_boxes=bx, pgt_classes=cl, pgt_weights=wt, seed_slots=slots, seed_counts=cnt)
    return out
