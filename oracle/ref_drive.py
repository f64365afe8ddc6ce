"""Drive the REAL reference code (loaded in place by oracle/ref_loader.py) through one head step,
mirroring OICRROIHeads._forward_box (roi_heads_oicr.py:404-449) call for call.  This container only
(needs /root/reference).  Output format matches oracle/ref_heads.head_step."""
import types

import torch

from . import ref_loader as rl


def reference_head_step(mil_cls, mil_det, ref_logits, ref_deltas, proposal_boxes, labels, num_classes, mean_loss=True,
                        reg=False, reg_weighted=True, thresholds=(0.5,), iou_labels=(0, 1)):
    m = rl.install()
    Boxes, Instances = m["boxes"].Boxes, m["instances"].Instances
    wsddn, oicr = m["wsddn"], m["oicr"]
    b2b = m["box_regression"].Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    n = len(proposal_boxes)
    K = num_classes
    C = mil_cls.clone().requires_grad_()
    D = mil_det.clone().requires_grad_()
    RL = [x.clone().requires_grad_() for x in ref_logits]
    RD = [x.clone().requires_grad_() for x in ref_deltas]
    proposals = []
    for b in proposal_boxes:
        inst = Instances((600, 800))
        inst.proposal_boxes = Boxes(b.clone())
        inst.objectness_logits = torch.zeros(b.shape[0])
        proposals.append(inst)
    oh = torch.zeros(n, K)
    for i, l in enumerate(labels):
        oh[i, l] = 1.0
    # --- WSDDNOutputLayers.forward (fast_rcnn_wsddn.py:546-596) with cls/det = the given logits
    fake = types.SimpleNamespace(cls=lambda x: x[:, :K], det=lambda x: x[:, K:], cmil=False, training=True,
                                 num_bbox_reg_classes=K, box_dim=4)
    scores, deltas0 = wsddn.WSDDNOutputLayers.forward(fake, torch.cat([C, D], dim=1), proposals)
    outs = wsddn.WSDDNOutputs(b2b, scores, deltas0, proposals, 0.0, "smooth_l1", mean_loss, oh)
    losses = dict(outs.losses())
    img = outs.predict_probs_img().clone().detach()
    # --- refinement loop (roi_heads_oicr.py:413-449)
    self_, f = rl.make_self(K, labels, img, thresholds, iou_labels)
    prev_scores = scores.detach()
    prev_boxes = [p.proposal_boxes for p in proposals]
    branches = []
    layer = types.SimpleNamespace(box2box_transform=b2b)
    for k in range(len(RL)):
        targets = f["get_pgt_top_k"](self_, prev_boxes, prev_scores, proposals, suffix="_r%d" % k)
        props_k = f["label_and_sample_proposals"](self_, proposals, targets, suffix="_r%d" % k)
        dl = RD[k] if reg else torch.zeros(RL[k].shape[0], 4 * K)
        o = oicr.OICROutputs(b2b, RL[k], dl, props_k, 0.0, "smooth_l1_weighted" if reg_weighted else "smooth_l1",
                             mean_loss, "_r%d" % k, has_reg=reg)
        for name, v in o.losses().items():
            losses[name] = v
        pred = (RL[k], dl)
        probs = oicr.OICROutputLayers.predict_probs(layer, pred, props_k)
        boxes = oicr.OICROutputLayers.predict_boxes(layer, pred, props_k)
        prev_scores = [p.detach() for p in probs]
        prev_boxes = [b.detach() for b in boxes]
        branches.append(dict(
            pgt=[(t.gt_scores, t.gt_boxes.tensor, t.gt_classes, t.gt_weights) for t in targets],
            gt_classes=torch.cat([p.gt_classes for p in props_k]),
            gt_weights=torch.cat([p.gt_weights for p in props_k]),
            gt_boxes=torch.cat([p.gt_boxes.tensor for p in props_k]),
            probs=torch.cat(prev_scores), boxes=torch.cat(prev_boxes)))
    total = sum(losses.values())
    total.backward()
    return dict(losses={k: v.detach() for k, v in losses.items()}, scores=scores.detach(), img_scores=img,
                branches=branches, grad_cls=C.grad, grad_det=D.grad, grad_ref_logits=[x.grad for x in RL],
                grad_ref_deltas=[x.grad if x.grad is not None else torch.zeros_like(x) for x in RD])
