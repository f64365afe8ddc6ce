"""Generate tests/golden/modules_*.npz: the REFERENCE's own `WSDDNOutputLayers`, `OICROutputLayers`, `OICROutputs`,
`get_pgt_top_k` and `label_and_sample_proposals` (run in place from /root/reference via oracle/ref_loader.py), wired as
`OICRROIHeads._forward_box` wires them (roi_heads_oicr.py:404-449), on seeded fc7 activations -> losses, pseudo-labels
and PARAMETER GRADIENTS after `loss.backward()`.  Run in the build container:
    python tests/golden/make_golden_modules.py
Inputs and parameters are regenerated from the seed by the tests (same `case()` / `params()`), so only the reference
outputs are stored.  Reference commit: a691dc59."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from jtsm_b200 import synth  # noqa: E402

CASES = {
    # name: (seed, proposals per image, fc7 width, K, Kr, box regression)
    "m0": (501, (260, 180), 48, 20, 3, False),
    "m1": (502, (300,), 32, 9, 2, True),
}


def case(name):
    seed, counts, d, k, kr, reg = CASES[name]
    g = synth.generator(seed)
    boxes = [synth.proposals(g, c, 600, 800) for c in counts]
    if reg:  # clustered proposals (jittered copies of a few boxes), so that the regression loss sees many foreground rows
        for b in boxes:
            n = b.shape[0] // 2
            src = b[torch.randint(0, 12, (n,), generator=g)]
            jit = (torch.rand(n, 4, generator=g) - 0.5) * 0.25 * (src[:, 2:] - src[:, :2]).repeat(1, 2)
            nb = src + jit
            nb[:, 0::2] = nb[:, 0::2].clamp(0, 800)
            nb[:, 1::2] = nb[:, 1::2].clamp(0, 600)
            b[-n:] = torch.cat([torch.minimum(nb[:, :2], nb[:, 2:] - 1), nb[:, 2:]], dim=1)
    labels = synth.image_labels(g, len(counts), k)
    x = torch.relu(torch.randn(sum(counts), d, generator=g))
    return dict(x=x, boxes=boxes, labels=labels, counts=counts, d=d, k=k, kr=kr, reg=reg)


def params(name):
    """Seeded parameter values (the reference's own initialisation gives near-zero refinement logits)."""
    seed, counts, d, k, kr, reg = CASES[name]
    g = synth.generator(seed + 50)
    p = {"box_predictor.cls.weight": torch.randn(k, d, generator=g) * 0.6, "box_predictor.cls.bias": torch.randn(k, generator=g) * 0.1,
         "box_predictor.det.weight": torch.randn(k, d, generator=g) * 0.6, "box_predictor.det.bias": torch.randn(k, generator=g) * 0.1}
    for i in range(kr):
        p["box_refinery_%d.cls_score.weight" % i] = torch.randn(k + 1, d, generator=g) * 0.4
        p["box_refinery_%d.cls_score.bias" % i] = torch.randn(k + 1, generator=g) * 0.1
        p["box_refinery_%d.bbox_pred.weight" % i] = torch.randn(4 * k, d, generator=g) * 0.02
        p["box_refinery_%d.bbox_pred.bias" % i] = torch.randn(4 * k, generator=g) * 0.01
    return p


def load_params(modules, p):
    """modules: {"box_predictor": m, "box_refinery_0": m, ...}"""
    with torch.no_grad():
        for key, v in p.items():
            mod, layer, what = key.split(".")
            getattr(getattr(modules[mod], layer), what).copy_(v)


def reference_run(name):
    from oracle import ref_loader as rl

    c = case(name)
    m = rl.install()
    Boxes, Instances = m["boxes"].Boxes, m["instances"].Instances
    wsddn, oicr = m["wsddn"], m["oicr"]
    ShapeSpec = sys.modules["detectron2.layers"].ShapeSpec
    b2b = m["box_regression"].Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    k, kr, reg = c["k"], c["kr"], c["reg"]
    pred = wsddn.WSDDNOutputLayers(ShapeSpec(channels=c["d"]), box2box_transform=b2b, num_classes=k, mean_loss=True)
    refs = [oicr.OICROutputLayers(ShapeSpec(channels=c["d"]), box2box_transform=b2b, num_classes=k, refine_k=i,
                                  refine_reg=[reg] * kr, box_reg_loss_type="smooth_l1_weighted" if reg else "smooth_l1",
                                  loss_weight={"loss_box_reg": 1.0})
            for i in range(kr)]
    modules = {"box_predictor": pred}
    modules.update({"box_refinery_%d" % i: r for i, r in enumerate(refs)})
    load_params(modules, params(name))
    proposals = []
    for b in c["boxes"]:
        inst = Instances((600, 800))
        inst.proposal_boxes = Boxes(b.clone())
        inst.objectness_logits = torch.zeros(b.shape[0])
        proposals.append(inst)
    oh = torch.zeros(len(c["boxes"]), k)
    for i, l in enumerate(c["labels"]):
        oh[i, l] = 1.0
    x = c["x"].clone().requires_grad_()
    predictions = pred(x, proposals)
    losses = dict(pred.losses(predictions, proposals, oh))
    img = pred.predict_probs_img(predictions, proposals).clone().detach()
    self_, f = rl.make_self(k, c["labels"], img)
    prev_scores = predictions[0].detach()
    prev_boxes = [p.proposal_boxes for p in proposals]
    out = {"scores": predictions[0].detach(), "img_scores": img}
    for i in range(kr):
        targets = f["get_pgt_top_k"](self_, prev_boxes, prev_scores, proposals, suffix="_r%d" % i)
        props_k = f["label_and_sample_proposals"](self_, proposals, targets, suffix="_r%d" % i)
        predictions_k = refs[i](x)
        losses.update(refs[i].losses(predictions_k, props_k))
        prev_scores = [s.detach() for s in refs[i].predict_probs(predictions_k, props_k)]
        prev_boxes = [b.detach() for b in refs[i].predict_boxes(predictions_k, props_k)]
        out["b%d/gt_classes" % i] = torch.cat([p.gt_classes for p in props_k])
        out["b%d/gt_weights" % i] = torch.cat([p.gt_weights for p in props_k])
        out["b%d/probs" % i] = torch.cat(prev_scores)
        out["b%d/boxes" % i] = torch.cat(prev_boxes)
    sum(losses.values()).backward()
    for n, v in losses.items():
        out["loss/" + n] = v.detach()
    out["grad/x"] = x.grad
    for mod_name, mod in modules.items():
        for pn, pv in mod.named_parameters():
            if pv.grad is not None:
                out["grad/%s.%s" % (mod_name, pn)] = pv.grad
    # inference path of the last refinery: mean of the branches' softmaxes / deltas (fast_rcnn_oicr.py:712-783)
    with torch.no_grad():
        pk = [r(c["x"]) for r in refs]
        out["infer/probs_K"] = torch.cat(refs[-1].predict_probs_K(pk, proposals))
        out["infer/boxes_K"] = torch.cat(refs[-1].predict_boxes_K(pk, proposals))
    return {n: v.detach().numpy() for n, v in out.items()}


if __name__ == "__main__":
    for name in CASES:
        np.savez_compressed(os.path.join(HERE, "modules_%s.npz" % name), **reference_run(name))
        print("wrote modules_%s.npz" % name)
