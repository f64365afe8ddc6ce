"""GPU parity of the head MODULES (jtsm_b200/roi_heads): `WSDDNOutputLayers`, `OICROutputLayers` and the predictor
chain of `OICRROIHeadsB200`, trained through with a plain `loss.backward()`, against fixtures produced by the
REFERENCE's own modules run in place (tests/golden/make_golden_modules.py): pseudo-labels bit-exact, losses, parameter
gradients and d loss / d fc7 to rel 1e-5 (+ an absolute floor of 1e-5 x the tensor's largest magnitude: the gradients
are sums of thousands of terms of both signs)."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden_modules as mg  # noqa: E402

pytestmark = pytest.mark.gpu


def _close(got, want, name, rtol=1e-5):
    want = np.asarray(want)
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * max(float(np.abs(want).max()), 1e-30), err_msg=name)


def _build(name, dev):
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.roi_heads import OICROutputLayers, OICRROIHeadsB200, WSDDNOutputLayers
    from jtsm_b200.structures import Boxes, Instances

    c = mg.case(name)
    k, kr, reg, d = c["k"], c["kr"], c["reg"], c["d"]
    b2b = Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    pred = WSDDNOutputLayers(d, box2box_transform=b2b, num_classes=k, mean_loss=True)
    refs = [OICROutputLayers(d, box2box_transform=b2b, num_classes=k, refine_k=i, refine_reg=[reg] * kr,
                             box_reg_loss_type="smooth_l1_weighted" if reg else "smooth_l1",
                             loss_weight={"loss_box_reg": 1.0}) for i in range(kr)]
    heads = OICRROIHeadsB200(num_classes=k, box_in_features=["res5"], box_pooler=None, box_head=torch.nn.Identity(),
                             box_predictor=pred, refine_K=kr, refine_reg=[reg] * kr, box_refinery=refs)
    mods = {"box_predictor": pred}
    mods.update({"box_refinery_%d" % i: r for i, r in enumerate(refs)})
    mg.load_params(mods, mg.params(name))
    heads.to(dev)
    proposals, targets = [], []
    for b, l in zip(c["boxes"], c["labels"]):
        proposals.append(Instances((600, 800), proposal_boxes=Boxes(b.to(dev)),
                                   objectness_logits=torch.zeros(b.shape[0], device=dev)))
        targets.append(Instances((600, 800), gt_classes=l.to(dev)))
    return c, heads, mods, proposals, targets


@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_head_modules_train_step_matches_reference_modules(cuda, name):
    from jtsm_b200.roi_heads import get_image_level_gt

    gold = np.load(os.path.join(os.path.dirname(mg.__file__), "modules_%s.npz" % name))
    c, heads, mods, proposals, targets = _build(name, cuda)
    heads.train()
    heads.gt_classes_img, heads.gt_classes_img_int, heads.gt_classes_img_oh = get_image_level_gt(targets, c["k"])
    x = c["x"].to(cuda).requires_grad_()
    seen = {}
    orig = heads._label_proposals

    def spy(prev_scores, prev_boxes, props, counts, suffix=""):
        out = orig(prev_scores, prev_boxes, props, counts, suffix)
        i = int(suffix[2:])
        seen["b%d/gt_classes" % i] = torch.cat([p.gt_classes for p in out]).cpu().numpy()
        seen["b%d/gt_weights" % i] = torch.cat([p.gt_weights for p in out]).cpu().numpy()
        if i > 0:
            seen["b%d/probs" % (i - 1)] = prev_scores.cpu().numpy()
            seen["b%d/boxes" % (i - 1)] = prev_boxes.cpu().numpy()
        return out

    heads._label_proposals = spy
    losses = heads._forward_predictors(x, proposals)
    sum(losses.values()).backward()

    want_losses = {n[5:] for n in gold.files if n.startswith("loss/")}
    assert set(losses) == want_losses
    for n in want_losses:
        _close(losses[n].detach().cpu().numpy().reshape(()), gold["loss/" + n], n)
    _close(heads.pred_class_img_logits.cpu().numpy(), gold["img_scores"], "img_scores")
    for n, v in seen.items():
        if n.endswith("gt_classes"):
            assert np.array_equal(v, gold[n]), n  # pseudo-labels: bit-exact
        elif n.endswith("gt_weights"):
            _close(v, gold[n], n)  # = the image score of the matched seed's class (a fixed-order sum on the device)
        else:
            _close(v, gold[n], n)
    _close(x.grad.cpu().numpy(), gold["grad/x"], "grad/x")
    checked = 0
    for mod_name, mod in mods.items():
        for pn, pv in mod.named_parameters():
            key = "grad/%s.%s" % (mod_name, pn)
            if key in gold.files:
                assert pv.grad is not None, key
                if key.endswith("det.bias"):
                    # analytically ZERO (softmax over the proposals is invariant to a per-class bias): both sides hold
                    # rounding noise of ~1e-8, five orders below the other gradients
                    assert float(np.abs(gold[key]).max()) < 1e-6 and float(pv.grad.abs().max()) < 1e-6, key
                else:
                    _close(pv.grad.cpu().numpy(), gold[key], key)
                checked += 1
            else:
                assert pv.grad is None or float(pv.grad.abs().max()) == 0.0, key  # bbox_pred without regression
    assert checked == sum(1 for n in gold.files if n.startswith("grad/")) - 1
    # device-side logging counters exist for every branch (no host sync inside the step)
    for i in range(c["kr"]):
        assert heads.last_counters["assign_r%d" % i].dtype == torch.int32
        acc = heads.last_counters["accuracy_r%d" % i].cpu().numpy()
        gt = gold["b%d/gt_classes" % i]
        assert acc[1] == int(((gt >= 0) & (gt < c["k"])).sum())


@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_head_modules_inference_means_match_reference(cuda, name):
    gold = np.load(os.path.join(os.path.dirname(mg.__file__), "modules_%s.npz" % name))
    c, heads, mods, proposals, _ = _build(name, cuda)
    heads.eval()
    with torch.no_grad():
        x = c["x"].to(cuda)
        pk = [r(x) for r in heads.box_refinery]
        probs = torch.cat(heads.box_refinery[-1].predict_probs_K(pk, proposals)).cpu().numpy()
        boxes = torch.cat(heads.box_refinery[-1].predict_boxes_K(pk, proposals)).cpu().numpy()
        _close(probs, gold["infer/probs_K"], "probs_K")
        _close(boxes, gold["infer/boxes_K"], "boxes_K")
        inst, kept, all_scores, all_boxes = heads.box_refinery[-1].inference(pk, proposals)
    assert len(inst) == len(proposals) and all(i["scores"].numel() <= 100 for i in inst)


def test_wsddn_layer_standalone_losses_and_autograd_without_fused_loss(cuda):
    """`losses()` on predictions whose forward did not know the labels (the reference's calling order) takes the
    autograd path through the unclamped image sums; same loss and gradients as the fused-loss path."""
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.roi_heads import WSDDNOutputLayers
    from jtsm_b200.structures import Boxes, Instances

    g = torch.Generator().manual_seed(5)
    layer = WSDDNOutputLayers(24, box2box_transform=Box2BoxTransform((10.0, 10.0, 5.0, 5.0)), num_classes=7).to(cuda)
    with torch.no_grad():
        layer.cls.weight.normal_(0, 0.5, generator=None)
        layer.det.weight.normal_(0, 0.5, generator=None)
    x = torch.rand(90, 24, generator=g).to(cuda)
    props = [Instances((600, 800), proposal_boxes=Boxes(torch.rand(n, 4, device=cuda) * 100)) for n in (50, 40)]
    oh = torch.zeros(2, 7, device=cuda)
    oh[0, 2] = oh[1, 5] = oh[1, 0] = 1.0
    res = []
    for fused in (True, False):
        layer.zero_grad()
        pred = layer(x, props, gt_classes_img_oh=oh if fused else None)
        loss = layer.losses(pred, props, oh)["loss_cls"]
        loss.backward()
        res.append((float(loss), layer.cls.weight.grad.clone(), layer.det.weight.grad.clone()))
    assert abs(res[0][0] - res[1][0]) <= 1e-6 * abs(res[1][0])
    for a, b in zip(res[0][1:], res[1][1:]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-7)
    # eager definition (fast_rcnn_wsddn.py:571-584, 346-379) on the same parameters
    C, D = layer.cls(x), layer.det(x)
    s = torch.cat([torch.softmax(c, 1) * torch.softmax(d, 0) for c, d in zip(C.split([50, 40]), D.split([50, 40]))])
    img = torch.cat([t.sum(0, keepdim=True) for t in s.split([50, 40])]).clamp(1e-6, 1 - 1e-6)
    want = torch.nn.functional.binary_cross_entropy(img, oh)
    assert abs(res[0][0] - float(want)) <= 1e-5 * float(want)


def test_wsddn_layer_contextlocnet_streams(cuda):
    """`forward(context=True)` (fast_rcnn_wsddn.py:560-561, 600-619): C = cls(x), D = det(frame) - det(context), then the
    usual per-image softmax product -- against the eager definition."""
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.roi_heads import WSDDNOutputLayers
    from jtsm_b200.structures import Boxes, Instances

    g = torch.Generator().manual_seed(8)
    layer = WSDDNOutputLayers(24, box2box_transform=Box2BoxTransform((10.0, 10.0, 5.0, 5.0)), num_classes=7).to(cuda)
    feats = tuple(torch.rand(90, 24, generator=g).to(cuda).requires_grad_() for _ in range(3))
    props = [Instances((600, 800), proposal_boxes=Boxes(torch.rand(n, 4, device=cuda) * 100)) for n in (50, 40)]
    scores, deltas = layer(feats, props, context=True)
    C = layer.cls(feats[0])
    D = layer.det(feats[1]) - layer.det(feats[2])
    want = torch.cat([torch.softmax(c, 1) * torch.softmax(d, 0) for c, d in zip(C.split([50, 40]), D.split([50, 40]))])
    torch.testing.assert_close(scores, want, rtol=1e-5, atol=1e-9)
    scores.sum().backward()
    assert all(f.grad is not None and float(f.grad.abs().sum()) > 0 for f in feats)
