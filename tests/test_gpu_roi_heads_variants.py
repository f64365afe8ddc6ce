"""The other two registered heads on the GPU: `WSDDNROIHeadsB200` (= the OICR heads without refinement branches,
roi_heads_wsddn.py:28-330) and the box branch of `JTSMROIHeadsB200` (roi_heads_jtsm.py:590-752: MOIPool over superpixel
sets, mask_scale x objectness fused into the pooling launch)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

gpu = pytest.mark.gpu


def _heads(kind, cuda, c=16, k=6, kr=2, d=24):
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.poolers import ROIPooler
    from jtsm_b200.roi_heads import (DiscriminativeAdaptionNeck, JTSMROIHeadsB200, OICROutputLayers, WSDDNOutputLayers,
                                     WSDDNROIHeadsB200)

    torch.manual_seed(3)
    shape = type("ShapeSpec", (), dict(channels=c, height=7, width=7, stride=None))()
    head = DiscriminativeAdaptionNeck(shape, fc_dims=(d, d))
    b2b = Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    pred = WSDDNOutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k)
    if kind == "wsddn":
        pooler = ROIPooler(output_size=7, scales=(1 / 8.0,), sampling_ratio=0, pooler_type="ROIPool")
        return WSDDNROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=head,
                                 box_predictor=pred).to(cuda)
    pooler = ROIPooler(output_size=7, scales=(1 / 8.0,), sampling_ratio=0, pooler_type="MOIPool")
    refs = [OICROutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k, refine_k=i,
                             refine_reg=[False] * kr) for i in range(kr)]
    return JTSMROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=head, box_predictor=pred,
                            refine_K=kr, refine_reg=[False] * kr, box_refinery=refs).to(cuda)


def _inputs(cuda, counts=(120, 90), c=16, k=6, with_sp=False):
    from jtsm_b200.structures import Boxes, Instances

    g = synth.generator(17)
    boxes = [synth.proposals(g, n, 600, 800) for n in counts]
    feat = synth.feature_map(g, len(counts), c, 75, 100).to(cuda).requires_grad_()
    labels = synth.image_labels(g, len(counts), k)
    sp = oh = None
    if with_sp:
        sp, oh, _, s = synth.superpixel_case(g, boxes, 600, 800, 22)
    props, tgts, off = [], [], 0
    for i, (b, l) in enumerate(zip(boxes, labels)):
        fields = dict(proposal_boxes=Boxes(b.to(cuda)), objectness_logits=torch.rand(b.shape[0], generator=g).to(cuda))
        if with_sp:
            fields["oh_labels"] = oh[off:off + b.shape[0]].to(cuda)
        props.append(Instances((600, 800), **fields))
        tgts.append(Instances((600, 800), gt_classes=l.to(cuda)))
        off += b.shape[0]
    return feat, props, tgts, (sp.to(cuda) if with_sp else None)


@gpu
def test_wsddn_heads_train_and_infer(cuda):
    heads = _heads("wsddn", cuda)
    feat, props, tgts, _ = _inputs(cuda)
    heads.train()
    for m in heads.modules():  # the neck's dropout would draw a different mask in the eager re-run below
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    _, losses = heads(None, {"conv5": feat}, props, tgts)
    assert set(losses) == {"loss_cls"}
    losses["loss_cls"].backward()
    assert feat.grad is not None and float(feat.grad.abs().sum()) > 0
    # eager definition of the same step (roi_heads_wsddn.py:300-311, fast_rcnn_wsddn.py:571-584, 346-379)
    with torch.no_grad():
        from jtsm_b200.layers import RoIPool
        from jtsm_b200.poolers import convert_boxes_to_pooler_format

        rois = convert_boxes_to_pooler_format([p.proposal_boxes for p in props])
        x = RoIPool((7, 7), 1 / 8.0)(feat.detach(), rois)
        x = x * torch.cat([p.objectness_logits + 1 for p in props]).view(-1, 1, 1, 1)
        x = heads.box_head(x)
        C, D = heads.box_predictor.cls(x), heads.box_predictor.det(x)
        n = [len(p) for p in props]
        s = torch.cat([torch.softmax(c_, 1) * torch.softmax(d_, 0) for c_, d_ in zip(C.split(n), D.split(n))])
        img = torch.cat([t.sum(0, keepdim=True) for t in s.split(n)]).clamp(1e-6, 1 - 1e-6)
        oh = torch.zeros_like(img)
        for i, t in enumerate(tgts):
            oh[i, t.gt_classes] = 1
        want = torch.nn.functional.binary_cross_entropy(img, oh)
    assert abs(float(losses["loss_cls"].detach()) - float(want)) <= 1e-5 * float(want)
    heads.eval()
    with torch.no_grad():
        inst, extra, all_scores, all_boxes = heads(None, {"conv5": feat.detach()}, props)
    assert len(inst) == 2 and extra == {} and tuple(all_scores[0].shape[-2:]) == (120, 7)


@gpu
def test_jtsm_heads_box_branch_pooling_and_step(cuda):
    heads = _heads("jtsm", cuda)
    feat, props, tgts, sp = _inputs(cuda, with_sp=True)
    heads.train()
    heads.superpixels = sp
    fused, scale = heads._pool({"conv5": feat.detach()}, props)
    # the reference's arithmetic (roi_heads_jtsm.py:597-625) on the un-fused op
    with torch.no_grad():
        out, argmax = heads.box_pooler([feat.detach()], [p.proposal_boxes for p in props],
                                       oh_labels_list=[p.oh_labels for p in props], superpixels=sp)
        mask_scale = argmax.size(2) * argmax.size(3) * (
            (argmax[:, 0, :, :] != -1).view(argmax.size(0), -1).sum(dim=1).to(dtype=torch.float32) + 1).reciprocal()
        want = out * mask_scale.view(-1, 1, 1, 1)
        want = want * torch.cat([p.objectness_logits + 1 for p in props]).view(-1, 1, 1, 1)
    assert torch.equal(fused, want)
    assert float((argmax[:, 0] == -1).float().mean()) > 0.05  # the superpixel masks really empty some bins
    _, losses = heads(None, {"conv5": feat}, props, tgts, superpixels=sp)
    assert set(losses) == {"loss_cls", "loss_cls_r0", "loss_cls_r1"}
    sum(losses.values()).backward()
    assert torch.isfinite(feat.grad).all() and float(feat.grad.abs().sum()) > 0
    assert heads.cnt_superpixels[0].shape[-1] == props[0].oh_labels.shape[1]
    assert int(heads.cnt_superpixels[0].sum()) == 600 * 800
    heads.eval()
    with torch.no_grad():
        inst, _, all_scores, _ = heads(None, {"conv5": feat.detach()}, props, superpixels=sp)
    assert len(inst) == 2 and tuple(all_scores[1].shape[-2:]) == (90, 7)


@gpu
def test_contextlocnet_heads_train_and_infer(cuda):
    """ROILoopPool -> box head on the three pools -> cls(RoI), det(frame) - det(context) (roi_heads_contextlocnet.py:
    284-335): loss equal to the eager composition over the un-fused ops, gradients reach the feature map."""
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.poolers import ROIPooler
    from jtsm_b200.roi_heads import ContextLocNetROIHeadsB200, DiscriminativeAdaptionNeck, WSDDNOutputLayers

    torch.manual_seed(4)
    c, k, d = 16, 6, 24
    shape = type("ShapeSpec", (), dict(channels=c, height=7, width=7, stride=None))()
    head = DiscriminativeAdaptionNeck(shape, fc_dims=(d, d))
    pred = WSDDNOutputLayers(head.output_shape, box2box_transform=Box2BoxTransform((10.0, 10.0, 5.0, 5.0)), num_classes=k)
    pooler = ROIPooler(output_size=7, scales=(1 / 8.0,), sampling_ratio=0, pooler_type="ROILoopPool")
    heads = ContextLocNetROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=head,
                                      box_predictor=pred).to(cuda)
    for m in heads.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    feat, props, tgts, _ = _inputs(cuda, c=c, k=k)
    heads.train()
    _, losses = heads(None, {"conv5": feat}, props, tgts)
    losses["loss_cls"].backward()
    assert float(feat.grad.abs().sum()) > 0
    with torch.no_grad():
        x = pooler([feat.detach()], [p.proposal_boxes for p in props])
        sc = torch.cat([p.objectness_logits + 1 for p in props])
        x = head(x * torch.cat([sc, sc, sc]).view(-1, 1, 1, 1))
        a, f, cx = torch.chunk(x, 3, dim=0)
        C, D = pred.cls(a), pred.det(f) - pred.det(cx)
        n = [len(p) for p in props]
        s = torch.cat([torch.softmax(c_, 1) * torch.softmax(d_, 0) for c_, d_ in zip(C.split(n), D.split(n))])
        img = torch.cat([t.sum(0, keepdim=True) for t in s.split(n)]).clamp(1e-6, 1 - 1e-6)
        oh = torch.zeros_like(img)
        for i, t in enumerate(tgts):
            oh[i, t.gt_classes] = 1
        want = torch.nn.functional.binary_cross_entropy(img, oh)
    assert abs(float(losses["loss_cls"].detach()) - float(want)) <= 1e-5 * float(want)
    heads.eval()
    with torch.no_grad():
        inst, _, all_scores, _ = heads(None, {"conv5": feat.detach()}, props)
    assert len(inst) == 2 and tuple(all_scores[0].shape[-2:]) == (120, k + 1)


@gpu
def test_cmil_heads_relabel_with_roi_label(cuda):
    """CMILROIHeadsB200: every branch's labels are ROILabel's (roi_heads_cmil.py:419-437), computed on the device from
    the branch's input scores, the proposals' IoU matrix, the image labels and the image scores -- compared with the
    restatement (oracle/ref_cmil.py, pinned to the reference's C++) under the same visiting order."""
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.layers import ROILabel
    from jtsm_b200.poolers import ROIPooler
    from jtsm_b200.roi_heads import CMILROIHeadsB200, DiscriminativeAdaptionNeck, OICROutputLayers, WSDDNOutputLayers
    from oracle import ref_cmil, ref_heads

    torch.manual_seed(6)
    c, k, d, kr = 16, 6, 24, 2
    shape = type("ShapeSpec", (), dict(channels=c, height=7, width=7, stride=None))()
    head = DiscriminativeAdaptionNeck(shape, fc_dims=(d, d))
    b2b = Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    pred = WSDDNOutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k)
    refs = [OICROutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k, refine_k=i, refine_reg=[False] * kr)
            for i in range(kr)]
    pooler = ROIPooler(output_size=7, scales=(1 / 8.0,), sampling_ratio=0, pooler_type="ROIPool")
    gen = torch.Generator(device=cuda).manual_seed(11)
    label = ROILabel(0.3, 0.3, 0.0, 32, 96, 1, 0, 1 << 20, generator=gen)
    heads = CMILROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=head, box_predictor=pred,
                             refine_K=kr, refine_reg=[False] * kr, box_refinery=refs, roi_label=label).to(cuda)
    feat, props, tgts, _ = _inputs(cuda, counts=(400,), c=c, k=k)
    heads.train()
    seen = []
    orig = label.forward

    def spy(S, U, L, CW, perm=None):
        perm = torch.randperm(S.shape[0], device=S.device, generator=gen)
        RL, RW = orig(S, U, L, CW, perm=perm)
        seen.append((S.cpu().numpy(), U.cpu().numpy(), L.cpu().numpy(), CW.cpu().numpy(), perm.cpu().numpy(),
                     RL.cpu().numpy(), RW.cpu().numpy()))
        return RL, RW

    label.forward = spy
    _, losses = heads(None, {"conv5": feat}, props, tgts)
    assert set(losses) == {"loss_cls", "loss_cls_r0", "loss_cls_r1"} and len(seen) == kr
    sum(losses.values()).backward()
    assert torch.isfinite(feat.grad).all() and float(feat.grad.abs().sum()) > 0
    boxes = props[0].proposal_boxes.tensor.cpu()
    for S, U, L, CW, perm, RL, RW in seen:
        assert np.array_equal(U, ref_heads.pairwise_iou(boxes, boxes).numpy())
        want = ref_cmil.roi_label_forward(S, U, L, CW.reshape(-1), perm, 0.3, 0.3, 0.0, 32, 96, 1)
        assert np.array_equal(RL, want["RL"]) and np.array_equal(RW, want["RW"])
        assert want["bg_rois"] == 97
    assert int(label.P[9]) == kr


@gpu
def test_pcl_heads_refinement_loss(cuda):
    """PCLROIHeadsB200 (roi_heads_pcl.py:329-349): per stream, PCL targets from the previous stream's scores + pcl_loss on
    this stream's softmax.  The first stream's loss is recomputed from the pieces that are pinned to the reference
    (host centre selection with the checker's IoU, oracle/ref_pcl.py proposal clusters, oracle C pcl_loss)."""
    from jtsm_b200 import pcl
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.poolers import ROIPooler
    from jtsm_b200.roi_heads import DiscriminativeAdaptionNeck, OICROutputLayers, PCLROIHeadsB200, WSDDNOutputLayers
    from oracle import capi, ref_heads, ref_pcl

    torch.manual_seed(9)
    c, k, d, kr = 16, 6, 24, 2
    shape = type("ShapeSpec", (), dict(channels=c, height=7, width=7, stride=None))()
    head = DiscriminativeAdaptionNeck(shape, fc_dims=(d, d))
    b2b = Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    pred = WSDDNOutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k)
    refs = [OICROutputLayers(head.output_shape, box2box_transform=b2b, num_classes=k, refine_k=i, refine_reg=[False] * kr)
            for i in range(kr)]
    pooler = ROIPooler(output_size=7, scales=(1 / 8.0,), sampling_ratio=0, pooler_type="ROIPool")
    heads = PCLROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=head, box_predictor=pred,
                            refine_K=kr, refine_reg=[False] * kr, box_refinery=refs).to(cuda)
    for m in heads.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    feat, props, tgts, _ = _inputs(cuda, counts=(500,), c=c, k=k)
    heads.train()
    _, losses = heads(None, {"conv5": feat}, props, tgts)
    assert set(losses) == {"loss_cls", "loss_cls_r0", "loss_cls_r1"}
    sum(losses.values()).backward()
    assert torch.isfinite(feat.grad).all() and float(feat.grad.abs().sum()) > 0
    # stream 0 from the pinned pieces
    with torch.no_grad():
        x, _ = heads._pool({"conv5": feat.detach()}, props)
        x = heads.box_head(x)
        mil = heads.box_predictor(x, props)[0]
        probs = torch.softmax(heads.box_refinery[0](x)[0], dim=-1)
    boxes = props[0].proposal_boxes.tensor.cpu().numpy()
    oh = np.zeros((1, k), np.float32)
    oh[0, tgts[0].gt_classes.cpu().numpy()] = 1
    cpu_iou = lambda b: ref_heads.pairwise_iou(torch.from_numpy(b), torch.from_numpy(b)).numpy()
    cb, cc, cs = pcl.graph_centers(boxes, np.clip(mil.cpu().numpy(), 1e-9, 1 - 1e-9), oh, iou_fn=cpu_iou)
    pn = np.clip(probs.cpu().numpy(), 1e-9, 1 - 1e-9)
    t = ref_pcl.proposal_clusters(boxes, cb, cc, cs, pn)
    real = np.concatenate([np.ones((1, 1), np.float32), oh], axis=1)
    _, want = capi.pcl_loss_forward(probs.cpu().numpy(), t["labels"].astype(np.float32), t["cls_loss_weights"],
                                    t["pc_labels"].astype(np.float32), t["pc_probs"], t["img_cls_loss_weights"], real)
    want = float(want)
    assert abs(float(losses["loss_cls_r0"].detach()) - want) <= 1e-4 * abs(want)
    heads.eval()
    with torch.no_grad():
        inst, _, all_scores, _ = heads(None, {"conv5": feat.detach()}, props)
    assert len(inst) == 1 and tuple(all_scores[0].shape[-2:]) == (500, k + 1)


def test_heads_register_in_a_registry():
    from jtsm_b200.roi_heads import register_all

    class Reg:
        def __init__(self):
            self.names = []

        def register(self, c):
            self.names.append(c.__name__)

    r = register_all(Reg())
    assert r.names == ["OICRROIHeadsB200", "WSDDNROIHeadsB200", "ContextLocNetROIHeadsB200", "CMILROIHeadsB200",
                       "PCLROIHeadsB200", "JTSMROIHeadsB200"]
