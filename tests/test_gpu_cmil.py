"""GPU parity of ROIMerge / ROILabel (csrc/cmil.cu, jtsm_b200/layers/roi_merge.py, roi_label.py) and the CMIL variant of
WSDDNOutputLayers against oracle/ref_cmil.py -- a restatement pinned bit for bit to the reference's own C++
(tests/test_ref_cmil.py).  Cluster ids, member counts, merged scores, gradients, labels and weights: bit-exact."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_cmil, ref_heads

pytestmark = pytest.mark.gpu


def _merge_case(seed, r, k):
    g = synth.generator(seed)
    boxes = synth.proposals(g, r, 600, 800)
    J = ref_heads.pairwise_iou(boxes, boxes)
    C = torch.randn(r, k, generator=g)
    D = torch.randn(r, k, generator=g)
    S = torch.rand(r, generator=g)
    return boxes, J, C, D, S


@pytest.mark.parametrize("seed,r,k,lam", [(1, 2000, 20, 0.0), (2, 2000, 20, 0.35), (3, 2000, 20, 0.62), (4, 150, 7, 0.2),
                                          (5, 40, 3, 0.5), (6, 1, 4, 0.3), (7, 700, 80, 0.1), (8, 201, 5, 1.0)])
def test_roi_merge_forward_backward_bit_exact(cuda, seed, r, k, lam):
    from jtsm_b200 import ops

    _, J, C, D, S = _merge_case(seed, r, k)
    want = ref_cmil.roi_merge_forward(S.numpy(), J.numpy(), C.numpy(), D.numpy(), lam)
    got = ops.roi_merge_forward(S.to(cuda), J.to(cuda), C.to(cuda), D.to(cuda), lam)
    stats = got["stats"].tolist()
    num_id = want["MC"].shape[0]
    assert stats == [num_id, want["num_top_id"], want["max_clique"], want["min_clique"]]
    assert np.array_equal(got["I"].cpu().numpy(), want["I"])
    assert np.array_equal(got["IC"].cpu().numpy(), want["IC"])
    assert np.array_equal(got["MC"][:num_id].cpu().numpy(), want["MC"])
    assert np.array_equal(got["MD"][:num_id].cpu().numpy(), want["MD"])
    assert float(got["MC"][num_id:].abs().sum()) == 0.0 and float(got["MD"][num_id:].abs().sum()) == 0.0
    if lam < 0.3 and r >= 150:
        assert want["max_clique"] > 1
    rs = np.random.RandomState(seed)
    gmc, gmd = rs.randn(num_id, k).astype(np.float32), rs.randn(num_id, k).astype(np.float32)
    wgc, wgd = ref_cmil.roi_merge_backward(gmc, gmd, want["I"], want["IC"])
    gc, gd = ops.roi_merge_backward(torch.from_numpy(gmc).to(cuda), torch.from_numpy(gmd).to(cuda), got["I"], got["IC"])
    assert np.array_equal(gc.cpu().numpy(), wgc) and np.array_equal(gd.cpu().numpy(), wgd)


def test_roi_merge_module_autograd_state_and_lambda_schedule(cuda):
    from jtsm_b200 import ops
    from jtsm_b200.layers import ROIMerge

    _, J, C, D, S = _merge_case(11, 600, 20)
    mod = ROIMerge(0, 1 << 20, 20, 100).to(cuda)
    Cg, Dg = C.to(cuda).requires_grad_(), D.to(cuda).requires_grad_()
    w = torch.randn(600, 20, generator=synth.generator(5)).to(cuda)
    sizes = []
    for it in range(3):
        lam = ref_cmil.merge_lambda(it, 100, 20)
        assert ops.roi_merge_lambda(it, 100, 20) == float(lam)
        want = ref_cmil.roi_merge_forward(S.numpy(), J.numpy(), C.numpy(), D.numpy(), lam)
        MC, MD = mod(S.to(cuda).view(-1, 1), J.to(cuda), Cg, Dg)
        assert MC.shape == want["MC"].shape and np.array_equal(MC.detach().cpu().numpy(), want["MC"])
        sizes.append(MC.shape[0])
        Cg.grad = Dg.grad = None
        ((MC * w[:MC.shape[0]]).sum() + (MD * w[:MD.shape[0]]).sum() * 2.0).backward()
        wgc, wgd = ref_cmil.roi_merge_backward(w[:MC.shape[0]].cpu().numpy(), 2.0 * w[:MC.shape[0]].cpu().numpy(),
                                               want["I"], want["IC"])
        assert np.array_equal(Cg.grad.cpu().numpy(), wgc) and np.array_equal(Dg.grad.cpu().numpy(), wgd)
    p = mod.P.tolist()
    assert p[2] == 3 and p[3] == 20 and p[4] == 100 and p[5] > 0 and p[6] >= p[7] > 0
    assert sizes[0] <= sizes[-1]  # lambda grows with the iteration: fewer boxes merge
    assert "cur_iter=3" in repr(mod)
    sd = mod.state_dict()
    fresh = ROIMerge(0, 1, 1, 1).to(cuda)
    fresh.load_state_dict(sd)
    MC2, _ = fresh(S.to(cuda).view(-1, 1), J.to(cuda), C.to(cuda), D.to(cuda))
    want = ref_cmil.roi_merge_forward(S.numpy(), J.numpy(), C.numpy(), D.numpy(), ref_cmil.merge_lambda(3, 100, 20))
    assert np.array_equal(MC2.cpu().numpy(), want["MC"])


@pytest.mark.parametrize("seed,r,k,ks,top_k,with_cw,thr,ident", [
    (1, 2000, 20, 21, 1, True, (0.6, 0.4, 0.1), False), (2, 2000, 20, 20, 1, True, (0.3, 0.3, 0.0), False),
    (3, 90, 6, 6, 2, True, (0.6, 0.4, 0.1), False), (4, 700, 20, 20, 1, False, (0.2, 0.2, 0.05), False),
    (5, 30, 4, 4, 1, True, (0.6, 0.4, 0.1), True), (6, 1500, 20, 20, 3, True, (0.1, 0.5, 0.0), False),
    (7, 3000, 80, 81, 1, True, (0.25, 0.25, 0.0), True)])
def test_roi_label_bit_exact(cuda, seed, r, k, ks, top_k, with_cw, thr, ident):
    from jtsm_b200 import ops

    g = synth.generator(100 + seed)
    boxes = synth.proposals(g, r, 600, 800)
    U = ref_heads.pairwise_iou(boxes, boxes)
    S = torch.rand(r, ks, generator=g)
    L = torch.zeros(k)
    L[torch.randperm(k, generator=g)[:3]] = 1
    CW = torch.rand(k, generator=g) if with_cw else None
    perm = torch.arange(r) if ident else torch.randperm(r, generator=g)
    want = ref_cmil.roi_label_forward(S.numpy(), U.numpy(), L.numpy(), CW.numpy() if with_cw else None, perm.numpy(),
                                      thr[0], thr[1], thr[2], 32, 96, top_k)
    rl, rw, stats = ops.roi_label_forward(S.to(cuda), U.to(cuda), L.to(cuda).view(1, -1),
                                          CW.to(cuda).view(1, -1) if with_cw else None,
                                          None if ident else perm.to(cuda), thr[0], thr[1], thr[2], 32, 96, top_k)
    assert np.array_equal(rl.cpu().numpy(), want["RL"])
    assert np.array_equal(rw.cpu().numpy(), want["RW"])
    st = stats.tolist()
    assert (int(st[0]), int(st[1])) == (want["fg_rois"], want["bg_rois"])
    np.testing.assert_allclose(st[2:], [want["fg_weight"], want["bg_weight"]], rtol=1e-5)
    if thr[0] < 0.5 and r >= 700:
        assert want["bg_rois"] == 97 or want["fg_rois"] == 33


def test_roi_label_module_reproducible_and_counts(cuda):
    from jtsm_b200.layers import ROILabel

    g = synth.generator(9)
    r, k = 1200, 20
    boxes = synth.proposals(g, r, 600, 800)
    U = ref_heads.pairwise_iou(boxes, boxes).to(cuda)
    S = torch.rand(r, k + 1, generator=g).to(cuda)
    L = torch.zeros(1, k)
    L[0, [2, 7]] = 1
    CW = torch.rand(1, k, generator=g)
    outs = []
    for _ in range(2):
        gen = torch.Generator(device=cuda).manual_seed(77)
        mod = ROILabel(0.3, 0.3, 0.0, 32, 96, 1, 0, 1 << 20, generator=gen).to(cuda)
        RL, RW = mod(S, U, L.to(cuda), CW.to(cuda))
        outs.append((RL.clone(), RW.clone()))
        p = mod.P.tolist()
        assert p[9] == 1 and p[11] == 97 and 0 < p[10] <= 33
        assert int((RL == k).sum()) == 97 and int((RW > 0).sum()) == int(p[10]) + 97
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert RL.dtype == torch.int32 and set(RL.unique().tolist()) <= {2, 7, k}


def test_wsddn_layer_cmil_training_forward_matches_composition(cuda):
    """WSDDNOutputLayers(cmil=True).forward in training mode (fast_rcnn_wsddn.py:567-568, 621-667): the merged C / D rows
    come from ROIMerge on (row sums of the scores, pairwise IoU); scores of the merged rows, the un-merged scores and
    gradients through both."""
    from jtsm_b200.box_regression import Box2BoxTransform
    from jtsm_b200.roi_heads import WSDDNOutputLayers
    from jtsm_b200.structures import Boxes, Instances

    g = synth.generator(21)
    r, k, d = 500, 20, 32
    layer = WSDDNOutputLayers(d, box2box_transform=Box2BoxTransform((10.0, 10.0, 5.0, 5.0)), num_classes=k, cmil=True,
                              display=1 << 20, max_epoch=20, size_epoch=100).to(cuda)
    layer.train()
    boxes = synth.proposals(g, r, 600, 800)
    x = torch.randn(r, d, generator=g).to(cuda).requires_grad_()
    props = [Instances((600, 800), proposal_boxes=Boxes(boxes.to(cuda)))]
    scores, deltas, oscores, odeltas = layer(x, props)
    with torch.no_grad():
        C, D = layer.cls(x), layer.det(x)
        eager = torch.softmax(C, 1) * torch.softmax(D, 0)
    torch.testing.assert_close(oscores, eager, rtol=1e-5, atol=1e-9)
    J = ref_heads.pairwise_iou(boxes, boxes)
    want = ref_cmil.roi_merge_forward(oscores.detach().sum(1).cpu().numpy(), J.numpy(), C.cpu().numpy(), D.cpu().numpy(),
                                      ref_cmil.merge_lambda(0, 100, 20))
    mc, md = torch.from_numpy(want["MC"]).to(cuda), torch.from_numpy(want["MD"]).to(cuda)
    assert scores.shape == (mc.shape[0], k) and mc.shape[0] < r and deltas.shape == (mc.shape[0], 4 * k)
    torch.testing.assert_close(scores, torch.softmax(mc, 1) * torch.softmax(md, 0), rtol=1e-5, atol=1e-9)
    oh = torch.zeros(1, k, device=cuda)
    oh[0, 3] = 1
    loss = layer.losses((scores, deltas), props, oh)["loss_cls"] + oscores.sum() * 0.1
    loss.backward()
    assert x.grad is not None and torch.isfinite(x.grad).all() and float(x.grad.abs().sum()) > 0
    # eager autograd of the same composition (cluster assignment fixed to the restatement's)
    x2 = x.detach().clone().requires_grad_()
    C2, D2 = layer.cls(x2), layer.det(x2)
    I = torch.from_numpy(want["I"]).long().to(cuda)
    cnt = torch.from_numpy(want["IC"]).to(cuda)[:mc.shape[0]].float().view(-1, 1)
    MC2 = torch.zeros_like(mc).index_add_(0, I, C2) / cnt
    MD2 = torch.zeros_like(md).index_add_(0, I, D2) / cnt
    s2 = torch.softmax(MC2, 1) * torch.softmax(MD2, 0)
    img = s2.sum(0, keepdim=True).clamp(1e-6, 1 - 1e-6)
    l2 = torch.nn.functional.binary_cross_entropy(img, oh) + (torch.softmax(C2, 1) * torch.softmax(D2, 0)).sum() * 0.1
    l2.backward()
    assert abs(float(loss) - float(l2)) <= 1e-5 * abs(float(l2))
    torch.testing.assert_close(x.grad, x2.grad, rtol=1e-4, atol=1e-6 * float(x2.grad.abs().max()))
    layer.eval()
    assert len(layer(x.detach(), props)) == 2  # inference: the plain 2-tuple, no merge
