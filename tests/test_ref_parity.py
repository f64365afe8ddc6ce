"""Pin oracle/ref_heads.py (the restatement that travels to the GPU box) against the reference's OWN
code executed in place from /root/reference (oracle/ref_loader.py).  Skipped where the checkout is absent."""
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads, ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference checkout not present")


def small_case(seed, counts=(50, 70), k=6, kr=2, sigma=3.0):
    g = synth.generator(seed)
    boxes = [synth.proposals(g, c, 600, 800) for c in counts]
    r = sum(counts)
    labels = synth.image_labels(g, len(counts), k)
    return dict(mil_cls=torch.randn(r, k, generator=g) * sigma, mil_det=torch.randn(r, k, generator=g) * sigma,
                ref_logits=[torch.randn(r, k + 1, generator=g) * 2 for _ in range(kr)],
                ref_deltas=[torch.randn(r, 4 * k, generator=g) * 0.1 for _ in range(kr)],
                proposal_boxes=boxes, labels=labels, num_classes=k)


def assert_same(a, b, exact=False):
    if exact:
        assert torch.equal(a, b)
    else:
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("reg", [False, True])
@pytest.mark.parametrize("mean_loss", [True, False])
def test_head_step_matches_reference(reg, mean_loss):
    from oracle import ref_drive

    case = small_case(5 + int(reg))
    ref = ref_drive.reference_head_step(**case, reg=reg, mean_loss=mean_loss)
    out = ref_heads.head_step(**case, reg=reg, mean_loss=mean_loss)
    assert set(ref["losses"]) == set(out["losses"])
    for k in ref["losses"]:
        assert_same(ref["losses"][k], out["losses"][k])
    assert_same(ref["scores"], out["scores"], exact=True)
    assert_same(ref["img_scores"], out["img_scores"], exact=True)
    for rb, ob in zip(ref["branches"], out["branches"]):
        for (rs, rbx, rc, rw), (os_, obx, oc, ow) in zip(rb["pgt"], ob["pgt"]):
            assert torch.equal(rs, os_) and torch.equal(rbx, obx) and torch.equal(rc, oc) and torch.equal(rw, ow)
        assert torch.equal(rb["gt_classes"], ob["gt_classes"])
        assert torch.equal(rb["gt_weights"], ob["gt_weights"])
        assert torch.equal(rb["gt_boxes"], ob["gt_boxes"])
        assert torch.equal(rb["probs"], ob["probs"])
        assert torch.equal(rb["boxes"], ob["boxes"])
    assert_same(ref["grad_cls"], out["grad_cls"])
    assert_same(ref["grad_det"], out["grad_det"])
    for a, b in zip(ref["grad_ref_logits"], out["grad_ref_logits"]):
        assert_same(a, b)
    for a, b in zip(ref["grad_ref_deltas"], out["grad_ref_deltas"]):
        assert_same(a, b)


def test_ignore_band_matcher_matches_reference():
    from oracle import ref_drive

    case = small_case(9, counts=(120,), k=5, kr=1)
    kw = dict(thresholds=(0.3, 0.6), iou_labels=(0, -1, 1))
    ref = ref_drive.reference_head_step(**case, **kw)
    out = ref_heads.head_step(**case, **kw)
    assert torch.equal(ref["branches"][0]["gt_classes"], out["branches"][0]["gt_classes"])
    assert (out["branches"][0]["gt_classes"] == -1).any()
    assert_same(ref["losses"]["loss_cls_r0"], out["losses"]["loss_cls_r0"])


def test_reference_golden_vectors_iou_and_matcher():
    # tests/structures/test_boxes.py:150-176 and tests/modeling/test_matcher.py:14-27 of the reference
    b1 = torch.tensor([[0.0, 0.0, 1.0, 1.0], [0.0, 0.0, 1.0, 1.0]])
    b2 = torch.tensor([[0.0, 0.0, 1.0, 1.0], [0.0, 0.0, 0.5, 1.0], [0.0, 0.0, 1.0, 0.5], [0.0, 0.0, 0.5, 0.5],
                       [0.5, 0.5, 1.0, 1.0], [0.5, 0.5, 1.5, 1.5]])
    expected = torch.tensor([[1.0, 0.5, 0.5, 0.25, 0.25, 0.25 / (2 - 0.25)]] * 2)
    assert torch.allclose(ref_heads.pairwise_iou(b1, b2), expected)
    m = ref_loader.install()
    assert torch.equal(m["boxes"].pairwise_iou(m["boxes"].Boxes(b1), m["boxes"].Boxes(b2)), ref_heads.pairwise_iou(b1, b2))
    q = torch.tensor([[0.15, 0.45, 0.2, 0.6], [0.3, 0.65, 0.05, 0.1], [0.05, 0.4, 0.25, 0.4]])
    matches, labels = ref_heads.matcher(q, (0.3, 0.7), (0, -1, 1))
    rm, rl_ = m["matcher"].Matcher([0.3, 0.7], [0, -1, 1], allow_low_quality_matches=False)(q)
    assert torch.equal(matches, rm) and torch.equal(labels, rl_)
    assert matches.tolist() == [1, 1, 2, 0]


def test_box2box_round_trip_and_reference():
    m = ref_loader.install()
    g = synth.generator(3)
    src = synth.random_boxes(g, 40)
    dst = synth.random_boxes(g, 40)
    t = m["box_regression"].Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    d = ref_heads.get_deltas(src, dst)
    assert torch.equal(d, t.get_deltas(src, dst))
    assert torch.equal(ref_heads.apply_deltas(d, src), t.apply_deltas(d, src))
    assert torch.allclose(ref_heads.apply_deltas(d, src), dst, atol=1e-3)


@pytest.mark.parametrize("mode", ["top_k3", "top_frac_thres", "mist"])
def test_pgt_variants_match_reference(mode):
    """General get_pgt_top_k (top_k > 1, fractional top_k + thres) and get_pgt_mist of the restatement vs the
    reference's own methods run in place (roi_heads_oicr.py:549-591, 675-811), then label_and_sample_proposals."""
    from oracle import ref_drive  # noqa: F401  (installs the stubs)

    g = synth.generator(31)
    counts, K = (120, 90), 7
    boxes = [synth.proposals(g, c, 600, 800) for c in counts]
    labels = synth.image_labels(g, len(counts), K)
    scores = torch.rand(sum(counts), K + 1, generator=g)
    img = torch.rand(len(counts), K, generator=g)
    m = ref_loader.install()
    Boxes, Instances = m["boxes"].Boxes, m["instances"].Instances
    proposals = []
    for b in boxes:
        inst = Instances((600, 800))
        inst.proposal_boxes = Boxes(b.clone())
        inst.objectness_logits = torch.zeros(b.shape[0])
        proposals.append(inst)
    self_, f = ref_loader.make_self(K, labels, img)
    pb = [p.proposal_boxes for p in proposals]
    if mode == "top_k3":
        targets = f["get_pgt_top_k"](self_, pb, scores, proposals, top_k=3)
        mine = ref_heads.get_pgt_top_k(boxes, list(scores.split(list(counts))), labels, img, K, top_k=3)
    elif mode == "top_frac_thres":
        targets = f["get_pgt_top_k"](self_, pb, scores, proposals, top_k=0.1, thres=0.9)
        mine = ref_heads.get_pgt_top_k(boxes, list(scores.split(list(counts))), labels, img, K, top_k=0.1, thres=0.9)
    else:
        targets = f["get_pgt_mist"](self_, pb, scores, proposals)
        mine = ref_heads.get_pgt_mist(boxes, list(scores.split(list(counts))), labels, img, K)
    for t, (sc, bx, cl, wt) in zip(targets, mine):
        assert torch.equal(t.gt_scores, sc) and torch.equal(t.gt_boxes.tensor, bx)
        assert torch.equal(t.gt_classes, cl) and torch.equal(t.gt_weights, wt)
    props = f["label_and_sample_proposals"](self_, proposals, targets)
    for p, seeds, b in zip(props, mine, boxes):
        lab = ref_heads.label_proposals(seeds, b, K)
        assert torch.equal(p.gt_classes, lab["gt_classes"]) and torch.equal(p.gt_weights, lab["gt_weights"])
        assert torch.equal(p.gt_boxes.tensor, lab["gt_boxes"])
