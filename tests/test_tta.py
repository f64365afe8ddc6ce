"""TTA-AVG merge: pin the CPU restatement (oracle/ref_tta.py) to the reference's own code run in
place (needs /root/reference), check the committed golden fixture, and check the host mirror's transform lists."""
import os
import types

import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_loader, ref_tta

GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_ref = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present")

CASES = [  # image (h, w), orig (h, w), min_sizes, max_size, flip, R, K
    ((375, 500), (375, 500), (480, 576, 688, 864, 1200), 4000, True, 300, 20),
    ((600, 800), (480, 640), (480, 576), 4000, True, 200, 20),  # pre_tfm is a resize
    ((500, 375), (500, 375), (480, 1200), 1000, False, 150, 7),  # portrait, max_size clamp
]


def _ref_transforms(ref, image_hw, orig_hw, min_sizes, max_size, flip):
    """DatasetMapperTTAAVG.__call__ (:147-196) with the reference's own ResizeShortestEdge.get_transform."""
    from oracle.fvcore_transforms import HFlipTransform, NoOpTransform, TransformList

    img = np.zeros(tuple(image_hw) + (3,), np.uint8)
    pre = ref["ResizeTransform"](orig_hw[0], orig_hw[1], image_hw[0], image_hw[1]) if tuple(image_hw) != tuple(orig_hw) \
        else NoOpTransform()
    out = []
    for size in min_sizes:
        aug = types.SimpleNamespace(is_range=False, short_edge_length=(size, size), max_size=max_size, interp=None)
        resize = ref["get_transform"](aug, img)
        out.append(pre + TransformList([resize]))
        if flip:
            out.append(pre + TransformList([resize, HFlipTransform(resize.new_w)]))
    return out


@needs_ref
@pytest.mark.parametrize("case", CASES)
def test_restatement_equals_reference_tta(case):
    image_hw, orig_hw, min_sizes, max_size, flip, r, k = case
    ref = ref_loader.tta_reference()
    g = synth.generator(77)
    tf_ref = _ref_transforms(ref, image_hw, orig_hw, min_sizes, max_size, flip)
    tf = ref_tta.tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip)
    assert len(tf) == len(tf_ref)
    for a, b in zip(tf, tf_ref):  # same resized shapes
        assert [(type(t).__name__, getattr(t, "new_h", None), getattr(t, "new_w", None), getattr(t, "width", None))
                for t in a.transforms] == \
               [(type(t).__name__, getattr(t, "new_h", None), getattr(t, "new_w", None), getattr(t, "width", None))
                for t in b.transforms]
    pb = synth.proposals(g, r, *orig_hw)
    boxes, scores = ref_tta.synth_augmented_outputs(g, pb, tf, k)
    # the reference's own method, unbound, with _batch_inference returning the synthetic passes
    self_ = types.SimpleNamespace(_batch_inference=lambda inputs: (None, [s.clone() for s in scores],
                                                                   [b.clone() for b in boxes]))
    rb, rs, _ = ref["get_augmented_boxes"](self_, [None] * len(tf_ref), tf_ref)
    mb, ms = ref_tta.get_augmented_boxes(boxes, scores, tf)
    assert torch.equal(mb, rb) and torch.equal(ms, rs)


@needs_ref
def test_restatement_equals_reference_transform_proposals():
    ref = ref_loader.tta_reference()
    g = synth.generator(78)
    image_hw = (375, 500)
    tf_ref = _ref_transforms(ref, image_hw, image_hw, (480, 864), 4000, True)
    tf = ref_tta.tta_transforms(image_hw, image_hw, (480, 864), 4000, True)
    pb = synth.proposals(g, 400, *image_hw)
    pb[::7, 2] = pb[::7, 0]  # some empty boxes
    pb[5::11] += 600.0  # some outside the image
    obj = torch.rand(400, generator=g)
    for t_ref, t in zip(tf_ref, tf):
        new_hw = (t.transforms[-1].new_h, t.transforms[-1].new_w) if hasattr(t.transforms[-1], "new_h") else \
            (t.transforms[-2].new_h, t.transforms[-2].new_w)
        inst = ref["Instances"](image_hw)
        inst.proposal_boxes = ref["Boxes"](pb.clone())
        inst.objectness_logits = obj.clone()
        d = {"proposals": inst}
        ref["transform_proposals"](d, new_hw, t_ref, proposal_topk=300)
        mine = ref_tta.transform_proposals(pb, obj, new_hw, t, 300)
        assert torch.equal(d["proposals"].proposal_boxes.tensor, mine["proposal_boxes"])
        assert torch.equal(d["proposals"].objectness_logits, mine["objectness_logits"])


def test_golden_tta_fixture():
    z = np.load(os.path.join(GOLD, "tta_case0.npz"))
    image_hw, orig_hw = tuple(int(v) for v in z["image_hw"]), tuple(int(v) for v in z["orig_hw"])  # Python ints, as
    # image.shape gives them: a numpy int64 would turn the scale factor into a strong float64 scalar
    tf = ref_tta.tta_transforms(image_hw, orig_hw, tuple(int(v) for v in z["min_sizes"]), int(z["max_size"]),
                                bool(z["flip"]))
    boxes = [torch.from_numpy(b)[None] for b in z["boxes"]]
    scores = [torch.from_numpy(s)[None] for s in z["scores"]]
    mb, ms = ref_tta.get_augmented_boxes(boxes, scores, tf)
    assert np.array_equal(mb.numpy(), z["mean_boxes"]) and np.array_equal(ms.numpy(), z["mean_scores"])


def test_sequential_mean_vs_torch_mean():
    """The device kernel's definition (ref_tta.mean_sequential) against torch.mean as the reference calls it: at most
    one ulp apart, and only in the scalar tail of ATen's vectorised outer sum."""
    g = synth.generator(5)
    for a, n in ((10, 4000 * 84), (16, 6300), (3, 37)):
        x = torch.randn(a, n, generator=g) * 50
        m, t = ref_tta.mean_sequential(x), torch.mean(x, dim=0)
        bound = 4 * 2.0 ** -24 * x.abs().sum(0)  # a different association of the same a-term sum
        assert bool(((m - t).abs() <= bound).all())
        body = n - n % 64
        assert torch.equal(m[:body], t[:body])


def test_host_mirror_transform_lists_match_restatement():
    """jtsm_b200.tta builds the same augmentations (shapes, order) and the inverse stage lists carry the factors the
    numpy code multiplies by."""
    from jtsm_b200 import ops, tta

    for image_hw, orig_hw, min_sizes, max_size, flip, _, _ in CASES:
        a = tta.tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip)
        b = ref_tta.tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip)
        assert len(a) == len(b)
        for ta, tb in zip(a, b):
            inv = tb.inverse().transforms
            st = ta.inverse().stages()
            assert len(st) == len(inv)
            for (ty, fa, fb), t in zip(st, inv):
                name = type(t).__name__
                if name == "HFlipTransform":
                    assert ty == ops.TTA_HFLIP and fa == float(t.width)
                elif name == "ResizeTransform":
                    assert ty == ops.TTA_SCALE
                    assert fa == float(np.float32(t.new_w * 1.0 / t.w)) and fb == float(np.float32(t.new_h * 1.0 / t.h))
                else:
                    assert ty == ops.TTA_NOOP


@needs_ref
def test_oracle_inference_single_image_equals_reference():
    """oracle/ref_heads.fast_rcnn_inference_single_image vs the reference's own function (fast_rcnn_oicr.py:95-163) run
    in place, at a size where detectron2's batched_nms takes the per-class branch (>= 10 000 candidates, nms.py:24-30)."""
    from oracle import ref_heads

    m = ref_loader.install()
    g = synth.generator(91)
    r, k = 1500, 20
    pb = synth.proposals(g, r, 600, 800)
    boxes = (pb[:, None, :] + (torch.rand(r, k, 4, generator=g) - 0.5) * 8).reshape(r, 4 * k)
    scores = torch.softmax(torch.randn(r, k + 1, generator=g), dim=1)
    want, kept, _, _ = m["oicr"].fast_rcnn_inference_single_image(boxes, scores, (600, 800), 1e-5, 0.3, 100)
    got = ref_heads.fast_rcnn_inference_single_image(boxes, scores, (600, 800), 1e-5, 0.3, 100)
    assert torch.equal(got["pred_boxes"], want.pred_boxes.tensor)
    assert torch.equal(got["scores"], want.scores)
    assert torch.equal(got["pred_classes"], want.pred_classes)
    assert torch.equal(got["pred_inds"], want.pred_inds)
    assert torch.equal(got["kept_rois"], kept)
