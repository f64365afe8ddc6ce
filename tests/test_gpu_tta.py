"""GPU parity of the device-side TTA-AVG merge through the C ABI: mean boxes / scores bit-equal to the
CPU restatement of GeneralizedRCNNWithTTAAVG._get_augmented_boxes (oracle/ref_tta.py, itself bit-equal to the
reference's own method run in place: tests/test_tta.py), the committed golden fixture, transform_proposals, and the
detections that follow (threshold + per-class NMS) at config-4 size."""
import os

import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads, ref_tta

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

CASES = [  # image (h, w), orig (h, w), min_sizes, max_size, flip, R, K
    ((375, 500), (375, 500), (480, 576, 688, 864, 1200), 4000, True, 300, 20),
    ((600, 800), (480, 640), (480, 576), 4000, True, 200, 20),
    ((500, 375), (500, 375), (480, 1200), 1000, False, 150, 7),
    ((375, 500), (375, 500), (480,), 4000, False, 1, 1),  # one augmentation, one box
]


@pytest.mark.parametrize("case", CASES)
def test_tta_merge_bit_exact(cuda, case):
    from jtsm_b200 import tta

    image_hw, orig_hw, min_sizes, max_size, flip, r, k = case
    g = synth.generator(500 + r)
    tf_o = ref_tta.tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip)
    tf = tta.tta_transforms(image_hw, orig_hw, min_sizes, max_size, flip)
    pb = synth.proposals(g, r, *orig_hw)
    boxes, scores = ref_tta.synth_augmented_outputs(g, pb, tf_o, k)
    rb, rs = ref_tta.get_augmented_boxes(boxes, scores, tf_o, mean=ref_tta.mean_sequential)
    mb, ms = tta.get_augmented_boxes([b.to(cuda) for b in boxes], [s.to(cuda) for s in scores], tf)
    assert torch.equal(mb.cpu(), rb)  # bit-exact against the sequential-sum definition
    assert torch.equal(ms.cpu(), rs)
    # torch.mean as the reference calls it: identical except (at most) one ulp in ATen's scalar tail
    tb, ts = ref_tta.get_augmented_boxes(boxes, scores, tf_o)
    np.testing.assert_allclose(mb.cpu().numpy(), tb.numpy(), rtol=1e-6, atol=1e-4)
    np.testing.assert_allclose(ms.cpu().numpy(), ts.numpy(), rtol=1e-6, atol=1e-7)


def test_tta_merge_golden_fixture(cuda):
    from jtsm_b200 import tta

    z = np.load(os.path.join(GOLD, "tta_case0.npz"))
    image_hw, orig_hw = tuple(int(v) for v in z["image_hw"]), tuple(int(v) for v in z["orig_hw"])
    tf = tta.tta_transforms(image_hw, orig_hw, tuple(int(v) for v in z["min_sizes"]), int(z["max_size"]), bool(z["flip"]))
    boxes = [torch.from_numpy(b)[None].to(cuda) for b in z["boxes"]]
    scores = [torch.from_numpy(s)[None].to(cuda) for s in z["scores"]]
    mb, ms = tta.get_augmented_boxes(boxes, scores, tf)
    # the fixture holds the reference's torch.mean output: equal up to one ulp in ATen's scalar tail (ref_tta.mean_sequential)
    np.testing.assert_allclose(mb.cpu().numpy(), z["mean_boxes"], rtol=1e-6, atol=1e-4)
    np.testing.assert_allclose(ms.cpu().numpy(), z["mean_scores"], rtol=1e-6, atol=1e-7)
    assert (mb.cpu().numpy() != z["mean_boxes"]).mean() < 0.02 and (ms.cpu().numpy() != z["mean_scores"]).mean() < 0.05


def test_tta_transform_proposals_bit_exact(cuda):
    from jtsm_b200 import tta

    g = synth.generator(78)
    image_hw = (375, 500)
    tf_o = ref_tta.tta_transforms(image_hw, image_hw, (480, 864), 4000, True)
    tf = tta.tta_transforms(image_hw, image_hw, (480, 864), 4000, True)
    pb = synth.proposals(g, 4000, *image_hw)
    pb[::7, 2] = pb[::7, 0]  # empty boxes
    pb[5::11] += 600.0  # outside the image
    obj = torch.rand(4000, generator=g)
    for t_o, t in zip(tf_o, tf):
        rz = [x for x in t_o.transforms if hasattr(x, "new_h")][-1]
        new_hw = (rz.new_h, rz.new_w)
        want = ref_tta.transform_proposals(pb, obj, new_hw, t_o, 2000)
        got = tta.transform_proposals(pb.to(cuda), obj.to(cuda), new_hw, t, 2000)
        assert torch.equal(got["proposal_boxes"].cpu(), want["proposal_boxes"])
        assert torch.equal(got["objectness_logits"].cpu(), want["objectness_logits"])
        # apply_box without clipping == fvcore's apply_box
        fwd = t.apply_box(pb.to(cuda))
        assert np.array_equal(fwd.cpu().numpy(), t_o.apply_box(pb.numpy().copy()))


def test_tta_config4_detections_equal_cpu_path(cuda):
    """c4: R = 4000, K = 20, 10 augmentations (5 scales x flip): merge -> threshold 1e-5 -> per-class NMS 0.3 ->
    top 100; the kept detections are identical to the CPU path's (boxes bit-equal, same classes and RoI indices)."""
    from jtsm_b200 import tta

    image_hw = (600, 800)
    g = synth.generator(synth.seed_for(4))
    sizes = (480, 576, 688, 864, 1200)
    tf_o = ref_tta.tta_transforms(image_hw, image_hw, sizes, 4000, True)
    tf = tta.tta_transforms(image_hw, image_hw, sizes, 4000, True)
    assert len(tf) == 10
    pb = synth.proposals(g, 4000, *image_hw)
    boxes, scores = ref_tta.synth_augmented_outputs(g, pb, tf_o, 20)
    rb, rs = ref_tta.get_augmented_boxes(boxes, scores, tf_o, mean=ref_tta.mean_sequential)
    want = ref_heads.fast_rcnn_inference_single_image(rb, rs, image_hw, 1e-5, 0.3, 100)
    got = tta.inference_one_image_boxes([b.to(cuda) for b in boxes], [s.to(cuda) for s in scores], tf, image_hw,
                                        1e-5, 0.3, 100)
    assert torch.equal(got["pred_boxes"].cpu(), want["pred_boxes"])
    assert torch.equal(got["pred_classes"].cpu(), want["pred_classes"])
    assert torch.equal(got["scores"].cpu(), want["scores"])


def test_tta_merge_argument_errors(cuda):
    from jtsm_b200 import ops

    b = torch.zeros(4, 8, device=cuda)
    with pytest.raises(RuntimeError):
        ops.tta_merge([b] * 33, None, [[]] * 33)
    with pytest.raises(RuntimeError):
        ops.tta_merge([b], None, [[(ops.TTA_NOOP, 0, 0)] * 5])
    with pytest.raises(RuntimeError, match="CUDA-only"):
        ops.tta_merge([b.cpu()], None, [[]])
    out, _ = ops.tta_merge([b[:0]], None, [[]])  # empty R
    assert out.shape == (0, 8)
