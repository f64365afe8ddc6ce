"""ROIPooler (host mirror of wsl/modeling/poolers.py) and the inference path (predict_probs_K, per-class NMS,
top-k) on the GPU vs CPU compositions of the oracle pieces."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads

pytestmark = pytest.mark.gpu


def test_roi_pooler_single_and_multi_level(cuda, oracle):
    from jtsm_b200.poolers import ROIPooler, assign_boxes_to_levels, convert_boxes_to_pooler_format

    g = synth.generator(81)
    boxes = [synth.proposals(g, 60, 600, 800), synth.proposals(g, 40, 600, 800)]
    feats = [synth.feature_map(g, 2, 5, 150, 200), synth.feature_map(g, 2, 5, 75, 100), synth.feature_map(g, 2, 5, 38, 50)]
    rois = convert_boxes_to_pooler_format(boxes)
    assert rois.shape == (100, 5) and rois[60, 0] == 1
    # single level, ROIPool with fused objectness scaling
    obj = torch.rand(100, generator=g) + 1
    p1 = ROIPooler(7, [1 / 8.0], 0, "ROIPool")
    out = p1([feats[1].to(cuda)], [b.to(cuda) for b in boxes], row_scale=obj.to(cuda))
    ref, _ = oracle.roi_pool_forward(feats[1].numpy(), rois.numpy(), 1 / 8.0, (7, 7))
    assert np.array_equal(out.cpu().numpy(), ref * obj.numpy()[:, None, None, None])
    # three levels, ROIAlignV2: each box pooled from its FPN level
    p3 = ROIPooler(7, [1 / 4.0, 1 / 8.0, 1 / 16.0], 0, "ROIAlignV2")
    out3 = p3([f.to(cuda) for f in feats], [b.to(cuda) for b in boxes]).cpu().numpy()
    lv = assign_boxes_to_levels(boxes, 2, 4, 224, 4)
    assert set(lv.tolist()) <= {0, 1, 2} and len(set(lv.tolist())) > 1
    for level, scale in enumerate([1 / 4.0, 1 / 8.0, 1 / 16.0]):
        idx = torch.nonzero(lv == level, as_tuple=True)[0]
        r = oracle.roi_align_forward(feats[level].numpy(), rois[idx].numpy(), scale, (7, 7), 0, True)
        assert np.array_equal(out3[idx.numpy()], r)
    # ROILoopPool stacks three row blocks
    pl = ROIPooler(7, [1 / 8.0], 0, "ROILoopPool")
    assert pl([feats[1].to(cuda)], [b.to(cuda) for b in boxes]).shape == (300, 5, 7, 7)
    with pytest.raises(ValueError):
        ROIPooler(7, [1 / 8.0], 0, "ROIAlignRotated")


def test_inference_single_image_matches_cpu_composition(cuda):
    from jtsm_b200 import inference

    g = synth.generator(83)
    r, k, kr = 1500, 20, 3
    props = synth.proposals(g, r, 600, 800)
    logits = [torch.randn(r, k + 1, generator=g) * 2 for _ in range(kr)]
    deltas = [torch.randn(r, 4 * k, generator=g) * 0.05 for _ in range(kr)]
    probs_ref = sum(torch.softmax(x, -1) for x in logits) / kr
    boxes_ref = ref_heads.apply_deltas(sum(deltas) / kr, props)
    probs = inference.predict_probs_K([x.to(cuda) for x in logits])
    boxes = inference.predict_boxes_K([d.to(cuda) for d in deltas], props.to(cuda))
    np.testing.assert_allclose(probs.cpu().numpy(), probs_ref.numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(boxes.cpu().numpy(), boxes_ref.numpy(), rtol=1e-5, atol=1e-3)
    # feed identical (CPU-computed) inputs to both sides: the keep list must be bit-exact
    res, kept, all_s, all_b = inference.fast_rcnn_inference_single_image(boxes_ref.to(cuda), probs_ref.to(cuda), (600, 800),
                                                                        1e-3, 0.3, 100)
    sc = probs_ref[:, :-1]
    bx = boxes_ref.reshape(-1, 4).clone()
    bx[:, 0::2] = bx[:, 0::2].clamp(0, 800)
    bx[:, 1::2] = bx[:, 1::2].clamp(0, 600)
    bx = bx.view(r, k, 4)
    mask = sc > 1e-3
    fi = mask.nonzero()
    keep = ref_heads.batched_nms(bx[mask], sc[mask], fi[:, 1], 0.3)[:100]
    assert torch.equal(res["pred_classes"].cpu(), fi[keep][:, 1])
    assert torch.equal(kept.cpu(), fi[keep][:, 0])
    assert torch.equal(res["scores"].cpu(), sc[mask][keep])
    assert all_s.shape == (1, r, k + 1) and all_b.shape == (1, r, 4 * k)
