"""GPU parity of the device-side `fast_rcnn_inference_single_image` (jtsm_fast_rcnn_inference: finite mask, background
drop, clip, threshold, ordered compaction, per-class NMS, top-k in one C-ABI call without a host sync) against
oracle/ref_heads.fast_rcnn_inference_single_image, which tests/test_tta.py pins bit-equal to the reference's own
function (fast_rcnn_oicr.py:95-163) run in place.  Every output is compared exactly: boxes, scores, classes,
pred_inds, the kept-row indices, and the number of detections."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads

pytestmark = pytest.mark.gpu


def _inputs(seed, r, k, per_class, bad_rows=0):
    g = synth.generator(seed)
    props = synth.proposals(g, r, 600, 800)
    if per_class:
        deltas = torch.randn(r, 4 * k, generator=g) * 0.08
        boxes = ref_heads.apply_deltas(deltas, props)  # some leave the image -> clipping matters
    else:
        boxes = props + torch.randn(r, 4, generator=g) * 3
    scores = torch.softmax(torch.randn(r, k + 1, generator=g) * 3, dim=1)
    if bad_rows:
        idx = torch.randperm(r, generator=g)[:bad_rows]
        boxes[idx[: bad_rows // 2], 1] = float("nan")
        scores[idx[bad_rows // 2:], k] = float("inf")  # a bad BACKGROUND score also drops the row
    return boxes, scores


def _check(cuda, boxes, scores, thresh, nms, topk):
    from jtsm_b200 import inference

    want = ref_heads.fast_rcnn_inference_single_image(boxes, scores, (600, 800), thresh, nms, topk)
    res, kept, all_s, all_b = inference.fast_rcnn_inference_single_image(boxes.to(cuda), scores.to(cuda), (600, 800),
                                                                        thresh, nms, topk)
    assert res["scores"].shape[0] == want["scores"].shape[0]
    assert torch.equal(res["pred_boxes"].cpu(), want["pred_boxes"])
    assert torch.equal(res["scores"].cpu(), want["scores"])
    assert torch.equal(res["pred_classes"].cpu(), want["pred_classes"])
    assert torch.equal(res["pred_inds"].cpu(), want["pred_inds"])
    assert torch.equal(kept.cpu(), want["kept_rois"])
    assert torch.equal(all_s.cpu()[0], scores) or bool(torch.isnan(scores).any() | torch.isinf(scores).any())
    return want["scores"].shape[0]


@pytest.mark.parametrize("r,k,per_class,bad,thresh,topk", [
    (4000, 20, True, 0, 1e-5, 100),    # config 4: 80 000 candidates, class-specific boxes
    (4000, 20, False, 0, 1e-5, 100),   # class-agnostic boxes (WSDDN / no box regression)
    (1500, 20, True, 40, 1e-3, 100),   # rows with NaN boxes / inf background scores are dropped, indices re-based
    (700, 80, True, 10, 0.02, -1),     # no top-k: every survivor comes back
    (300, 5, False, 0, 0.9999, 100),   # nothing (or almost nothing) passes the threshold
    (64, 1, False, 0, 0.0, 10),        # a single class
])
def test_device_inference_equals_cpu_reference_function(cuda, r, k, per_class, bad, thresh, topk):
    boxes, scores = _inputs(900 + r + k, r, k, per_class, bad)
    _check(cuda, boxes, scores, thresh, 0.3, topk)


def test_device_inference_padded_outputs_and_graph_capture(cuda):
    """The padded device form never synchronises: it can be captured into a CUDA graph and replayed on new inputs."""
    from jtsm_b200 import inference

    r, k = 2000, 20
    b0, s0 = _inputs(31, r, k, True)
    b1, s1 = _inputs(32, r, k, True)
    bd, sd = b0.to(cuda).clone(), s0.to(cuda).clone()
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        inference.fast_rcnn_inference_device(bd, sd, (600, 800), 1e-5, 0.3, 100)
    torch.cuda.current_stream().wait_stream(stream)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = inference.fast_rcnn_inference_device(bd, sd, (600, 800), 1e-5, 0.3, 100)
    for b, s in ((b0, s0), (b1, s1)):
        bd.copy_(b)
        sd.copy_(s)
        graph.replay()
        torch.cuda.synchronize()
        want = ref_heads.fast_rcnn_inference_single_image(b, s, (600, 800), 1e-5, 0.3, 100)
        n = int(out["num"].item())
        assert n == want["scores"].shape[0]
        assert torch.equal(out["pred_boxes"][:n].cpu(), want["pred_boxes"])
        assert torch.equal(out["pred_classes"][:n].cpu(), want["pred_classes"])
        assert bool((out["pred_classes"][n:] == -1).all())


def test_batched_nms_many_small_and_one_huge_class(cuda):
    """Segment bookkeeping of the NMS kernel: 300 classes of a few boxes each, and one class of 30 000 boxes next to
    empty class ids; keep lists bit-exact vs the CPU restatement."""
    from jtsm_b200 import ops

    g = synth.generator(77)
    for n, ncls in ((3000, 300), (30000, 1), (5000, 7)):
        boxes = synth.proposals(g, n, 600, 800)
        scores = torch.rand(n, generator=g)
        cls = torch.randint(0, ncls, (n,), generator=g) * 3 + 5  # sparse class ids
        keep = ops.batched_nms(boxes.to(cuda), scores.to(cuda), cls.to(cuda), 0.4)
        assert torch.equal(keep.cpu(), ref_heads.batched_nms(boxes, scores, cls, 0.4)), (n, ncls)
