"""GPU parity of the device-side batched NMS: keep indices bit-exact vs the CPU oracle (torchvision CPU
kernel order, per class, keep in descending-score order) and vs torchvision's CPU nms itself."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads

pytestmark = pytest.mark.gpu


def _boxes(g, n, dense=False):
    if dense:  # detectron2/utils/testing.py random_boxes: heavy overlap
        return synth.random_boxes(g, n, 100)
    return synth.proposals(g, n, 600, 800)


@pytest.mark.parametrize("n,classes,thr,dense", [(2000, 50, 0.2, True), (2000, 50, 0.5, True), (2000, 50, 0.8, True),
                                                 (4000, 1, 0.3, False), (20000, 20, 0.3, False), (1, 1, 0.5, False),
                                                 (65, 3, 0.5, True), (80000, 20, 0.3, False)])
def test_batched_nms_keep_indices_bit_exact(cuda, n, classes, thr, dense):
    from jtsm_b200 import ops

    g = synth.generator(7 + n)
    boxes = _boxes(g, n, dense)
    scores = torch.rand(n, generator=g)
    idxs = torch.randint(0, classes, (n,), generator=g)
    ref = ref_heads.batched_nms(boxes, scores, idxs, thr)
    keep = ops.batched_nms(boxes.to(cuda), scores.to(cuda), idxs.to(cuda), thr)
    assert torch.equal(keep.cpu(), ref)


def test_nms_vs_torchvision_cpu_and_ties(cuda):
    import torchvision

    from jtsm_b200 import ops

    g = synth.generator(99)
    boxes = synth.random_boxes(g, 1500, 100)
    boxes = torch.cat([boxes, boxes[:200]])  # exact duplicates
    scores = (torch.rand(boxes.shape[0], generator=g) * 50).floor() / 50  # tied scores
    ref = torchvision.ops.nms(boxes, scores, 0.5)
    keep = ops.nms(boxes.to(cuda), scores.to(cuda), 0.5)
    assert torch.equal(keep.cpu(), ref)


def test_nms_empty(cuda):
    from jtsm_b200 import ops

    keep = ops.batched_nms(torch.zeros(0, 4, device=cuda), torch.zeros(0, device=cuda),
                           torch.zeros(0, dtype=torch.int64, device=cuda), 0.5)
    assert keep.numel() == 0 and keep.dtype == torch.int64
