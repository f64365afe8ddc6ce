"""Superpixel-set IoU: the memory-lean CPU restatement equals the reference's own pairwise_iou_sp run
in place (CPU, needs /root/reference); GPU kernel bit-equal to the restatement, Matcher outputs included."""
import types

import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads, ref_loader


def _case(seed, n1, n2, s, p1=0.3, p2=0.2):
    g = synth.generator(seed)
    cnt = torch.randint(1, 3000, (s,), generator=g, dtype=torch.int64)
    oh1 = (torch.rand(n1, s, generator=g) < p1).to(torch.int32)
    oh2 = (torch.rand(n2, s, generator=g) < p2).to(torch.int32)
    if n2 > 3:
        oh2[1] = 0  # a proposal without superpixels
    if n1 > 1 and n2 > 3:
        oh2[2] = oh1[1]  # a perfect match
    return oh1, oh2, cnt


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present")
@pytest.mark.parametrize("n1,n2,s", [(3, 200, 256), (7, 150, 1024), (1, 50, 37), (12, 64, 300)])
def test_restatement_equals_reference_pairwise_iou_sp(n1, n2, s):
    f = ref_loader.pairwise_iou_sp_reference()
    oh1, oh2, cnt = _case(n1 * 1000 + n2, n1, n2, s)
    want = f(types.SimpleNamespace(oh_labels=oh1), types.SimpleNamespace(oh_labels=oh2), cnt)
    got = ref_heads.pairwise_iou_sp(oh1, oh2, cnt)
    assert torch.equal(torch.nan_to_num(got, nan=-7.0), torch.nan_to_num(want, nan=-7.0))
    assert torch.equal(torch.isnan(got), torch.isnan(want))


@pytest.mark.gpu
@pytest.mark.parametrize("n1,n2,s,thr,lab", [(3, 4000, 1024, (0.5,), (0, 1)), (7, 8000, 1024, (0.5,), (0, 1)),
                                             (1, 50, 37, (0.5,), (0, 1)), (12, 333, 300, (0.3, 0.6), (0, -1, 1)),
                                             (20, 2000, 256, (0.5,), (0, 1)), (0, 100, 64, (0.5,), (0, 1))])
def test_gpu_pairwise_iou_sp_bit_exact(cuda, n1, n2, s, thr, lab):
    from jtsm_b200 import ops

    oh1, oh2, cnt = _case(n1 * 1000 + n2, n1, n2, s)
    want = ref_heads.pairwise_iou_sp(oh1, oh2, cnt)
    mi_w, ml_w = ref_heads.matcher(torch.nan_to_num(want, nan=float("nan")), thr, lab)
    iou, mi, ml = ops.pairwise_iou_sp(oh1.to(cuda), oh2.to(cuda), cnt.to(cuda), thr, lab)
    got = iou.cpu()
    assert torch.equal(torch.isnan(got), torch.isnan(want))
    assert torch.equal(torch.nan_to_num(got, nan=-7.0), torch.nan_to_num(want, nan=-7.0))
    finite = ~torch.isnan(want).any(dim=0) if n1 > 0 else torch.ones(n2, dtype=torch.bool)
    assert torch.equal(mi.cpu()[finite], mi_w[finite])
    assert torch.equal(ml.cpu()[finite], ml_w[finite])
    only = ops.pairwise_iou_sp(oh1.to(cuda), oh2.to(cuda), cnt.to(cuda))
    assert torch.equal(torch.nan_to_num(only.cpu(), nan=-7.0), torch.nan_to_num(want, nan=-7.0))
