"""Pin the CPU oracle (oracle/jtsm_oracle.c) for the pooling ops:
 - reference golden vectors (tests/layers/test_roi_align.py:26-39 of the reference),
 - bit-equality with torchvision's CPU ops (the dependency the reference calls),
 - edge cases the reference tests (empty boxes / empty batch)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

tv = pytest.importorskip("torchvision")


def _case(seed=0, n=2, c=6, h=38, w=50, r=200):
    g = synth.generator(seed)
    x = synth.feature_map(g, n, c, h, w)
    boxes = [synth.proposals(g, r // n, 600, 800) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    perm = torch.randperm(rois.shape[0], generator=g)  # un-grouped batch indices
    return x, rois[perm]


def test_roi_align_reference_golden_vectors(oracle):
    # reference: tests/layers/test_roi_align.py:12-45
    inp = np.arange(25, dtype=np.float32).reshape(1, 1, 5, 5)
    rois = np.array([[0, 1, 1, 3, 3]], np.float32)
    old_results = [[7.5, 8, 8.5, 9], [10, 10.5, 11, 11.5], [12.5, 13, 13.5, 14], [15, 15.5, 16, 16.5]]
    correct_results = [[4.5, 5.0, 5.5, 6.0], [7.0, 7.5, 8.0, 8.5], [9.5, 10.0, 10.5, 11.0], [12.0, 12.5, 13.0, 13.5]]
    out_old = oracle.roi_align_forward(inp, rois, 1.0, (4, 4), 0, False)
    out_new = oracle.roi_align_forward(inp, rois, 1.0, (4, 4), 0, True)
    assert np.allclose(out_old[0, 0], np.array(old_results, np.float32))
    assert np.allclose(out_new[0, 0], np.array(correct_results, np.float32))


def test_roi_align_empty_box(oracle):
    # reference: tests/layers/test_roi_align.py:94-104
    inp = np.random.RandomState(0).rand(1, 1, 5, 5).astype(np.float32)
    rois = np.array([[0, 3, 3, 3, 3]], np.float32)
    out = oracle.roi_align_forward(inp, rois, 1.0, (7, 7), 0, True)
    assert (out == 0).all()
    gi = oracle.roi_align_backward(np.ones_like(out), rois, 1.0, inp.shape, 0, True)
    assert (gi == 0).all()


@pytest.mark.parametrize("scale,pooled", [(1 / 16.0, (7, 7)), (1 / 8.0, (7, 7)), (1 / 16.0, (14, 14)), (1 / 16.0, (3, 5))])
def test_roi_pool_bit_equal_torchvision(oracle, scale, pooled):
    x, rois = _case(1)
    xg = x.clone().requires_grad_()
    ref = tv.ops.roi_pool(xg, rois, pooled, scale)
    g = torch.randn(ref.shape, generator=synth.generator(5))
    ref.backward(g)
    out, arg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), scale, pooled)
    assert np.array_equal(out, ref.detach().numpy())
    gi = oracle.max_pool_backward(g.numpy(), arg, rois.numpy(), tuple(x.shape))
    assert np.array_equal(gi, xg.grad.numpy())


@pytest.mark.parametrize("aligned", [False, True])
@pytest.mark.parametrize("sampling_ratio", [0, 2])
def test_roi_align_bit_equal_torchvision(oracle, aligned, sampling_ratio):
    x, rois = _case(2, r=120)
    xg = x.clone().requires_grad_()
    ref = tv.ops.roi_align(xg, rois, (7, 7), 1 / 16.0, sampling_ratio, aligned)
    g = torch.randn(ref.shape, generator=synth.generator(6))
    ref.backward(g)
    out = oracle.roi_align_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7), sampling_ratio, aligned)
    assert np.array_equal(out, ref.detach().numpy())
    gi = oracle.roi_align_backward(g.numpy(), rois.numpy(), 1 / 16.0, tuple(x.shape), sampling_ratio, aligned)
    assert np.allclose(gi, xg.grad.numpy(), rtol=1e-6, atol=1e-6)


def test_roi_pool_degenerate_and_out_of_image(oracle):
    x = synth.feature_map(synth.generator(3), 1, 3, 10, 12)
    rois = torch.tensor(
        [[0, 5, 5, 5, 5], [0, -50, -50, -20, -20], [0, 500, 500, 600, 600], [0, 30, 40, 10, 20], [0, 0, 0, 191, 159]],
        dtype=torch.float32,
    )
    ref = tv.ops.roi_pool(x, rois, (7, 7), 1 / 16.0)
    out, arg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    assert np.array_equal(out, ref.numpy())
    assert (arg[1] == -1).all() and (out[1] == 0).all()


def test_loop_pool_roi_rows_equal_roi_pool_on_nonneg_input(oracle):
    """Rows [0,R) of ROILoopPool are plain RoIPool when the input is >= 0 (maxval init 0,
    ROILoopPool_cuda.cu:109) -- except argmax stays -1 where the max is exactly 0."""
    x, rois = _case(4, n=1, c=4, r=60)
    out3, arg3 = oracle.roi_loop_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    out, arg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    r = rois.shape[0]
    assert np.array_equal(out3[:r], out)
    pos = out > 0
    assert np.array_equal(arg3[:r][pos], arg[pos])
    assert (arg3[:r][~pos] == -1).all()
    # frame <= roi <= ... and context >= 0
    assert (out3[r:2 * r] <= out3[:r]).all()
    assert (out3[2 * r:] >= 0).all()


def test_moi_pool_full_mask_equals_roi_pool(oracle):
    """With every superpixel switched on, MOIPool's mask is the (inclusive) RoI rectangle,
    which contains every RoIPool bin cell -> identical to RoIPool."""
    x, rois = _case(7, n=2, c=3, r=40)
    sp = np.zeros((2, 600 // 4, 800 // 4), np.int32)
    oh = np.ones((rois.shape[0], 4), np.int32)
    out, arg = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh, sp, 1 / 16.0, (7, 7))
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    assert np.array_equal(out, ref) and np.array_equal(arg, rarg)
    out0, arg0 = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh * 0, sp, 1 / 16.0, (7, 7))
    assert (out0 == 0).all() and (arg0 == -1).all()
