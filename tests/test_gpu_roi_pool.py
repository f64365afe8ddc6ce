"""GPU parity: jtsm_b200 RoIPool (through the C ABI) vs the CPU oracle on identical seeded inputs.
Forward values and argmax are bit-exact (max is exact); backward is an fp32 sum whose order
differs from the oracle's -> rel 1e-5 (north_star tolerance)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _case(seed, n, c, h, w, r, img=(600, 800), shuffle=True):
    g = synth.generator(seed)
    x = synth.feature_map(g, n, c, h, w)
    boxes = [synth.proposals(g, r, img[0], img[1]) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    if shuffle:
        rois = rois[torch.randperm(rois.shape[0], generator=g)]
    return x, rois


@pytest.mark.parametrize(
    "n,c,h,w,r,scale,pooled,shuffle",
    [
        (1, 16, 38, 50, 300, 1 / 16.0, (7, 7), False),
        (2, 8, 75, 100, 200, 1 / 8.0, (7, 7), True),
        (3, 5, 19, 25, 100, 1 / 32.0, (14, 14), True),
        (1, 4, 75, 100, 150, 1 / 8.0, (3, 5), False),
        (1, 2, 300, 400, 64, 1 / 2.0, (7, 7), False),  # plane too big for smem -> gmem kernel
    ],
)
def test_roi_pool_fwd_bwd_matches_oracle(cuda, oracle, n, c, h, w, r, scale, pooled, shuffle):
    from jtsm_b200 import ops

    x, rois = _case(11, n, c, h, w, r, shuffle=shuffle)
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), scale, pooled)
    for adt in (torch.int32, torch.int16):
        if adt == torch.int16 and h * w >= 65535:
            continue
        out, arg = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled, argmax_dtype=adt)
        assert np.array_equal(out.cpu().numpy(), ref)
        a = arg.cpu().numpy()
        if adt == torch.int16:
            a = a.view(np.uint16).astype(np.int32)
            a[a == 0xFFFF] = -1
        assert np.array_equal(a, rarg)
        g = torch.randn(out.shape, generator=synth.generator(12))
        gi_ref = oracle.max_pool_backward(g.numpy(), rarg, rois.numpy(), tuple(x.shape))
        gi = ops.roi_pool_backward(g.to(cuda), rois.to(cuda), arg, x.shape, pooled)
        np.testing.assert_allclose(gi.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5)


def test_roi_pool_module_autograd_and_row_scale(cuda, oracle):
    from jtsm_b200.layers import RoIPool

    x, rois = _case(13, 1, 8, 38, 50, 128)
    obj = torch.rand(rois.shape[0], generator=synth.generator(14)) + 1.0
    xg = x.to(cuda).requires_grad_()
    pool = RoIPool((7, 7), 1 / 16.0)
    out = pool(xg, rois.to(cuda), row_scale=obj.to(cuda))
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    ref_scaled = ref * obj.numpy()[:, None, None, None]
    assert np.array_equal(out.detach().cpu().numpy(), ref_scaled)
    g = torch.randn(out.shape, generator=synth.generator(15))
    out.backward(g.to(cuda))
    gi_ref = oracle.max_pool_backward(g.numpy() * obj.numpy()[:, None, None, None], rarg, rois.numpy(), tuple(x.shape))
    np.testing.assert_allclose(xg.grad.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5)


def test_roi_pool_noncontiguous_grad_and_edge_rois(cuda, oracle):
    from jtsm_b200 import ops

    x = synth.feature_map(synth.generator(3), 1, 3, 10, 12)
    rois = torch.tensor(
        [[0, 5, 5, 5, 5], [0, -50, -50, -20, -20], [0, 500, 500, 600, 600], [0, 30, 40, 10, 20], [0, 0, 0, 191, 159]],
        dtype=torch.float32,
    )
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    out, arg = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), 1 / 16.0, (7, 7))
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(arg.cpu().numpy(), rarg)
    g = torch.randn(5, 3, 7, 14, generator=synth.generator(4))[:, :, :, ::2]  # stride-2 view
    gi_ref = oracle.max_pool_backward(g.contiguous().numpy(), rarg, rois.numpy(), tuple(x.shape))
    gd = torch.randn(5, 3, 7, 14, device=cuda)
    gd.copy_(torch.zeros_like(gd))
    gd[:, :, :, ::2] = g.to(cuda)
    gi = ops.roi_pool_backward(gd[:, :, :, ::2], rois.to(cuda), arg, x.shape, (7, 7))
    np.testing.assert_allclose(gi.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5)


def test_roi_pool_empty_inputs(cuda):
    from jtsm_b200 import ops

    x = torch.rand(2, 3, 8, 8, device=cuda)
    out, arg = ops.roi_pool_forward(x, torch.zeros(0, 5, device=cuda), 1.0, (7, 7))
    assert out.shape == (0, 3, 7, 7)
    gi = ops.roi_pool_backward(out, torch.zeros(0, 5, device=cuda), arg, x.shape, (7, 7))
    assert gi.shape == x.shape and float(gi.abs().sum()) == 0.0


def test_roi_pool_bf16(cuda, oracle):
    from jtsm_b200 import ops

    x, rois = _case(21, 1, 8, 38, 50, 100)
    xb = x.bfloat16()
    ref, rarg = oracle.roi_pool_forward(xb.float().numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    out, arg = ops.roi_pool_forward(xb.to(cuda), rois.to(cuda), 1 / 16.0, (7, 7))
    assert out.dtype == torch.bfloat16
    assert np.array_equal(out.float().cpu().numpy(), ref)  # max of bf16 values is exact
    assert np.array_equal(arg.cpu().numpy(), rarg)
    g = torch.randn(out.shape, generator=synth.generator(22)).bfloat16()
    gi_ref = oracle.max_pool_backward(g.float().numpy(), rarg, rois.numpy(), tuple(x.shape))
    gi = ops.roi_pool_backward(g.to(cuda), rois.to(cuda), arg, x.shape, (7, 7))
    np.testing.assert_allclose(gi.float().cpu().numpy(), gi_ref, rtol=1e-2, atol=1e-2)
