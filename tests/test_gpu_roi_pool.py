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


def _decode_u16(arg):
    a = arg.cpu().numpy().view(np.uint16).astype(np.int32)
    a[a == 0xFFFF] = -1
    return a


def test_roi_pool_full_config2_bit_exact(cuda, oracle):
    """BASELINE.json configs[1] at full size (512 x 75 x 100, 2000 RoIs): forward bit-exact, backward 1e-5."""
    from jtsm_b200 import ops

    cfg, x, _, rois, obj = synth.pool_case("c2")
    scale = obj + 1.0
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), cfg["scale"], (7, 7))
    out, arg = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), cfg["scale"], (7, 7), row_scale=scale.to(cuda),
                                    argmax_dtype=torch.int16)
    assert np.array_equal(out.cpu().numpy(), ref * scale.numpy()[:, None, None, None])
    assert np.array_equal(_decode_u16(arg), rarg)
    g = synth.expand_grad_tile(torch.randn(64, cfg["C"], 7, 7, generator=synth.generator(3)), rois.shape[0])
    gi_ref = oracle.max_pool_backward((g * scale[:, None, None, None]).numpy(), rarg, rois.numpy(), tuple(x.shape))
    gi = ops.roi_pool_backward(g.to(cuda), rois.to(cuda), arg, x.shape, (7, 7), row_scale=scale.to(cuda))
    np.testing.assert_allclose(gi.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))


def test_roi_pool_table_kernel_equals_direct_scan(cuda):
    """Size-independent property at config-3 geometry (2 images, 4000 RoIs each, 128 of its channels): the
    range-max-table kernel and the direct-scan kernel give identical bits."""
    from jtsm_b200 import _lib, ops

    cfg, x, _, rois, _ = synth.pool_case("c3", C=128)
    xd, rd = x.to(cuda), rois.to(cuda)
    try:
        _lib.set_tuning("roi_pool_fwd_algo", 3)
        o3, a3 = ops.roi_pool_forward(xd, rd, cfg["scale"], (7, 7))
        _lib.set_tuning("roi_pool_fwd_algo", 1)
        o1, a1 = ops.roi_pool_forward(xd, rd, cfg["scale"], (7, 7))
    finally:
        _lib.set_tuning("roi_pool_fwd_algo", 0)
    assert torch.equal(o3, o1) and torch.equal(a3, a1)
    # every output is the value at its arg-max cell, and 0 / -1 only for empty bins
    flat = xd.view(xd.shape[0], xd.shape[1], -1)
    b = rd[:, 0].long()
    gathered = flat[b].gather(2, a3.clamp(min=0).view(a3.shape[0], a3.shape[1], -1).long()).view_as(o3)
    assert torch.equal(torch.where(a3 >= 0, gathered, torch.zeros_like(gathered)), o3)


def test_roi_pool_extreme_boxes(cuda, oracle):
    """Degenerate, elongated, huge, out-of-image and reversed boxes (the slow-bin list of the table kernel)."""
    from jtsm_b200 import ops

    g = synth.generator(21)
    x = synth.feature_map(g, 2, 6, 75, 100)
    boxes = torch.tensor([
        [0, 0, 0, 799, 599], [1, 0, 0, 799, 599],          # whole image
        [0, 10, 10, 10, 10], [1, 795, 595, 799, 599],      # one pixel, corner
        [0, 0, 300, 799, 303], [1, 400, 0, 403, 599],      # 1 x 100 and 75 x 1 cells
        [0, 100, 100, 700, 140], [1, 120, 50, 150, 590],   # elongated
        [0, -200, -100, 50, 60], [1, 700, 500, 1200, 900], # partly outside
        [0, 900, 700, 1000, 800], [1, -300, -300, -10, -10],  # fully outside -> empty bins
        [0, 500, 400, 100, 50], [1, 8, 8, 15, 15],         # reversed, single cell
    ], dtype=torch.float32)
    for pooled in ((7, 7), (14, 14), (2, 3)):
        ref, rarg = oracle.roi_pool_forward(x.numpy(), boxes.numpy(), 1 / 8.0, pooled)
        out, arg = ops.roi_pool_forward(x.to(cuda), boxes.to(cuda), 1 / 8.0, pooled)
        assert np.array_equal(out.cpu().numpy(), ref)
        assert np.array_equal(arg.cpu().numpy(), rarg)
