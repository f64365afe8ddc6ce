"""GPU parity: RoIAlign (C ABI) vs the CPU oracle (pinned to the reference's golden vectors and bit-equal
to torchvision / the reference's ROIAlign_cpu.cpp, tests/test_oracle_pool.py).
Forward: bit-exact in fp32 (same operation order, no FMA).  Backward: fp32 sum order differs -> rel 1e-5."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _case(seed, n, c, h, w, r, img=(600, 800)):
    g = synth.generator(seed)
    x = synth.feature_map(g, n, c, h, w) - 0.3  # signed values
    boxes = [synth.proposals(g, r, img[0], img[1]) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    return x, rois[torch.randperm(rois.shape[0], generator=g)]


def test_reference_golden_vectors(cuda):
    # reference: tests/layers/test_roi_align.py:12-45
    from jtsm_b200.layers import ROIAlign

    inp = torch.arange(25, dtype=torch.float32).reshape(1, 1, 5, 5).to(cuda)
    rois = torch.tensor([[0, 1, 1, 3, 3]], dtype=torch.float32, device=cuda)
    old = [[7.5, 8, 8.5, 9], [10, 10.5, 11, 11.5], [12.5, 13, 13.5, 14], [15, 15.5, 16, 16.5]]
    new = [[4.5, 5.0, 5.5, 6.0], [7.0, 7.5, 8.0, 8.5], [9.5, 10.0, 10.5, 11.0], [12.0, 12.5, 13.0, 13.5]]
    assert torch.allclose(ROIAlign((4, 4), 1.0, 0, aligned=False)(inp, rois)[0, 0].cpu(), torch.tensor(old))
    assert torch.allclose(ROIAlign((4, 4), 1.0, 0, aligned=True)(inp, rois)[0, 0].cpu(), torch.tensor(new))


def test_empty_box_and_empty_batch(cuda):
    # reference: tests/layers/test_roi_align.py:94-111
    from jtsm_b200.layers import ROIAlign

    x = torch.rand(1, 1, 5, 5, device=cuda, requires_grad=True)
    out = ROIAlign((7, 7), 1.0, 0, aligned=True)(x, torch.tensor([[0, 3, 3, 3, 3]], dtype=torch.float32, device=cuda))
    assert (out == 0).all()
    out.sum().backward()
    assert (x.grad == 0).all()
    x0 = torch.zeros(0, 3, 10, 10, device=cuda)
    out0 = ROIAlign((7, 7), 1.0, 0, aligned=True)(x0, torch.zeros(0, 5, device=cuda))
    assert out0.shape == (0, 3, 7, 7)


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled,sr,aligned", [
    (2, 6, 38, 50, 100, 1 / 16.0, (7, 7), 0, True),
    (1, 4, 75, 100, 120, 1 / 8.0, (7, 7), 0, False),
    (1, 3, 75, 100, 80, 1 / 8.0, (14, 14), 2, True),
    (3, 5, 19, 25, 60, 1 / 32.0, (3, 5), 0, True),
    (1, 2, 300, 400, 40, 1 / 2.0, (7, 7), 2, True),  # plane too big for smem -> global-memory path
])
def test_roi_align_matches_oracle(cuda, oracle, n, c, h, w, r, scale, pooled, sr, aligned):
    from jtsm_b200 import ops

    x, rois = _case(61, n, c, h, w, r)
    ref = oracle.roi_align_forward(x.numpy(), rois.numpy(), scale, pooled, sr, aligned)
    out = ops.roi_align_forward(x.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], sr, aligned)
    assert np.array_equal(out.cpu().numpy(), ref)
    g = torch.randn(out.shape, generator=synth.generator(62))
    gi_ref = oracle.roi_align_backward(g.numpy(), rois.numpy(), scale, tuple(x.shape), sr, aligned)
    gi = ops.roi_align_backward(g.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], n, c, h, w, sr, aligned)
    tol = 1e-5 * float(np.abs(gi_ref).max())
    np.testing.assert_allclose(gi.cpu().numpy(), gi_ref, rtol=1e-5, atol=tol)


def test_roi_align_module_autograd_noncontiguous_and_bf16(cuda, oracle):
    from jtsm_b200.layers import ROIAlign

    x, rois = _case(63, 1, 4, 38, 50, 64)
    xg = x.to(cuda).requires_grad_()
    out = ROIAlign((7, 7), 1 / 16.0, 0, aligned=True)(xg, rois.to(cuda))
    g = torch.randn(64, 4, 7, 14, generator=synth.generator(64))
    gd = g.to(cuda)
    out.backward(gd[:, :, :, ::2])
    gi_ref = oracle.roi_align_backward(g[:, :, :, ::2].contiguous().numpy(), rois.numpy(), 1 / 16.0, tuple(x.shape), 0, True)
    np.testing.assert_allclose(xg.grad.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))
    xb = x.bfloat16()
    # the module casts rois to the feature dtype like the reference does (roi_align.py:108)
    ref = oracle.roi_align_forward(xb.float().numpy(), rois.bfloat16().float().numpy(), 1 / 16.0, (7, 7), 0, True)
    outb = ROIAlign((7, 7), 1 / 16.0, 0, aligned=True)(xb.to(cuda), rois.to(cuda))
    assert outb.dtype == torch.bfloat16
    np.testing.assert_allclose(outb.float().cpu().numpy(), ref, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled,sr,aligned", [
    (2, 6, 38, 50, 200, 1 / 16.0, (7, 7), 2, True), (1, 37, 75, 100, 150, 1 / 8.0, (7, 7), 2, False),
    (1, 4, 19, 25, 60, 1 / 32.0, (14, 14), 1, True), (2, 3, 38, 50, 90, 1 / 16.0, (3, 5), 3, True)])
def test_roi_align_deterministic_backward_bit_equal_to_reference_cpu_order(cuda, oracle, n, c, h, w, r, scale, pooled,
                                                                           sr, aligned):
    """The segmented-reduction backward adds the taps of a cell in the order of the reference's serial CPU backward
    (ROIAlign_cpu.cpp:168-263) with the same operands -> BIT-equal to it (the oracle C restatement is itself bit-equal
    to the reference's ROIAlign_cpu.cpp built in place, tests/test_ref_native.py), and bit-identical run to run."""
    from jtsm_b200 import ops

    g = synth.generator(900 + r)
    x = synth.feature_map(g, n, c, h, w)
    boxes = [synth.proposals(g, r // n, 600, 800) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    rois = rois[torch.randperm(rois.shape[0], generator=g)].contiguous()  # un-grouped batch indices
    rois[3, 1:] = torch.tensor([-300.0, -200.0, -250.0, -150.0])  # outside the map
    go = torch.randn(rois.shape[0], c, *pooled, generator=g)
    want = oracle.roi_align_backward(go.numpy(), rois.numpy(), scale, tuple(x.shape), sr, aligned)
    a = ops.roi_align_backward(go.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], n, c, h, w, sr, aligned,
                               deterministic=True)
    b = ops.roi_align_backward(go.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], n, c, h, w, sr, aligned,
                               deterministic=True)
    assert torch.equal(a, b)
    assert np.array_equal(a.cpu().numpy(), want)
    # non-contiguous upstream gradient (explicit strides, ROIAlign_cpu.cpp:478-482)
    gnc = go.to(cuda).permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
    assert not gnc.is_contiguous()
    assert torch.equal(ops.roi_align_backward(gnc, rois.to(cuda), scale, pooled[0], pooled[1], n, c, h, w, sr, aligned,
                                              deterministic=True), a)
    # the RED kernels agree within the atomic-sum tolerance
    red = ops.roi_align_backward(go.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], n, c, h, w, sr, aligned,
                                 deterministic=False)
    np.testing.assert_allclose(red.cpu().numpy(), want, rtol=1e-5, atol=1e-5 * float(np.abs(want).max()))


def test_roi_align_deterministic_backward_follows_torch_flag_and_rejects_adaptive_grid(cuda):
    from jtsm_b200 import ops
    from jtsm_b200.layers import ROIAlign

    g = synth.generator(77)
    x = synth.feature_map(g, 1, 8, 38, 50).to(cuda).requires_grad_()
    rois = synth.rois_from_boxes([synth.proposals(g, 64, 600, 800)]).to(cuda)
    go = torch.randn(64, 8, 7, 7, generator=g).to(cuda)
    with pytest.raises(RuntimeError, match="sampling_ratio"):
        ops.roi_align_backward(go, rois, 1 / 16.0, 7, 7, 1, 8, 38, 50, 0, True, deterministic=True)
    prev = torch.are_deterministic_algorithms_enabled()
    try:
        torch.use_deterministic_algorithms(True)
        outs = []
        for _ in range(2):
            x.grad = None
            ROIAlign((7, 7), 1 / 16.0, 2, True)(x, rois).backward(go)
            outs.append(x.grad.clone())
        assert torch.equal(outs[0], outs[1])
    finally:
        torch.use_deterministic_algorithms(prev)
    want = ops.roi_align_backward(go, rois, 1 / 16.0, 7, 7, 1, 8, 38, 50, 2, True, deterministic=True)
    assert torch.equal(outs[0], want)
