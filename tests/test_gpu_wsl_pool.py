"""GPU parity of the WSL project's pooling ops (ROILoopPool, MOIPool) through the C ABI vs the CPU oracle
restated from the reference CUDA kernels (the reference has no runnable CPU path and no test for these ops:
parity unpinned by the reference itself; oracle properties are pinned in tests/test_oracle_pool.py).
Forward values / argmax: bit-exact.  Backward: fp32 atomics -> rel 1e-5."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _case(seed, n, c, h, w, r):
    g = synth.generator(seed)
    x = synth.feature_map(g, n, c, h, w)
    boxes = [synth.proposals(g, r, 600, 800) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    return g, x, rois[torch.randperm(rois.shape[0], generator=g)]


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled", [(1, 6, 38, 50, 80, 1 / 16.0, (7, 7)), (2, 4, 75, 100, 60, 1 / 8.0, (7, 7)),
                                                   (1, 3, 75, 100, 40, 1 / 8.0, (3, 5)), (1, 2, 300, 400, 24, 1 / 2.0, (7, 7))])
def test_roi_loop_pool_matches_oracle(cuda, oracle, n, c, h, w, r, scale, pooled):
    from jtsm_b200 import ops
    from jtsm_b200.layers import ROILoopPool

    g, x, rois = _case(71, n, c, h, w, r)
    ref, rarg = oracle.roi_loop_pool_forward(x.numpy(), rois.numpy(), scale, pooled)
    out, arg = ops.roi_loop_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1])
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(arg.cpu().numpy(), rarg)
    # module + autograd: backward == shared max-pool backward over 3R rows
    xg = x.to(cuda).requires_grad_()
    y = ROILoopPool(pooled, scale)(xg, rois.to(cuda))
    go = torch.randn(y.shape, generator=synth.generator(72))
    y.backward(go.to(cuda))
    gi_ref = oracle.max_pool_backward(go.numpy(), rarg, np.tile(rois.numpy(), (3, 1)), tuple(x.shape))
    np.testing.assert_allclose(xg.grad.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled,block", [(1, 6, 38, 50, 80, 1 / 16.0, (7, 7), 24), (2, 4, 75, 100, 60, 1 / 8.0, (7, 7), 40),
                                                         (1, 3, 75, 100, 40, 1 / 8.0, (14, 14), 12)])
def test_moi_pool_matches_oracle(cuda, oracle, n, c, h, w, r, scale, pooled, block):
    from jtsm_b200.layers import MOIPool

    g, x, rois = _case(73, n, c, h, w, r)
    sp, s = synth.superpixels(g, n, 600, 800, block)
    oh = synth.oh_labels(g, rois.shape[0], s)
    ref, rarg = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh.numpy(), sp.numpy(), scale, pooled)
    xg = x.to(cuda).requires_grad_()
    out, arg = MOIPool(pooled, scale)(xg, rois.to(cuda), oh.to(cuda), sp.to(cuda))
    assert arg.dtype == torch.int32
    assert np.array_equal(out.detach().cpu().numpy(), ref)
    assert np.array_equal(arg.cpu().numpy(), rarg)
    assert (rarg == -1).any() and (rarg >= 0).any()
    go = torch.randn(out.shape, generator=synth.generator(74))
    out.backward(go.to(cuda))
    gi_ref = oracle.max_pool_backward(go.numpy(), rarg, rois.numpy(), tuple(x.shape))
    np.testing.assert_allclose(xg.grad.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))


def test_wsl_pools_empty(cuda):
    from jtsm_b200 import ops

    x = torch.rand(1, 3, 8, 8, device=cuda)
    out, arg = ops.roi_loop_pool_forward(x, torch.zeros(0, 5, device=cuda), 1.0, 7, 7)
    assert out.shape == (0, 3, 7, 7) and arg.shape == (0, 3, 7, 7)
    out, arg = ops.moi_pool_forward(x, torch.zeros(0, 5, device=cuda), 1.0, 7, 7,
                                    torch.zeros(0, 4, dtype=torch.int32, device=cuda),
                                    torch.zeros(1, 16, 16, dtype=torch.int32, device=cuda))
    assert out.shape == (0, 3, 7, 7)


def test_moi_pool_fused_jtsm_scales(cuda, oracle):
    """Fused mask_scale * (objectness + 1) epilogue == the JTSM heads' eager arithmetic on the oracle's output
    (roi_heads_jtsm.py:615-633), bit for bit; backward carries the same factors."""
    from jtsm_b200.layers import MOIPool

    g, x, rois = _case(81, 2, 5, 75, 100, 70)
    sp, s = synth.superpixels(g, 2, 600, 800, 40)
    oh = synth.oh_labels(g, rois.shape[0], s, p=0.2)
    obj = torch.rand(rois.shape[0], generator=g) + 1.0
    ref, rarg = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh.numpy(), sp.numpy(), 1 / 8.0, (7, 7))
    ref, rarg = torch.from_numpy(ref), torch.from_numpy(rarg)
    mask_scale = rarg.size(2) * rarg.size(3) * ((rarg[:, 0, :, :] != -1).view(rarg.size(0), -1).sum(dim=1)
                                                .to(dtype=torch.float32) + 1).reciprocal()
    want = (ref * mask_scale.view(-1, 1, 1, 1)) * obj.view(-1, 1, 1, 1)
    assert mask_scale.min() < mask_scale.max()
    xg = x.to(cuda).requires_grad_()
    out, arg = MOIPool((7, 7), 1 / 8.0)(xg, rois.to(cuda), oh.to(cuda), sp.to(cuda), row_scale=obj.to(cuda),
                                        mask_scale=True)
    assert torch.equal(out.detach().cpu(), want) and torch.equal(arg.cpu(), rarg)
    go = torch.randn(out.shape, generator=synth.generator(82))
    out.backward(go.to(cuda))
    gi_ref = oracle.max_pool_backward((go * (mask_scale * obj).view(-1, 1, 1, 1)).numpy(), rarg.numpy(), rois.numpy(),
                                      tuple(x.shape))
    np.testing.assert_allclose(xg.grad.cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled,img,block", [
    (1, 5, 75, 100, 300, 1 / 8.0, (7, 7), (600, 800), 24),   # visit lists / row bit masks (W <= 128), RoIs wider than 32 cells
    (2, 4, 75, 100, 120, 1 / 8.0, (7, 7), (600, 800), 24),   # two images, two channels per CTA
    (1, 3, 75, 100, 50, 1 / 8.0, (1, 2), (600, 800), 24),    # bins wider than 32 cells: several 32-bit chunks per row
    (1, 3, 75, 100, 40, 1 / 8.0, (7, 7), (600, 800), 2),     # > 8 superpixels under a cell: footprint rescan
    (1, 2, 60, 150, 60, 1 / 8.0, (7, 7), (480, 1200), 24),   # W > 128: byte-mask kernel
    (1, 3, 32, 128, 60, 1 / 8.0, (7, 7), (256, 1024), 24),   # W == 128 exactly, odd channel count
])
def test_moi_pool_list_bits_and_byte_paths_equal_oracle(cuda, oracle, n, c, h, w, r, scale, pooled, img, block):
    from jtsm_b200 import _lib, ops

    g = synth.generator(75 + w)
    x = synth.feature_map(g, n, c, h, w)
    rois = synth.rois_from_boxes([synth.proposals(g, r, *img) for _ in range(n)])
    rois[0, 1:] = torch.tensor([0.0, 0.0, img[1] - 1.0, img[0] - 1.0])  # whole image
    sp, s = synth.superpixels(g, n, img[0], img[1], block)
    oh = synth.oh_labels(g, rois.shape[0], s)
    ref, rarg = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh.numpy(), sp.numpy(), scale, pooled)
    res = {}
    modes = {"gated": (0, 50), "rmq": (0, 100), "lists": (0, -1), "bits": (1, -1), "bytes": (2, -1)}
    try:
        for name, (algo, rmq) in modes.items():
            _lib.set_tuning("moi_algo", algo)
            _lib.set_tuning("moi_rmq", rmq)
            out, arg = ops.moi_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1], oh.to(cuda), sp.to(cuda))
            res[name] = (out.cpu().numpy(), arg.cpu().numpy())
    finally:
        _lib.set_tuning("moi_algo", 0)
        _lib.set_tuning("moi_rmq", 60)
    for name in modes:
        assert np.array_equal(res[name][0], ref), name
        assert np.array_equal(res[name][1], rarg), name


def test_moi_pool_range_max_mode_dense_masks_and_scales(cuda, oracle):
    """Box-consistent superpixels (most bins fully switched on -> range-max fast path) with the fused JTSM scales,
    two images with interleaved batch indices; forced range-max, forced lists and auto agree with the oracle bit for bit."""
    from jtsm_b200 import _lib, ops

    g = synth.generator(311)
    n, c, h, w = 2, 6, 75, 100
    x = synth.feature_map(g, n, c, h, w)
    boxes = [synth.proposals(g, 200, 600, 800) for _ in range(n)]
    rois = synth.rois_from_boxes(boxes)
    sp, oh, _, s = synth.superpixel_case(g, boxes, 600, 800, 22)
    perm = torch.randperm(rois.shape[0], generator=g)
    rois, oh = rois[perm].contiguous(), oh[perm].contiguous()
    obj = torch.rand(rois.shape[0], generator=g)
    ref, rarg = oracle.moi_pool_forward(x.numpy(), rois.numpy(), oh.numpy(), sp.numpy(), 1 / 8.0, (7, 7))
    valid = (rarg[:, 0].reshape(rarg.shape[0], -1) != -1).sum(1).astype(np.float32)
    ms = np.float32(49) * (np.float32(1.0) / (valid + np.float32(1.0)))
    want = (ref * ms[:, None, None, None]) * (obj.numpy() + np.float32(1.0))[:, None, None, None]
    full = (rarg != -1).mean()
    assert full > 0.5
    try:
        for rmq in (100, -1, 60):
            _lib.set_tuning("moi_rmq", rmq)
            out, arg, scales = ops.moi_pool_forward(x.to(cuda), rois.to(cuda), 1 / 8.0, 7, 7, oh.to(cuda), sp.to(cuda),
                                                    row_scale=(obj + 1.0).to(cuda), mask_scale=True, return_scales=True)
            assert np.array_equal(arg.cpu().numpy(), rarg), rmq
            assert np.array_equal(out.cpu().numpy(), want), rmq
            assert np.array_equal(scales[0].cpu().numpy(), ms)
            plain, parg = ops.moi_pool_forward(x.to(cuda), rois.to(cuda), 1 / 8.0, 7, 7, oh.to(cuda), sp.to(cuda))
            assert np.array_equal(plain.cpu().numpy(), ref) and np.array_equal(parg.cpu().numpy(), rarg)
    finally:
        _lib.set_tuning("moi_rmq", 60)


@pytest.mark.parametrize("n,c,h,w,r,scale,pooled", [(1, 6, 75, 100, 200, 1 / 8.0, (7, 7)), (2, 4, 38, 50, 120, 1 / 16.0, (7, 7)),
                                                    (2, 3, 75, 100, 60, 1 / 8.0, (14, 14)), (1, 5, 38, 50, 40, 1 / 16.0, (3, 5)),
                                                    (1, 2, 60, 150, 50, 1 / 8.0, (7, 7))])
def test_roi_loop_pool_virtual_roi_path_equals_direct_scans_and_oracle(cuda, oracle, n, c, h, w, r, scale, pooled):
    """ROILoopPool as 3R virtual RoIs on the range-max + visit-list kernels (default when W <= 128) against the direct
    scans and the oracle (itself bit-equal to the reference's kernel, tests/test_ref_native.py): values and argmax
    bit-exact, incl. extreme boxes (context rectangle leaves the map, degenerate boxes) and shuffled batch indices."""
    from jtsm_b200 import _lib, ops

    g = synth.generator(400 + r)
    img = (int(h / scale), int(w / scale))
    x = synth.feature_map(g, n, c, h, w)
    x[:, 0] = 0.0  # an all-zero channel: nothing beats the 0 the running maximum starts at -> argmax -1
    rois = synth.rois_from_boxes([synth.proposals(g, r // n, *img) for _ in range(n)])
    rois = rois[torch.randperm(rois.shape[0], generator=g)].contiguous()
    rois[0, 1:] = torch.tensor([0.0, 0.0, img[1] - 1.0, img[0] - 1.0])
    rois[1, 1:] = torch.tensor([100.0, 100.0, 100.0, 100.0])
    rois[2, 1:] = torch.tensor([-50.0, -40.0, 30.0, 20.0])
    rois[3, 1:] = torch.tensor([img[1] - 100.0, img[0] - 100.0, img[1] + 100.0, img[0] + 100.0])
    ref, rarg = oracle.roi_loop_pool_forward(x.numpy(), rois.numpy(), scale, pooled)
    try:
        for algo in (0, 1):
            _lib.set_tuning("loop_algo", algo)
            out, arg = ops.roi_loop_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled[0], pooled[1])
            assert np.array_equal(out.cpu().numpy(), ref), algo
            assert np.array_equal(arg.cpu().numpy(), rarg), algo
    finally:
        _lib.set_tuning("loop_algo", 0)
