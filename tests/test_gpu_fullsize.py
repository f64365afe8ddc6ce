"""GPU parity at the FULL sizes of BASELINE.json's configurations: the CUDA path (through the C ABI) against the CPU
oracle (oracle/jtsm_oracle.c, host threads over RoI chunks) on the same seeded inputs.

  c1   1 x 512 x 38 x 50 (stride 16), 2000 RoIs, RoIPool 7x7                         values / argmax bit-exact, bwd 1e-5
  c3   2 x 2048 x 75 x 100, 8000 RoIs, RoIPool 7x7                                    values / argmax bit-exact, bwd 1e-5
  c3m  the same map, MOIPool + mask_scale x (objectness + 1) epilogue (JTSM's pooler) values / argmax bit-exact
  c5   six points of the stress sweep (stride 8/16/32, 7x7 / 14x14 bins, 1k..16k RoIs, fp32 and bf16):
       RoIPool bit-exact, RoIAlign forward bit-exact in fp32 (1e-2 in bf16), backward by tolerance

(c2 at full size: tests/test_gpu_roi_pool.py::test_roi_pool_full_config2_bit_exact; the head math at every config's
R and K: tests/test_gpu_heads.py.)  The oracle runs one RoI chunk per host thread and the comparison walks the same
chunks, so the 3.2 GB outputs of c3 never exist twice on the host."""
import math
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu

THREADS = max(1, min(32, os.cpu_count() or 1))


def _chunks(r, size):
    return [(a, min(a + size, r)) for a in range(0, r, size)]


def _decode(arg):
    a = arg.cpu().numpy()
    if a.dtype == np.int16:
        a = a.view(np.uint16).astype(np.int32)
        a[a == 0xFFFF] = -1
    return a


def _check_pool_chunks(ref_fn, out, arg, chunks, scale_rows=None):
    """ref_fn(a, b) -> (values, argmax) of RoIs [a, b); compared bit for bit with the device tensors' rows."""
    def work(ab):
        a, b = ab
        ref, rarg = ref_fn(a, b)
        if scale_rows is not None:
            ref = scale_rows(ref, rarg, a, b)
        o = out[a:b].float().cpu().numpy()
        ok_v = np.array_equal(o, ref)
        ok_a = np.array_equal(_decode(arg[a:b]), rarg)
        return ok_v, ok_a, rarg

    with ThreadPoolExecutor(THREADS) as ex:
        res = list(ex.map(work, chunks))
    assert all(r[0] for r in res), "pooled values differ from the oracle"
    assert all(r[1] for r in res), "argmax differs from the oracle"
    return [r[2] for r in res]


def _check_backward(oracle, grad_fn, rargs, chunks, rois, shape, gi_dev, row_scale=None):
    """grad_input == sum over RoI chunks of the oracle's max-pool backward (linear in the rows)."""
    def work(i):
        a, b = chunks[i]
        g = grad_fn(a, b)
        if row_scale is not None:
            g = g * row_scale[a:b, None, None, None]
        return oracle.max_pool_backward(g, rargs[i], rois[a:b], shape)

    with ThreadPoolExecutor(THREADS) as ex:
        parts = list(ex.map(work, range(len(chunks))))
    gi_ref = np.zeros(shape, np.float32)
    for p in parts:
        gi_ref += p
    np.testing.assert_allclose(gi_dev.float().cpu().numpy(), gi_ref, rtol=1e-5, atol=1e-5 * float(np.abs(gi_ref).max()))


def _tile_grad(tile, a, b):
    idx = np.arange(a, b) % tile.shape[0]
    return tile[idx]


@pytest.mark.parametrize("name", ["c1", "c3"])
def test_roi_pool_full_size_vs_oracle(cuda, oracle, name):
    from jtsm_b200 import ops

    cfg, x, _, rois, obj = synth.pool_case(name)
    P = (cfg["P"], cfg["P"])
    scale = (obj + 1.0)
    xn, rn, sn = x.numpy(), rois.numpy(), scale.numpy()
    xd, rd, sd = x.to(cuda), rois.to(cuda), scale.to(cuda)
    out, arg = ops.roi_pool_forward(xd, rd, cfg["scale"], P, row_scale=sd, argmax_dtype=torch.int16)
    chunks = _chunks(rois.shape[0], 250 if name == "c1" else 500)
    rargs = _check_pool_chunks(lambda a, b: oracle.roi_pool_forward(xn, rn[a:b], cfg["scale"], P), out, arg, chunks,
                               scale_rows=lambda ref, rarg, a, b: ref * sn[a:b, None, None, None])
    del out
    tile = torch.randn(64, cfg["C"], P[0], P[1], generator=synth.generator(5))
    g = synth.expand_grad_tile(tile.to(cuda), rois.shape[0])
    gi = ops.roi_pool_backward(g, rd, arg, x.shape, P, row_scale=sd)
    tn = tile.numpy()
    _check_backward(oracle, lambda a, b: _tile_grad(tn, a, b), rargs, chunks, rn, tuple(x.shape), gi, row_scale=sn)


def test_moi_pool_full_size_c3m_vs_oracle(cuda, oracle):
    """JTSM's pooler at config 3: 2 x 2048 x 75 x 100, 8000 RoIs, ~1000 superpixels per image, box-consistent
    oh_labels, fused mask_scale x (objectness + 1) epilogue (roi_heads_jtsm.py:607-633)."""
    from jtsm_b200 import ops

    wl = synth.step_case("c3m")
    cfg, x, rois, obj = wl["cfg"], wl["feat"], wl["rois"], wl["objectness"]
    P = (cfg["P"], cfg["P"])
    xn, rn = x.numpy(), rois.numpy()
    oh, sp = wl["oh_labels"].numpy(), wl["superpixels"].numpy()
    rs = (obj + 1.0)
    out, arg, scales = ops.moi_pool_forward(x.to(cuda), rois.to(cuda), cfg["scale"], P[0], P[1], wl["oh_labels"].to(cuda),
                                            wl["superpixels"].to(cuda), row_scale=rs.to(cuda), mask_scale=True,
                                            return_scales=True)
    rsn = rs.numpy()

    def scale_rows(ref, rarg, a, b):
        valid = (rarg[:, 0].reshape(rarg.shape[0], -1) != -1).sum(axis=1).astype(np.float32)
        ms = np.float32(P[0] * P[1]) * (np.float32(1.0) / (valid + np.float32(1.0)))  # P*P * (n + 1).reciprocal()
        return (ref * ms[:, None, None, None]) * rsn[a:b, None, None, None]

    chunks = _chunks(rois.shape[0], 250)
    _check_pool_chunks(lambda a, b: oracle.moi_pool_forward(xn, rn[a:b], oh[a:b], sp, cfg["scale"], P), out, arg, chunks,
                       scale_rows=scale_rows)
    # the scales the backward multiplies by: products of the two factors
    assert torch.equal(scales[2], scales[0] * scales[1])


C5_POINTS = [
    # stride, P, R, dtype
    (8, 7, 16000, "fp32"),
    (16, 14, 4000, "fp32"),
    (32, 7, 8000, "fp32"),
    (32, 14, 16000, "bf16"),
    (8, 14, 1000, "bf16"),
    (16, 7, 2000, "bf16"),
]


@pytest.mark.parametrize("stride,p,r,dtype", C5_POINTS)
def test_c5_sweep_points_vs_oracle(cuda, oracle, stride, p, r, dtype):
    """BASELINE.json configs[4] (RoI-op stress sweep), six points against the ORACLE (tools/sweep_c5.py compares all
    60 with torchvision's CUDA ops): C = 512 for RoIPool; RoIAlign (sampling_ratio 2, aligned) on 64 channels."""
    from jtsm_b200 import ops

    g = synth.generator(synth.seed_for(5, stride * 100 + p))
    h, w = int(math.ceil(600 / stride)), int(math.ceil(800 / stride))
    x = synth.feature_map(g, 1, 512, h, w)
    rois = synth.rois_from_boxes([synth.proposals(g, r, 600, 800)])
    if dtype == "bf16":
        x = x.bfloat16()
    xn = x.float().numpy()  # the oracle sees the values the kernel sees
    rn = rois.numpy()
    P = (p, p)
    scale = 1.0 / stride
    xd, rd = x.to(cuda), rois.to(cuda)
    out, arg = ops.roi_pool_forward(xd, rd, scale, P, argmax_dtype=torch.int16)
    chunks = _chunks(r, 125 if p == 14 else 500)
    rargs = _check_pool_chunks(lambda a, b: oracle.roi_pool_forward(xn, rn[a:b], scale, P), out, arg, chunks)
    tile = torch.randn(64, 512, p, p, generator=synth.generator(6))
    if dtype == "bf16":
        tile = tile.bfloat16()
    gi = ops.roi_pool_backward(synth.expand_grad_tile(tile.to(cuda), r), rd, arg, x.shape, P)
    tn = tile.float().numpy()
    if dtype == "fp32":
        _check_backward(oracle, lambda a, b: _tile_grad(tn, a, b), rargs, chunks, rn, tuple(x.shape), gi)
    else:  # bf16 gradients: fp32 accumulation, one rounding at the end -> 1e-2
        with ThreadPoolExecutor(THREADS) as ex:
            parts = list(ex.map(lambda i: oracle.max_pool_backward(_tile_grad(tn, *chunks[i]), rargs[i],
                                                                   rn[chunks[i][0]:chunks[i][1]], tuple(x.shape)),
                                range(len(chunks))))
        ref = np.sum(parts, axis=0)
        np.testing.assert_allclose(gi.float().cpu().numpy(), ref, rtol=1e-2, atol=1e-2 * float(np.abs(ref).max()))
    del out, arg, gi
    # RoIAlign on a 64-channel slice (the oracle's adaptive-grid forward is ~50x the pool's cost per channel)
    xa, xan = xd[:, :64].contiguous(), xn[:, :64]
    ra = 2000 if r > 2000 else r
    oa = ops.roi_align_forward(xa, rd[:ra], scale, p, p, 2, True)
    ach = _chunks(ra, 125)
    with ThreadPoolExecutor(THREADS) as ex:
        refs = list(ex.map(lambda ab: oracle.roi_align_forward(xan, rn[ab[0]:ab[1]], scale, P, 2, True), ach))
    ref = np.concatenate(refs)
    if dtype == "fp32":
        assert np.array_equal(oa.cpu().numpy(), ref)
    else:
        np.testing.assert_allclose(oa.float().cpu().numpy(), ref, rtol=1e-2, atol=1e-2)
    ga = torch.randn(ra, 64, p, p, generator=synth.generator(7))
    gia = ops.roi_align_backward(ga.to(cuda).to(x.dtype), rd[:ra], scale, p, p, 1, 64, h, w, 2, True)
    gref = oracle.roi_align_backward(ga.to(x.dtype).float().numpy(), rn[:ra], scale, (1, 64, h, w), 2, True)
    tol = 1e-5 if dtype == "fp32" else 1e-2
    np.testing.assert_allclose(gia.float().cpu().numpy(), gref, rtol=tol, atol=tol * float(np.abs(gref).max()))


def test_roi_pool_nonfinite_cells_match_the_scan(cuda, oracle):
    """NaN, +-inf and -FLT_MAX cells: a cell that cannot win the reference's strict `>` scan from -FLT_MAX is never
    selected, a bin of such cells reports (-FLT_MAX, -1), +inf wins (first occurrence).  Both range-max table formats
    (planes are sanitised when they are unpacked) and the direct scan agree with the oracle AND with torchvision's CPU op."""
    import torchvision

    from jtsm_b200 import _lib, ops

    g = synth.generator(77)
    x = synth.feature_map(g, 1, 6, 38, 50) - 0.3
    fmax = torch.finfo(torch.float32).max
    x[0, 0, 5:20, 5:30] = float("nan")
    x[0, 1, ::3, ::2] = float("-inf")
    x[0, 2, 10:12, :] = float("inf")
    x[0, 3, :, :] = -fmax
    x[0, 3, 7, 9] = float("nan")
    x[0, 4, :, :] = float("nan")
    x[0, 5, 3, 4] = float("inf")
    x[0, 5, 20, 33] = float("-inf")
    rois = synth.rois_from_boxes([synth.proposals(g, 200, 600, 800)])
    ref, rarg = oracle.roi_pool_forward(x.numpy(), rois.numpy(), 1 / 16.0, (7, 7))
    tv = torchvision.ops.roi_pool(x, rois, (7, 7), 1 / 16.0).numpy()
    assert np.array_equal(ref, tv, equal_nan=True)
    try:
        for keys, algo in ((0, 0), (1, 0), (0, 1)):
            _lib.set_tuning("roi_pool_fwd_keys", keys)
            _lib.set_tuning("roi_pool_fwd_algo", algo)
            out, arg = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), 1 / 16.0, (7, 7))
            assert np.array_equal(out.cpu().numpy(), ref, equal_nan=True), (keys, algo)
            assert np.array_equal(arg.cpu().numpy(), rarg), (keys, algo)
    finally:
        _lib.set_tuning("roi_pool_fwd_keys", 0)
        _lib.set_tuning("roi_pool_fwd_algo", 0)


def test_roi_pool_keyed_tables_equal_index_tables(cuda):
    """Size-independent property at config-2 geometry: the keyed (5 look-ups) and the index (8 look-ups) range-max
    kernels and the direct scan return identical bits, with and without the bulk-copy plane fetch, fp32 and bf16
    (planes whose byte size is not a multiple of 16 exercise the aligned-window fetch and its fall-back)."""
    from jtsm_b200 import _lib, ops

    for shape, scale in (((1, 96, 75, 100), 1 / 8.0), ((2, 37, 38, 51), 1 / 16.0)):
        g = synth.generator(91)
        x = synth.feature_map(g, *shape)
        x[:, ::5] *= 0.0  # flat planes: every comparison ties at the plane minimum
        x[:, 1::7] = torch.round(x[:, 1::7] * 2) / 2  # heavy exact ties at a few levels
        rois = synth.rois_from_boxes([synth.proposals(g, 700, 600, 800) for _ in range(shape[0])])
        for dt in (torch.float32, torch.bfloat16, torch.float16):
            xd, rd = x.to(cuda).to(dt), rois.to(cuda)
            res = {}
            try:
                for name, (keys, tma, algo, rows) in {"keyed+tma": (1, 1, 0, 0), "keyed": (1, 0, 0, 0), "index": (0, 1, 0, 0),
                                                      "index-no-tma": (0, 0, 0, 0), "index-rows": (0, 1, 0, 1),
                                                      "scan": (0, 1, 1, 0)}.items():
                    _lib.set_tuning("roi_pool_fwd_keys", keys)
                    _lib.set_tuning("roi_pool_fwd_tma", tma)
                    _lib.set_tuning("roi_pool_fwd_algo", algo)
                    _lib.set_tuning("roi_pool_fwd_rows", rows)  # row-mapped main loop (single-image RoI order only)
                    res[name] = ops.roi_pool_forward(xd, rd, scale, (7, 7))
            finally:
                _lib.set_tuning("roi_pool_fwd_keys", 0)
                _lib.set_tuning("roi_pool_fwd_tma", 1)
                _lib.set_tuning("roi_pool_fwd_algo", 0)
                _lib.set_tuning("roi_pool_fwd_rows", 0)
            for name in ("keyed+tma", "keyed", "index", "index-no-tma", "index-rows"):
                assert torch.equal(res[name][0], res["scan"][0]), (name, dt, shape)
                assert torch.equal(res[name][1], res["scan"][1]), (name, dt, shape)


@pytest.mark.parametrize("shape,scale,pooled", [((1, 64, 75, 100), 1 / 8.0, (7, 7)),    # 40 B per cell: two row bands
                                                ((2, 33, 38, 50), 1 / 16.0, (7, 7)),   # one band, bulk-copy staging
                                                ((1, 24, 19, 25), 1 / 32.0, (14, 14)),
                                                ((1, 16, 120, 67), 1 / 8.0, (7, 7))])   # odd width, three bands
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_roi_pool_packed_16bit_tables_equal_index_tables_and_scan(cuda, shape, scale, pooled, dtype):
    """16-bit feature maps take the packed (value, index) tables (4 look-ups per bin, roi_pool_rmqp.cuh): same values
    and arg-max as the index tables and the direct scan, including exact ties, -0.0 / +0.0, NaN, +-inf cells, boxes
    reaching outside the map and un-grouped RoI order."""
    from jtsm_b200 import _lib, ops

    g = synth.generator(123)
    n, c, h, w = shape
    x = synth.feature_map(g, *shape) - 0.3
    x[:, ::4] = torch.round(x[:, ::4] * 4) / 4  # exact ties after rounding to 16 bits
    x[:, 1::6] *= 0.0
    x[:, 1::6, ::3] = -0.0  # -0.0 next to +0.0: equal under `>`, first index wins
    xd = x.to(dtype)
    flat = xd.view(-1)
    pos = torch.randperm(flat.numel(), generator=g)[:600]
    flat[pos[:200]] = float("nan")
    flat[pos[200:400]] = float("-inf")
    flat[pos[400:600]] = float("inf")
    xd[:, 2] = float("nan")  # whole planes that can never win
    img_h, img_w = int(h / scale), int(w / scale)
    boxes = [synth.proposals(g, 500, img_h, img_w) for _ in range(n)]
    boxes[0][:6] = torch.tensor([[-300.0, -200.0, img_w + 400.0, img_h + 300.0], [0, 0, 1, 1], [5, 5, 5, 5],
                                 [img_w - 2.0, img_h - 2.0, img_w + 50.0, img_h + 50.0], [0, 0, img_w, 3.0],
                                 [3.0, 0, 4.0, img_h]])
    rois = synth.rois_from_boxes(boxes)
    rois = rois[torch.randperm(rois.shape[0], generator=g)]  # un-grouped batch indices
    xd, rd = xd.to(cuda), rois.to(cuda)
    res = {}
    try:
        for name, (pack, tma, algo) in {"packed": (2, 1, 0), "packed-no-tma": (2, 0, 0), "index": (0, 1, 0),
                                        "scan": (0, 1, 1)}.items():
            _lib.set_tuning("roi_pool_fwd_pack16", pack)
            _lib.set_tuning("roi_pool_fwd_tma", tma)
            _lib.set_tuning("roi_pool_fwd_algo", algo)
            res[name] = ops.roi_pool_forward(xd, rd, scale, pooled)
    finally:
        _lib.set_tuning("roi_pool_fwd_pack16", 1)
        _lib.set_tuning("roi_pool_fwd_tma", 1)
        _lib.set_tuning("roi_pool_fwd_algo", 0)
    for name in ("packed", "packed-no-tma", "index"):
        a, b = res[name][0].float(), res["scan"][0].float()
        assert torch.equal(torch.isnan(a), torch.isnan(b)), (name, dtype)
        assert torch.equal(torch.nan_to_num(a, nan=0.0), torch.nan_to_num(b, nan=0.0)), (name, dtype)  # -0.0 == +0.0
        assert torch.equal(res[name][1], res["scan"][1]), (name, dtype)
    assert int((res["packed"][1] < 0).sum()) > 0  # bins that nothing can win really occur (int32 argmax: -1)


@pytest.mark.parametrize("h,w,scale,pooled,dtype", [
    (150, 200, 1 / 8.0, (7, 7), torch.float32),    # c4's largest augmentation: 8 row bands
    (108, 144, 1 / 8.0, (7, 7), torch.bfloat16),   # 4 bands, 16-bit planes through the staging buffer
    (86, 115, 1 / 8.0, (14, 14), torch.float32),   # odd width (unaligned band starts), 14 x 14 bins
    (200, 60, 1 / 8.0, (3, 5), torch.float16),     # tall narrow map, few tall bins
])
def test_roi_pool_banded_planes_vs_oracle(cuda, oracle, h, w, scale, pooled, dtype):
    """Planes too large for shared memory are served in overlapping row bands by the same range-max kernel; boxes that
    reach far outside the map (bins taller than a band) take the global-memory scan.  Values / argmax bit-exact."""
    from jtsm_b200 import _lib, ops

    img_h, img_w = int(h / scale), int(w / scale)
    g = synth.generator(1000 + h + w)
    x = (synth.feature_map(g, 2, 5, h, w) - 0.2).to(dtype)
    boxes = [synth.proposals(g, 400, img_h, img_w) for _ in range(2)]
    extreme = torch.tensor([[0, 0, img_w - 1, img_h - 1], [-3.0 * img_w, -2.0 * img_h, 2.0 * img_w, 4.0 * img_h],
                            [10, -5.0 * img_h, 60, 6.0 * img_h], [-500, 20, img_w + 900, 40], [5, 5, 5, 5],
                            [img_w + 50, img_h + 50, img_w + 90, img_h + 90]], dtype=torch.float32)
    boxes[1][:6] = extreme
    rois = synth.rois_from_boxes(boxes)
    ref, rarg = oracle.roi_pool_forward(x.float().numpy(), rois.numpy(), scale, pooled)
    res = {}
    try:
        # band lists (default) / band filter; index, keyed and (16-bit maps) packed tables; un-grouped RoI order too
        for name, (keys, tma, algo, lists, pack) in {"bands": (0, 1, 0, 1, 0), "bands-filter": (0, 1, 0, 0, 0),
                                                     "bands-no-tma": (0, 0, 0, 1, 0), "bands-keyed": (1, 1, 0, 1, 0),
                                                     "bands-keyed-filter": (1, 1, 0, 0, 0), "bands-packed": (0, 1, 0, 1, 2),
                                                     "bands-packed-filter": (0, 1, 0, 0, 2), "scan": (0, 1, 1, 1, 0)}.items():
            _lib.set_tuning("roi_pool_fwd_keys", keys)
            _lib.set_tuning("roi_pool_fwd_tma", tma)
            _lib.set_tuning("roi_pool_fwd_algo", algo)
            _lib.set_tuning("roi_pool_fwd_bandlist", lists)
            _lib.set_tuning("roi_pool_fwd_pack16", pack)
            res[name] = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled)
        perm = torch.randperm(rois.shape[0], generator=g)
        _lib.set_tuning("roi_pool_fwd_keys", 0)
        _lib.set_tuning("roi_pool_fwd_pack16", 1)
        shuffled = ops.roi_pool_forward(x.to(cuda), rois[perm].to(cuda), scale, pooled)
    finally:
        _lib.set_tuning("roi_pool_fwd_keys", 0)
        _lib.set_tuning("roi_pool_fwd_tma", 1)
        _lib.set_tuning("roi_pool_fwd_algo", 0)
        _lib.set_tuning("roi_pool_fwd_bandlist", 1)
        _lib.set_tuning("roi_pool_fwd_pack16", 1)
    assert np.array_equal(shuffled[0].float().cpu().numpy(), ref[perm.numpy()]), "un-grouped RoIs"
    assert np.array_equal(shuffled[1].cpu().numpy(), rarg[perm.numpy()]), "un-grouped RoIs"
    for name, (out, arg) in res.items():
        assert np.array_equal(out.float().cpu().numpy(), ref), name
        assert np.array_equal(arg.cpu().numpy(), rarg), name
    # u16 argmax + row scale + backward through the banded forward's argmax
    sc = torch.rand(rois.shape[0], generator=g) + 0.5
    out, arg = ops.roi_pool_forward(x.to(cuda), rois.to(cuda), scale, pooled, row_scale=sc.to(cuda), argmax_dtype=torch.int16)
    assert np.array_equal(_decode(arg), rarg)
    if dtype == torch.float32:
        assert np.array_equal(out.cpu().numpy(), ref * sc.numpy()[:, None, None, None])
