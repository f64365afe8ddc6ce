"""Allocation-free entry point `jtsm_refine_ce_static` (include/jtsm_b200.h; no counterpart in the reference):
called through the C ABI and compared with the existing `jtsm_refine_ce` on the
same inputs, on two calls in a row (the static scratch keeps the first call's contents when the second starts).
Same kernel, same grid, same fixed-order grid sum -> the results are bit-identical; the assertion uses the
tolerance of the other head tests (rel 1e-5) for the floating-point outputs and exact equality for the counters.

Named test_zz_* so that pytest collects it last."""
import ctypes

import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _case(seed, r, k, dev):
    g = synth.generator(seed)
    logits = torch.randn(r, k + 1, generator=g) * 2
    gt = torch.randint(-1, k + 1, (r,), generator=g, dtype=torch.int64)  # -1 = ignore, k = background
    w = torch.rand(r, generator=g)
    w[gt == -1] = 0.0
    w[::7] = 0.0
    valid = torch.tensor([int((w > 1e-12).sum())], dtype=torch.int32)
    return logits.to(dev), gt.to(dev), w.to(dev), valid.to(dev)


def _run(fn_static, L, logits, gt, w, valid, loss_scale):
    r, kc = logits.shape
    dev = logits.device
    probs, grad = torch.empty_like(logits), torch.empty_like(logits)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    acc = torch.empty(4, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    if fn_static:
        rc = L.jtsm_refine_ce_static(logits.data_ptr(), r, kc, gt.data_ptr(), w.data_ptr(), valid.data_ptr(), 0,
                                     float(loss_scale), probs.data_ptr(), grad.data_ptr(), loss.data_ptr(),
                                     acc.data_ptr(), st)
    else:
        wsb = L.jtsm_refine_workspace_bytes()
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        rc = L.jtsm_refine_ce(logits.data_ptr(), r, kc, gt.data_ptr(), w.data_ptr(), valid.data_ptr(), 0,
                              float(loss_scale), probs.data_ptr(), grad.data_ptr(), loss.data_ptr(), acc.data_ptr(),
                              ws.data_ptr(), wsb, st)
    assert rc == 0, L.jtsm_error_string(rc)
    torch.cuda.synchronize()
    return probs.cpu().numpy(), grad.cpu().numpy(), loss.cpu().numpy(), acc.cpu().numpy()


def test_refine_ce_static_matches_refine_ce_on_two_calls_in_a_row(cuda):
    from jtsm_b200 import _lib

    L = _lib.lib()
    # two different problems back to back: the second starts on the scratch contents the first left behind
    for seed, r, k, scale in ((901, 2000, 20, 1.0), (902, 8000, 80, 3.0)):
        inp = _case(seed, r, k, cuda)
        want = _run(False, L, *inp, scale)
        got = _run(True, L, *inp, scale)
        for a, b, name in zip(got[:3], want[:3], ("probs", "grad_logits", "loss")):
            np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-7, err_msg=name)
        assert np.array_equal(got[3], want[3]), "accuracy counters"
        assert np.isfinite(got[2]).all() and got[2][0] > 0


def test_refine_ce_static_rejects_sizes_beyond_its_maximum(cuda):
    from jtsm_b200 import _lib

    L = _lib.lib()
    limit = L.jtsm_refine_ce_static_max_rois()
    assert limit == 262144
    one = torch.zeros(8, device=cuda)
    rc = L.jtsm_refine_ce_static(one.data_ptr(), limit + 1, 21, one.data_ptr(), one.data_ptr(), one.data_ptr(), 0, 1.0,
                                 0, 0, one.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 4  # JTSM_ERR_STATIC_LIMIT, before anything is launched
    assert b"static scratch limit" in L.jtsm_error_string(rc)
    assert isinstance(ctypes.c_int(rc).value, int)
