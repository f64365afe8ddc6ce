"""RoIPool -> fc6 fused on the tensor cores (jtsm_pool_fc6_forward: tcgen05.mma + TMEM accumulators + TMA weight tiles;
SURVEY 8f-2) against the unfused chain on the same bf16-rounded operands:
    relu(bias + flatten(roi_pool(x_bf16, rois) * row_scale -> bf16) @ W_bf16^T)       accumulated in fp32
Tolerance: rel 1e-2 of the largest output (north_star's bf16 bar; the operands are identical, only the fp32 summation
order differs, so the observed error is ~1e-6)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _reference(x_nchw_bf16, rois, scale, P, w, bias, row_scale, relu):
    from jtsm_b200 import ops

    pooled, _ = ops.roi_pool_forward(x_nchw_bf16, rois, scale, P, row_scale=row_scale)  # bf16 [R, C, P, P], bit-exact vs the oracle
    out = pooled.flatten(1).float() @ w.to(torch.bfloat16).float().t() + bias
    return torch.relu(out) if relu else out


@pytest.mark.parametrize("c,h,w,r,nout,p,relu", [
    (64, 38, 50, 200, 512, 7, True),      # two M tiles (128 + 72 rows), one N tile, 49 K slices
    (128, 30, 40, 90, 1024, 7, False),    # one partial M tile, two N tiles, 98 K slices, signed outputs
    (64, 20, 24, 300, 512, 3, True),      # 3 x 3 bins
])
def test_pool_fc6_fused_matches_unfused_chain(cuda, c, h, w, r, nout, p, relu):
    from jtsm_b200 import ops

    g = synth.generator(7000 + c + r)
    x = (synth.feature_map(g, 2, c, h, w) - 0.1).to(torch.bfloat16)
    boxes = [synth.proposals(g, r // 2, h * 16, w * 16), synth.proposals(g, r - r // 2, h * 16, w * 16)]
    boxes[1][0] = torch.tensor([-50.0, -50.0, -10.0, -10.0])  # outside the map: empty bins -> zeros
    rois = synth.rois_from_boxes(boxes)
    wt = torch.randn(nout, c * p * p, generator=g) * 0.05
    bias = torch.randn(nout, generator=g)
    rs = torch.rand(r, generator=g) + 1.0
    xd, rd, wd, bd, sd = x.to(cuda), rois.to(cuda), wt.to(cuda), bias.to(cuda), rs.to(cuda)
    want = _reference(xd, rd, 1 / 16.0, (p, p), wd, bd, sd, relu)
    x_nhwc = xd.permute(0, 2, 3, 1).contiguous()
    wb = ops.fc6_weight_binmajor(wd, c, p)
    got = ops.pool_fc6_forward(x_nhwc, rd, 1 / 16.0, (p, p), wb, bd, row_scale=sd, relu=relu)
    torch.cuda.synchronize()
    scale = float(want.abs().max())
    assert scale > 1.0
    err = float((got - want).abs().max())
    assert err <= 1e-2 * scale, (err, scale)
    assert err <= 1e-4 * scale, (err, scale)  # same operands, fp32 accumulation: far inside the bf16 bar
    # a second call (barriers / TMEM re-initialised per launch) gives the same bits
    again = ops.pool_fc6_forward(x_nhwc, rd, 1 / 16.0, (p, p), wb, bd, row_scale=sd, relu=relu)
    assert torch.equal(got, again)


def test_pool_fc6_rejects_unsupported_shapes(cuda):
    from jtsm_b200 import ops

    x = torch.zeros(1, 8, 8, 48, dtype=torch.bfloat16, device=cuda)  # channels not a multiple of 64
    rois = torch.tensor([[0, 0, 0, 50, 50.0]], device=cuda)
    w = torch.zeros(512, 48 * 49, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(RuntimeError):
        ops.pool_fc6_forward(x, rois, 1 / 16.0, 7, w, torch.zeros(512, device=cuda))
