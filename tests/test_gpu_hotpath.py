"""The whole hot-path step (jtsm_b200/hotpath.py:HotPathStep, what bench.py times) at reduced sizes against the CPU
oracle composition, for both poolers the step supports: RoIPool (c2, WSDDN+OICR) and MOIPool with the JTSM epilogue
(c3m).  Pooled features bit-exact, pseudo-labels bit-exact, losses rel 1e-5, gradient to the feature map within the
atomic-sum tolerance; the CUDA-graph replay reproduces the eager step."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import ref_heads

pytestmark = pytest.mark.gpu


def _oracle_pool(oracle, wl):
    cfg = wl["cfg"]
    x, rois = wl["feat"].numpy(), wl["rois"].numpy()
    P = (cfg["P"], cfg["P"])
    rs = (wl["objectness"] + 1.0).numpy()
    if cfg["pooler"] == "MOIPool":
        out, arg = oracle.moi_pool_forward(x, rois, wl["oh_labels"].numpy(), wl["superpixels"].numpy(), cfg["scale"], P)
        valid = (arg[:, 0].reshape(arg.shape[0], -1) != -1).sum(1).astype(np.float32)
        ms = np.float32(P[0] * P[1]) * (np.float32(1.0) / (valid + np.float32(1.0)))  # roi_heads_jtsm.py:615-633
        pooled = (out * ms[:, None, None, None]) * rs[:, None, None, None]
        bwd = ms * rs
    else:
        out, arg = oracle.roi_pool_forward(x, rois, cfg["scale"], P)
        pooled = out * rs[:, None, None, None]
        bwd = rs
    return pooled, arg, bwd


@pytest.mark.parametrize("name,over", [("c2", dict(C=8, R=300)), ("c3m", dict(C=6, R=150, K=12, Kr=2)),
                                       ("c3m", dict(C=5, R=90, K=12, Kr=1, N=1))])
def test_hot_path_step_matches_oracle(cuda, oracle, name, over):
    from jtsm_b200 import hotpath

    wl = synth.step_case(name, **over)
    cfg = wl["cfg"]
    step = hotpath.HotPathStep(cfg, cuda)
    host = step.pin_host_inputs(wl)
    step.run_from_host(host)
    torch.cuda.synchronize()
    last = step.last
    pooled, arg, bwd = _oracle_pool(oracle, wl)
    assert np.array_equal(last["pooled"].cpu().numpy(), pooled)
    # heads
    r = ref_heads.head_step(wl["mil_cls"], wl["mil_det"], wl["ref_logits"], wl["ref_deltas"], wl["boxes"], wl["labels"],
                            cfg["K"], reg=cfg["reg"])
    h = last["heads"]
    for k, v in r["losses"].items():
        np.testing.assert_allclose(h["losses"][k].cpu().numpy().reshape(()), v.detach().numpy(), rtol=1e-5, atol=1e-7)
    for rb, hb in zip(r["branches"], h["branches"]):
        assert torch.equal(rb["gt_classes"], hb["gt_classes"].cpu())
    # backward to the feature map
    g = synth.expand_grad_tile(wl["grad_tile"], wl["rois"].shape[0]).numpy() * bwd[:, None, None, None]
    gi = oracle.max_pool_backward(g, arg, wl["rois"].numpy(), tuple(wl["feat"].shape))
    np.testing.assert_allclose(last["grad_feat"].cpu().numpy(), gi, rtol=1e-5, atol=1e-5 * float(np.abs(gi).max()))
    # CUDA graph of the step reproduces it
    d = step.upload(host)
    torch.cuda.synchronize()
    step.capture(d)
    step.replay()
    torch.cuda.synchronize()
    assert torch.equal(step.graph_last["pooled"], last["pooled"])
    np.testing.assert_allclose(step.graph_last["grad_feat"].cpu().numpy(), last["grad_feat"].cpu().numpy(), rtol=1e-5,
                               atol=1e-5 * float(np.abs(gi).max()))
    for k in h["losses"]:
        assert torch.equal(step.graph_last["heads"]["losses"][k], h["losses"][k])


@pytest.mark.parametrize("depth,overlap", [(2, "any"), (3, "any"), (3, "after_forward")])
def test_streamed_steps_deliver_every_steps_results_in_order(cuda, depth, overlap):
    """HotPathStep.run_stream (what bench.py's e2e times): distinct host batches streamed through the captured
    graphs; every step's losses / pseudo-labels arrive, in order, and equal the un-pipelined path's -- including the image-level
    labels, which travel with the batch (device-side class lists), not with the captured graph."""
    from jtsm_b200 import hotpath

    wls = [synth.step_case("c2", sweep_index=i, C=4, R=200) for i in range(7)]
    for w in wls[1:]:  # the upstream gradient (fc6's backward) is device-resident state of the step, not a host input
        w["grad_tile"] = wls[0]["grad_tile"]
    step = hotpath.HotPathStep(wls[0]["cfg"], cuda)
    hosts = [step.pin_host_inputs(w) for w in wls]
    want0 = [[t.clone() for t in step.run_from_host(h)] for h in hosts]
    assert any(not torch.equal(want0[0][2], w[2]) for w in want0[1:])  # the batches really differ (labels too)
    step.prepare_stream(hosts[0], depth=depth, overlap=overlap)
    got = {}
    n = step.run_stream(hosts, consume=lambda i, out: got.__setitem__(i, [t.clone() for t in out]))
    assert n == 7 and sorted(got) == list(range(7))
    for i in range(7):
        for a, b in zip(got[i], want0[i]):
            if a.dtype.is_floating_point:
                np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-5 * max(1.0, float(b.abs().max())))
            else:
                assert torch.equal(a, b)
