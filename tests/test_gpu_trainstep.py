"""The ROI heads as a whole (`OICRROIHeadsB200.forward`: pool -> fc6/fc7 -> MIL -> Kr refinement branches) trained with
`loss.backward()` + SGD on the device (jtsm_b200/trainstep.py, the step bench.py times for train_images_per_sec)."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


def _small(name="c2", **over):
    kw = dict(C=16, R=300)
    kw.update(over)
    return synth.step_case(name, **kw)


def test_roi_heads_fused_objectness_equals_reference_order_of_operations(cuda):
    """`box_features * (objectness_logits + 1)` (roi_heads_oicr.py:393-394) fused into the pooling kernel and its
    backward == the reference's explicit multiply on the pooled tensor: identical losses, d loss / d feature map to 1e-5."""
    from jtsm_b200 import trainstep

    wl = _small()
    res = []
    for fused in (True, False):
        ts = trainstep.TrainStep(wl["cfg"], cuda, fc_dims=(64, 48), tf32=False)
        ts.heads.fuse_objectness = fused
        ts.heads.eval()  # no dropout: the two runs must see the same activations ...
        ts.heads.training = True  # ... but take the training branch of forward()
        for m in ts.heads.modules():
            if isinstance(m, torch.nn.Dropout):
                m.eval()
        batch = ts.make_batch(wl)
        _, losses = ts.heads(None, {"conv5": batch["feat"]}, batch["proposals"], batch["targets"])
        sum(losses.values()).backward()
        res.append(({k: float(v) for k, v in losses.items()}, batch["feat"].grad.clone(),
                    ts.heads.box_head.fc1.weight.grad.clone()))
    assert set(res[0][0]) == {"loss_cls", "loss_cls_r0", "loss_cls_r1", "loss_cls_r2"}
    for k in res[0][0]:
        assert abs(res[0][0][k] - res[1][0][k]) <= 1e-5 * abs(res[1][0][k]), k
    for a, b in zip(res[0][1:], res[1][1:]):
        np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-5, atol=1e-5 * float(b.abs().max()))
    assert float(res[0][1].abs().max()) > 0


@pytest.mark.parametrize("name,over", [("c2", {}), ("c3", dict(C=12, R=200, K=80))])
def test_train_step_runs_and_learns(cuda, name, over):
    """A few SGD iterations on one fixed batch: finite losses, every trainable parameter receives a gradient, the MIL
    loss falls by more than 10 % (c3: two images per step, box regression on)."""
    from jtsm_b200 import trainstep

    wl = _small(name, **over)
    ts = trainstep.TrainStep(wl["cfg"], cuda, fc_dims=(96, 64), lr=1e-2)
    batch = ts.make_batch(wl)
    hist = []
    for i in range(12):
        loss = ts.step(batch)
        hist.append(float(ts.last_losses["loss_cls"]))
        assert np.isfinite(float(loss))
        if i == 0:
            missing = [n for n, p in ts.heads.named_parameters() if p.requires_grad and p.grad is None]
            assert not missing, missing
            assert batch["feat"].grad is not None and float(batch["feat"].grad.abs().max()) > 0
    assert min(hist[1:]) < hist[0] and (name != "c2" or min(hist) < 0.9 * hist[0]), hist
    expect = {"loss_cls"} | {"loss_cls_r%d" % k for k in range(wl["cfg"]["Kr"])}
    if wl["cfg"]["reg"]:
        expect |= {"loss_box_reg_r%d" % k for k in range(wl["cfg"]["Kr"])}
    assert set(ts.last_losses) == expect


def test_roi_heads_inference_returns_reference_tuple(cuda):
    from jtsm_b200 import trainstep

    wl = _small()
    ts = trainstep.TrainStep(wl["cfg"], cuda, fc_dims=(64, 48))
    ts.heads.eval()
    batch = ts.make_batch(wl)
    with torch.no_grad():
        pred, empty, all_scores, all_boxes = ts.heads(None, {"conv5": batch["feat"]}, batch["proposals"], None)
    assert empty == {} and len(pred) == 1 and pred[0]["pred_boxes"].shape[1] == 4
    assert all_scores[0].shape == (1, 300, 21) and all_boxes[0].shape == (1, 300, 80)


def test_train_step_captured_graph_equals_eager(cuda):
    """The whole step (forward, backward through the autograd Functions, SGD) replays from one CUDA graph: same loss
    trajectory as the eager step (dropout off so that both draw nothing from the RNG)."""
    from jtsm_b200 import synth
    from jtsm_b200.trainstep import TrainStep

    wl = synth.step_case("c2", C=16, R=300)
    hist = {}
    for mode in ("eager", "graph"):
        ts = TrainStep(wl["cfg"], cuda, world=1, fc_dims=(64, 64), lr=1e-2, fused_sgd=(mode == "graph"))
        for m in ts.heads.modules():
            if isinstance(m, torch.nn.Dropout):
                m.p = 0.0
        batch = ts.make_batch(wl)
        if mode == "eager":
            hist[mode] = [float(ts.step(batch)) for _ in range(8)]
        else:
            ts.capture(batch)  # three eager warm-up steps, then the capture (whose forward is not executed)
            hist[mode] = [float(ts.replay()) for _ in range(5)]
    # eager steps 4..8 line up with replays 1..5
    np.testing.assert_allclose(hist["graph"], hist["eager"][3:], rtol=2e-4)
