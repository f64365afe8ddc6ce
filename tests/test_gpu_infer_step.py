"""BASELINE.json configs[3] (c4) at full size: 10 augmentations (5 scales x flip) x 4000 proposals x 20 classes through
jtsm_b200.inferpath.InferenceStep (pool -> mean of Kr softmaxes -> decoded mean deltas -> TTA merge -> threshold ->
per-class NMS -> top-100), eager and as a replayed CUDA graph, against the CPU composition of the reference's pieces
(oracle/cpu_infer.py: restatements pinned to the reference's own functions in tests/test_tta.py):
  * merged boxes / scores: rel 1e-5 (device exp / sum order in the softmax);
  * detections: EXACT, given the device's own merged boxes / scores (80 000 candidates through the CPU function);
  * pooled features of the largest augmentation (150 x 200 map): bit-exact vs the C oracle on a channel slice."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth
from oracle import cpu_infer, ref_heads

pytestmark = pytest.mark.gpu


def test_c4_inference_step_matches_cpu_composition(cuda, oracle):
    from jtsm_b200 import inferpath, ops

    wl = synth.infer_case("c4")
    cfg = wl["cfg"]
    step = inferpath.InferenceStep(cfg, cuda)
    d = step.upload(step.host_inputs(wl))
    torch.cuda.synchronize()
    det = step.run(d)
    want_det, want_mb, want_ms = cpu_infer.heads_and_detections(wl)
    mb, ms = step.last["mean_boxes"].cpu(), step.last["mean_scores"].cpu()
    np.testing.assert_allclose(ms.numpy(), want_ms.numpy(), rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(mb.numpy(), want_mb.numpy(), rtol=1e-5, atol=1e-3)
    # exact detections for the device's own merged tensors
    exact = ref_heads.fast_rcnn_inference_single_image(mb, ms, (cfg["img_h"], cfg["img_w"]), cfg["score_thresh"],
                                                       cfg["nms_thresh"], cfg["topk"])
    n = int(det["num"].item())
    assert n == exact["scores"].shape[0] == 100
    assert torch.equal(det["pred_boxes"][:n].cpu(), exact["pred_boxes"])
    assert torch.equal(det["scores"][:n].cpu(), exact["scores"])
    assert torch.equal(det["pred_classes"][:n].cpu(), exact["pred_classes"])
    assert torch.equal(det["pred_inds"][:n].cpu(), exact["pred_inds"])
    # and they are the CPU chain's detections up to the softmax's last bits: same classes / proposals for the top ones
    same = (det["pred_inds"][:n].cpu() == want_det["pred_inds"]) & (det["pred_classes"][:n].cpu() == want_det["pred_classes"])
    assert float(same.float().mean()) > 0.9
    # pooled features of the largest augmentation: 16 channels x 300 proposals vs the C oracle
    a = wl["augs"][-1]
    xs = a["feat"][:, 40:56].contiguous()
    ref, _ = oracle.roi_pool_forward(xs.numpy(), a["rois"][:300].numpy(), cfg["scale"], (7, 7))
    ref = ref * (wl["objectness"][:300] + 1.0).numpy()[:, None, None, None]
    assert np.array_equal(step.last["pooled"][-1][:300, 40:56].cpu().numpy(), ref)
    # the captured graph gives the same detections, and serves a second image through the same buffers
    step.capture(d)
    step.graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(step.last["det"]["pred_boxes"][:n].cpu(), exact["pred_boxes"])
    wl2 = synth.infer_case("c4", sweep_index=1)
    step.upload(step.host_inputs(wl2), into=d)
    step.graph.replay()
    torch.cuda.synchronize()
    mb2, ms2 = step.last["mean_boxes"].cpu(), step.last["mean_scores"].cpu()
    exact2 = ref_heads.fast_rcnn_inference_single_image(mb2, ms2, (cfg["img_h"], cfg["img_w"]), cfg["score_thresh"],
                                                        cfg["nms_thresh"], cfg["topk"])
    n2 = int(step.last["det"]["num"].item())
    assert n2 == exact2["scores"].shape[0]
    assert torch.equal(step.last["det"]["pred_boxes"][:n2].cpu(), exact2["pred_boxes"])
    assert torch.equal(step.last["det"]["pred_classes"][:n2].cpu(), exact2["pred_classes"])
    assert not torch.equal(exact2["pred_boxes"], exact["pred_boxes"])
