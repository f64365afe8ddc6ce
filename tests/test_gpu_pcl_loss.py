"""GPU parity of the device PCL loss through the C ABI against the CPU restatement of pcl_loss_cpu.cpp (oracle C,
itself bit-equal to the reference's own file built in place: tests/test_ref_native.py): backward bit-exact, forward
sums within rel 1e-5 (fp32 sum order + device logf), module = the reference's `PCLLoss` call signature."""
import numpy as np
import pytest
import torch

from jtsm_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("r,k,p", [(300, 20, 6), (2000, 20, 12), (8000, 80, 30), (7, 3, 4), (50, 80, 1)])
def test_pcl_loss_forward_backward(cuda, oracle, r, k, p):
    from jtsm_b200 import ops

    c = synth.pcl_case(synth.generator(60 + r), r, k, p)
    n = {kk: v.numpy() for kk, v in c.items()}
    want_out, want_loss = oracle.pcl_loss_forward(n["probs"], n["labels"], n["cls_loss_weights"], n["pc_labels"],
                                                  n["pc_probs"], n["img_cls_loss_weights"], n["im_labels"])
    want_g = oracle.pcl_loss_backward(n["probs"], n["labels"], n["cls_loss_weights"], n["gt_assignment"],
                                      n["pc_labels"], n["pc_probs"], n["pc_count"], n["img_cls_loss_weights"],
                                      n["im_labels"])
    probs = c["probs"].to(cuda)
    loss, out = ops.pcl_loss_forward(probs, c["labels"], c["cls_loss_weights"], c["pc_labels"], c["pc_probs"],
                                     c["img_cls_loss_weights"], c["im_labels"])
    np.testing.assert_allclose(out.cpu().numpy(), want_out, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(loss.cpu().numpy().reshape(()), want_loss, rtol=1e-5)
    g = ops.pcl_loss_backward(probs, c["labels"], c["cls_loss_weights"], c["gt_assignment"], c["pc_labels"],
                              c["pc_probs"], c["pc_count"], c["img_cls_loss_weights"], c["im_labels"])
    assert np.array_equal(g.cpu().numpy(), want_g)  # elementwise, same operation order -> bit-exact


def test_pcl_loss_module_autograd(cuda, oracle):
    from jtsm_b200.layers import PCLLoss

    c = synth.pcl_case(synth.generator(9), 500, 20, 8)
    n = {kk: v.numpy() for kk, v in c.items()}
    probs = c["probs"].to(cuda).requires_grad_()
    loss = PCLLoss()(probs, c["labels"], c["cls_loss_weights"], c["gt_assignment"], c["pc_labels"], c["pc_probs"],
                     c["pc_count"], c["img_cls_loss_weights"], c["im_labels"])
    assert loss.dim() == 0
    (loss * 3.0).backward()  # like the reference, the upstream factor is not applied
    want_g = oracle.pcl_loss_backward(n["probs"], n["labels"], n["cls_loss_weights"], n["gt_assignment"],
                                      n["pc_labels"], n["pc_probs"], n["pc_count"], n["img_cls_loss_weights"],
                                      n["im_labels"])
    assert np.array_equal(probs.grad.cpu().numpy(), want_g)
