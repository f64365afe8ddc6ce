"""Generate tests/golden/heads_*.npz by running the REFERENCE's own code (in place from /root/reference,
via oracle/ref_loader.py + oracle/ref_drive.py) on small seeded inputs.  Run in the build container:
    python tests/golden/make_golden.py
Reference commit: a691dc59 (.SUBMODULES.json:3).  Inputs are regenerated from the seed by the tests
(tests/test_golden.py: same `case()` below), so only the reference OUTPUTS are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from jtsm_b200 import synth  # noqa: E402

CASES = {
    # name: (seed, counts, K, Kr, reg, mean_loss, thresholds, labels)
    "a": (101, (60, 90), 6, 2, False, True, (0.5,), (0, 1)),
    "b": (102, (150,), 20, 3, False, True, (0.5,), (0, 1)),
    "c": (103, (80, 40, 70), 9, 2, True, False, (0.5,), (0, 1)),
    "d": (104, (200,), 5, 1, False, True, (0.3, 0.6), (0, -1, 1)),
}


def case(name):
    seed, counts, k, kr, reg, mean_loss, thr, lab = CASES[name]
    g = synth.generator(seed)
    boxes = [synth.proposals(g, c, 600, 800) for c in counts]
    r = sum(counts)
    labels = synth.image_labels(g, len(counts), k)
    inputs = dict(mil_cls=torch.randn(r, k, generator=g) * 3, mil_det=torch.randn(r, k, generator=g) * 3,
                  ref_logits=[torch.randn(r, k + 1, generator=g) * 2 for _ in range(kr)],
                  ref_deltas=[torch.randn(r, 4 * k, generator=g) * 0.1 for _ in range(kr)],
                  proposal_boxes=boxes, labels=labels, num_classes=k)
    return inputs, dict(reg=reg, mean_loss=mean_loss, thresholds=thr, iou_labels=lab)


def flatten(out):
    d = {"scores": out["scores"], "img_scores": out["img_scores"], "grad_cls": out["grad_cls"], "grad_det": out["grad_det"]}
    for k, v in out["losses"].items():
        d["loss/" + k] = v
    for i, b in enumerate(out["branches"]):
        for key in ("gt_classes", "gt_weights", "gt_boxes", "probs", "boxes"):
            d["b%d/%s" % (i, key)] = b[key]
        d["b%d/pgt_boxes" % i] = torch.cat([p[1] for p in b["pgt"]])
        d["b%d/pgt_classes" % i] = torch.cat([p[2] for p in b["pgt"]])
        d["b%d/pgt_scores" % i] = torch.cat([p[0] for p in b["pgt"]])
        d["b%d/grad_logits" % i] = out["grad_ref_logits"][i]
        d["b%d/grad_deltas" % i] = out["grad_ref_deltas"][i]
    return {k: v.detach().numpy() for k, v in d.items()}


def main():
    from oracle import ref_drive

    for name in CASES:
        inputs, kw = case(name)
        out = ref_drive.reference_head_step(**inputs, **kw)
        np.savez_compressed(os.path.join(HERE, "heads_%s.npz" % name), **flatten(out))
        print("wrote heads_%s.npz" % name)


if __name__ == "__main__":
    main()
