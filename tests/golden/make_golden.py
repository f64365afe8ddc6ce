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
    make_tta()


TTA_CASE = dict(seed=301, image_hw=(375, 500), orig_hw=(375, 500), min_sizes=(480, 688, 1200), max_size=4000, flip=True,
                r=120, k=6)


def make_tta():
    """tests/golden/tta_case0.npz: the reference's own GeneralizedRCNNWithTTAAVG._get_augmented_boxes (run in place;
    ResizeShortestEdge.get_transform and ResizeTransform are the reference's too) on synthetic augmented passes."""
    import types

    from oracle import ref_loader, ref_tta
    from oracle.fvcore_transforms import HFlipTransform, TransformList

    c = TTA_CASE
    ref = ref_loader.tta_reference()
    g = synth.generator(c["seed"])
    img = np.zeros(c["image_hw"] + (3,), np.uint8)
    tf_ref = []
    for size in c["min_sizes"]:
        aug = types.SimpleNamespace(is_range=False, short_edge_length=(size, size), max_size=c["max_size"], interp=None)
        resize = ref["get_transform"](aug, img)
        tf_ref.append(TransformList([resize]))
        tf_ref.append(TransformList([resize, HFlipTransform(resize.new_w)]))
    pb = synth.proposals(g, c["r"], *c["orig_hw"])
    boxes, scores = ref_tta.synth_augmented_outputs(g, pb, tf_ref, c["k"])
    self_ = types.SimpleNamespace(_batch_inference=lambda inputs: (None, [s.clone() for s in scores],
                                                                   [b.clone() for b in boxes]))
    mb, ms, _ = ref["get_augmented_boxes"](self_, [None] * len(tf_ref), tf_ref)
    np.savez_compressed(os.path.join(HERE, "tta_case0.npz"), image_hw=np.array(c["image_hw"]),
                        orig_hw=np.array(c["orig_hw"]), min_sizes=np.array(c["min_sizes"]), max_size=c["max_size"],
                        flip=c["flip"], boxes=torch.cat(boxes).numpy(), scores=torch.cat(scores).numpy(),
                        mean_boxes=mb.numpy(), mean_scores=ms.numpy())
    print("wrote tta_case0.npz")


if __name__ == "__main__":
    main()
