"""File-level loader for the reference's own Python (TEST INFRASTRUCTURE; this container only).

`import detectron2` / `import wsl` fail here (no fvcore / yacs / compiled _C), so the
reference sources are executed IN PLACE from /root/reference with a handful of stub modules for
what cannot be imported.  Nothing is copied into the repo.  Used by
  - tests/test_ref_parity.py  (skipped when /root/reference is absent, e.g. on the GPU box)
  - tests/golden/make_golden.py (generates the committed fixtures)
to check oracle/ref_heads.py (the restatement that travels) against the real reference code.

Stubs (and only these): fvcore.nn.smooth_l1_loss / giou_loss (fvcore is an absent third-party
dependency, fvcore>=0.1.2 in the reference's setup.py:214; restated from its published
definition -> "parity unpinned" for that function), detectron2.config.configurable (identity),
detectron2.utils.events.get_event_storage (recorder), detectron2.utils.env.TORCH_VERSION,
wsl.layers / PCL placeholders that the loaded code paths never call.
"""
import ast
import importlib.util
import os
import sys
import types

import torch

REF = os.environ.get("JTSM_REFERENCE", "/root/reference")
D2 = os.path.join(REF, "detectron2")
WSL = os.path.join(REF, "projects", "WSL", "wsl")


def available():
    return os.path.isdir(D2) and os.path.isdir(WSL)


class _Storage:
    def __init__(self):
        self.scalars = {}
        self.iter = 0

    def put_scalar(self, k, v, **kw):
        self.scalars[k] = float(v)

    def put_image(self, *a, **k):
        pass


STORAGE = _Storage()
_loaded = {}


def _smooth_l1_loss(input, target, beta, reduction="none"):
    # fvcore.nn.smooth_l1_loss (published definition): beta < 1e-5 -> L1
    if beta < 1e-5:
        loss = torch.abs(input - target)
    else:
        n = torch.abs(input - target)
        cond = n < beta
        loss = torch.where(cond, 0.5 * n ** 2 / beta, n - 0.5 * beta)
    if reduction == "mean":
        loss = loss.mean() if loss.numel() > 0 else 0.0 * loss.sum()
    elif reduction == "sum":
        loss = loss.sum()
    return loss


def _load_file(modname, path):
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []  # behave as a package so dotted stubs resolve, but block real imports
    sys.modules[name] = m
    return m


def install():
    """Populate sys.modules with the stubbed detectron2 / wsl / fvcore namespace (idempotent)."""
    if _loaded:
        return _loaded
    if not available():
        raise RuntimeError("reference checkout not found at %s" % REF)
    _stub("fvcore")
    _stub("fvcore.nn", smooth_l1_loss=_smooth_l1_loss, giou_loss=None)
    _stub("detectron2")
    _stub("detectron2.utils")
    _stub("detectron2.utils.env", TORCH_VERSION=tuple(int(x) for x in torch.__version__.split(".")[:2]))
    _stub("detectron2.utils.events", get_event_storage=lambda: STORAGE)
    _stub("detectron2.config", configurable=lambda f=None, **kw: f if f is not None else (lambda g: g))
    wrappers = _load_file("detectron2.layers.wrappers", os.path.join(D2, "layers", "wrappers.py"))
    nms_mod = _load_file("detectron2.layers.nms", os.path.join(D2, "layers", "nms.py"))
    shape_spec = _load_file("detectron2.layers.shape_spec", os.path.join(D2, "layers", "shape_spec.py"))
    _stub("detectron2.layers", cat=wrappers.cat, nonzero_tuple=wrappers.nonzero_tuple, Linear=torch.nn.Linear,
          batched_nms=nms_mod.batched_nms, ShapeSpec=shape_spec.ShapeSpec)
    boxes = _load_file("detectron2.structures.boxes", os.path.join(D2, "structures", "boxes.py"))
    inst = _load_file("detectron2.structures.instances", os.path.join(D2, "structures", "instances.py"))
    _stub("detectron2.structures", Boxes=boxes.Boxes, Instances=inst.Instances, pairwise_iou=boxes.pairwise_iou,
          BoxMode=boxes.BoxMode)
    _stub("detectron2.modeling")
    breg = _load_file("detectron2.modeling.box_regression", os.path.join(D2, "modeling", "box_regression.py"))
    matcher = _load_file("detectron2.modeling.matcher", os.path.join(D2, "modeling", "matcher.py"))
    _stub("wsl")
    _stub("wsl.layers", pcl_loss=None, ROIMerge=None)
    _stub("wsl.modeling")
    _stub("wsl.modeling.roi_heads")
    _stub("wsl.modeling.roi_heads.third_party")
    _stub("wsl.modeling.roi_heads.third_party.pcl", PCL=None)
    rh = os.path.join(WSL, "modeling", "roi_heads")
    wsddn = _load_file("wsl.modeling.roi_heads.fast_rcnn_wsddn", os.path.join(rh, "fast_rcnn_wsddn.py"))
    oicr = _load_file("wsl.modeling.roi_heads.fast_rcnn_oicr", os.path.join(rh, "fast_rcnn_oicr.py"))
    _loaded.update(boxes=boxes, instances=inst, box_regression=breg, matcher=matcher, nms=nms_mod, wsddn=wsddn,
                   oicr=oicr, wrappers=wrappers)
    return _loaded


def extract_methods(path, class_name, names, namespace):
    """Compile selected methods of `class_name` from the reference file `path` (in place) into
    plain functions -- for classes whose module cannot be imported as a whole."""
    with open(path) as fh:
        tree = ast.parse(fh.read(), filename=path)
    out = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == class_name:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name in names:
                    item.decorator_list = [d for d in item.decorator_list
                                           if not (isinstance(d, ast.Name) and d.id == "configurable")]
                    mod = ast.Module(body=[item], type_ignores=[])
                    code = compile(mod, path, "exec")
                    ns = dict(namespace)
                    exec(code, ns)
                    out[item.name] = ns[item.name]
    missing = set(names) - set(out)
    if missing:
        raise KeyError("not found in %s:%s: %s" % (path, class_name, sorted(missing)))
    return out


def roi_heads_methods():
    """get_pgt_top_k / get_pgt_mist (OICRROIHeads) and label_and_sample_proposals / _sample_proposals
    (ROIHeads) as callables taking an explicit `self`."""
    m = install()
    import numpy as np
    from typing import Dict, List, Optional, Tuple

    ns = dict(torch=torch, np=np, List=List, Tuple=Tuple, Dict=Dict, Optional=Optional, Boxes=m["boxes"].Boxes,
              Instances=m["instances"].Instances, pairwise_iou=m["boxes"].pairwise_iou,
              get_event_storage=lambda: STORAGE, batched_nms=m["nms"].batched_nms,
              add_ground_truth_to_proposals=None)
    rh = os.path.join(WSL, "modeling", "roi_heads")
    f = {}
    f.update(extract_methods(os.path.join(rh, "roi_heads_oicr.py"), "OICRROIHeads", ["get_pgt_top_k", "get_pgt_mist"], ns))
    f.update(extract_methods(os.path.join(rh, "roi_heads.py"), "ROIHeads",
                             ["label_and_sample_proposals", "_sample_proposals"], ns))
    return f


def make_self(num_classes, gt_classes_img_int, pred_class_img_logits, iou_thresholds=(0.5,), iou_labels=(0, 1),
              cls_agnostic_bbox_reg=False):
    """The attributes the extracted ROIHeads methods read."""
    m = install()
    f = roi_heads_methods()
    s = types.SimpleNamespace()
    s.num_classes = num_classes
    s.gt_classes_img_int = gt_classes_img_int
    s.pred_class_img_logits = pred_class_img_logits
    s.cls_agnostic_bbox_reg = cls_agnostic_bbox_reg
    s.proposal_append_gt = False
    s.proposal_matcher = m["matcher"].Matcher(list(iou_thresholds), list(iou_labels), allow_low_quality_matches=False)
    s._vis_pgt = lambda *a, **k: None
    s.vis_period = 0
    s.iter = 0
    s._sample_proposals = lambda *a: f["_sample_proposals"](s, *a)
    s.get_pgt_top_k = lambda *a, **k: f["get_pgt_top_k"](s, *a, **k)
    return s, f


def extract_function(path, name, namespace):
    """Compile one module-level function of the reference file `path` (in place) into a callable."""
    with open(path) as fh:
        tree = ast.parse(fh.read(), filename=path)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            code = compile(ast.Module(body=[node], type_ignores=[]), path, "exec")
            ns = dict(namespace)
            exec(code, ns)
            return ns[name]
    raise KeyError("%s not found in %s" % (name, path))


_tta = {}


def tta_reference():
    """The reference's own TTA-AVG code, executed in place on top of the restated fvcore base classes
    (oracle/fvcore_transforms.py):
      ResizeTransform              detectron2/data/transforms/transform.py:94-159 (whole module loaded)
      ResizeShortestEdge.get_transform   detectron2/data/transforms/augmentation_impl.py:152-172
      GeneralizedRCNNWithTTAAVG._get_augmented_boxes   projects/WSL/wsl/modeling/test_time_augmentation_avg.py:364-390
      transform_proposals          test_time_augmentation_avg.py:27-70
    -> dict(ResizeTransform, get_transform, get_augmented_boxes, transform_proposals, Boxes, Instances)."""
    if _tta:
        return _tta
    m = install()
    import numpy as np

    from . import fvcore_transforms as fv

    _stub("fvcore.transforms", HFlipTransform=fv.HFlipTransform, NoOpTransform=fv.NoOpTransform)
    _stub("fvcore.transforms.transform", CropTransform=fv.CropTransform, HFlipTransform=fv.HFlipTransform,
          NoOpTransform=fv.NoOpTransform, Transform=fv.Transform, TransformList=fv.TransformList)
    from PIL import Image

    if not hasattr(Image, "LINEAR"):  # removed from Pillow >= 10; the reference names it in a default argument only
        Image.LINEAR = Image.BILINEAR
    tr = _load_file("detectron2.data.transforms.transform", os.path.join(D2, "data", "transforms", "transform.py"))
    ns = dict(np=np, torch=torch, ResizeTransform=tr.ResizeTransform, NoOpTransform=fv.NoOpTransform,
              Boxes=m["boxes"].Boxes, Instances=m["instances"].Instances)
    gt = extract_methods(os.path.join(D2, "data", "transforms", "augmentation_impl.py"), "ResizeShortestEdge",
                         ["get_transform"], ns)["get_transform"]
    tta_path = os.path.join(WSL, "modeling", "test_time_augmentation_avg.py")
    gab = extract_methods(tta_path, "GeneralizedRCNNWithTTAAVG", ["_get_augmented_boxes"], ns)["_get_augmented_boxes"]
    tp = extract_function(tta_path, "transform_proposals", ns)
    _tta.update(ResizeTransform=tr.ResizeTransform, get_transform=gt, get_augmented_boxes=gab,
                transform_proposals=tp, Boxes=m["boxes"].Boxes, Instances=m["instances"].Instances)
    return _tta


def pairwise_iou_sp_reference():
    """The reference's own pairwise_iou_sp (roi_heads_jtsm.py:49-73), compiled in place; takes objects with
    an `.oh_labels` attribute like the reference's Instances."""
    install()
    path = os.path.join(WSL, "modeling", "roi_heads", "roi_heads_jtsm.py")
    return extract_function(path, "pairwise_iou_sp", dict(torch=torch))
