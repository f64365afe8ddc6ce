"""ctypes/numpy front end of oracle/jtsm_oracle.c (the CPU restatement).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never imported by jtsm_b200/.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "jtsm_oracle.c")
BUILD_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(BUILD_DIR, "libjtsm_oracle.so")

_lib = None


def build(force=False):
    os.makedirs(BUILD_DIR, exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-o", LIB, SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_nms.restype = ctypes.c_int
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def roi_pool_forward(inp, rois, spatial_scale, pooled, roi_range=None):
    inp, rois = _f32(inp), _f32(rois)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.zeros((r, c, ph, pw), np.float32)
    arg = np.full((r, c, ph, pw), -1, np.int32)
    b, e = roi_range or (0, r)
    lib().oracle_roi_pool_forward(_p(inp), _p(rois), c, h, w, ph, pw, ctypes.c_float(spatial_scale),
                                  b, e, _p(out), _p(arg))
    return out, arg


def max_pool_backward(grad_out, argmax, rois, input_shape):
    grad_out = _f32(grad_out)
    argmax = np.ascontiguousarray(argmax, dtype=np.int32)
    rois = _f32(rois)
    n, c, h, w = input_shape
    rows, _, ph, pw = grad_out.shape
    gi = np.zeros(input_shape, np.float32)
    lib().oracle_max_pool_backward(_p(grad_out), _p(argmax), _p(rois), rois.shape[0], c, h, w, ph, pw,
                                   0, rows, _p(gi))
    return gi


def roi_align_forward(inp, rois, spatial_scale, pooled, sampling_ratio, aligned, roi_range=None):
    inp, rois = _f32(inp), _f32(rois)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.zeros((r, c, ph, pw), np.float32)
    b, e = roi_range or (0, r)
    lib().oracle_roi_align_forward(_p(inp), _p(rois), c, h, w, ph, pw, ctypes.c_float(spatial_scale),
                                   int(sampling_ratio), int(bool(aligned)), b, e, _p(out))
    return out


def roi_align_backward(grad_out, rois, spatial_scale, input_shape, sampling_ratio, aligned):
    grad_out, rois = _f32(grad_out), _f32(rois)
    n, c, h, w = input_shape
    r, _, ph, pw = grad_out.shape
    gi = np.zeros(input_shape, np.float32)
    lib().oracle_roi_align_backward(_p(grad_out), _p(rois), c, h, w, ph, pw,
                                    ctypes.c_float(spatial_scale), int(sampling_ratio),
                                    int(bool(aligned)), 0, r, _p(gi))
    return gi


def roi_loop_pool_forward(inp, rois, spatial_scale, pooled, context_ratio=1.8):
    inp, rois = _f32(inp), _f32(rois)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.zeros((3 * r, c, ph, pw), np.float32)
    arg = np.full((3 * r, c, ph, pw), -1, np.int32)
    lib().oracle_roi_loop_pool_forward(_p(inp), _p(rois), r, c, h, w, ph, pw,
                                       ctypes.c_float(spatial_scale), ctypes.c_float(context_ratio),
                                       0, r, _p(out), _p(arg))
    return out, arg


def moi_pool_forward(inp, rois, oh_labels, superpixels, spatial_scale, pooled):
    inp, rois = _f32(inp), _f32(rois)
    oh = np.ascontiguousarray(oh_labels, dtype=np.int32)
    sp = np.ascontiguousarray(superpixels, dtype=np.int32)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.zeros((r, c, ph, pw), np.float32)
    arg = np.full((r, c, ph, pw), -1, np.int32)
    lib().oracle_moi_pool_forward(_p(inp), _p(rois), _p(oh), oh.shape[1], _p(sp), sp.shape[1],
                                  sp.shape[2], c, h, w, ph, pw, ctypes.c_float(spatial_scale), 0, r,
                                  _p(out), _p(arg))
    return out, arg


def pairwise_iou(boxes1, boxes2):
    b1, b2 = _f32(boxes1).reshape(-1, 4), _f32(boxes2).reshape(-1, 4)
    out = np.zeros((b1.shape[0], b2.shape[0]), np.float32)
    lib().oracle_pairwise_iou(_p(b1), b1.shape[0], _p(b2), b2.shape[0], _p(out))
    return out


def match(gt_boxes, prop_boxes, thr=0.5):
    gt, pr = _f32(gt_boxes).reshape(-1, 4), _f32(prop_boxes).reshape(-1, 4)
    r = pr.shape[0]
    idx = np.zeros(r, np.int64)
    val = np.zeros(r, np.float32)
    lab = np.zeros(r, np.int8)
    lib().oracle_match(_p(gt), gt.shape[0], _p(pr), r, ctypes.c_float(thr), _p(idx), _p(val), _p(lab))
    return idx, val, lab


def nms(boxes, scores, thr):
    b, s = _f32(boxes).reshape(-1, 4), _f32(scores).reshape(-1)
    keep = np.zeros(b.shape[0], np.int64)
    nk = lib().oracle_nms(_p(b), _p(s), b.shape[0], ctypes.c_float(thr), _p(keep))
    return keep[:nk].copy()


def pcl_loss_forward(probs, labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels):
    """-> (output [C], loss = output.sum() / R)  (pcl_loss_cpu.cpp:8-57, pcl_loss.py:52)."""
    a = [_f32(x).reshape(-1) for x in (labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels)]
    probs = _f32(probs)
    r, c = probs.shape
    out = np.zeros(c, np.float32)
    lib().oracle_pcl_loss_forward(_p(probs), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(a[5]), r, c,
                                  a[2].shape[0], _p(out))
    return out, np.float32(out.sum() / np.float32(r))


def pcl_loss_backward(probs, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                      img_cls_loss_weights, im_labels):
    """-> grad [R, C] / R  (pcl_loss_cpu.cpp:59-115, pcl_loss.py:84)."""
    a = [_f32(x).reshape(-1) for x in (labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                                       img_cls_loss_weights, im_labels)]
    probs = _f32(probs)
    r, c = probs.shape
    g = np.zeros((r, c), np.float32)
    lib().oracle_pcl_loss_backward(_p(probs), *[_p(x) for x in a], r, c, _p(g))
    return g / np.float32(r)
