"""oracle/_ref: the reference's OWN native code for the path, built in place from /root/reference and run on the CPU.

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.build()): never imported by jtsm_b200/.

 * libref_wslpool.so -- the device kernels of projects/WSL/wsl/layers/csrc/{ROILoopPool,MOIPool}/*_cuda.cu compiled
   as serial host C++ (oracle/Makefile.ref, oracle/ref_shim/, oracle/ref_wslpool_driver.cpp).  The reference has no
   runnable CPU path for these two ops (ROILoopPool.h:62, MOIPool.h:74 raise), so this is the only way to execute
   the reference's arithmetic for them without a GPU; it pins oracle/jtsm_oracle.c's restatement bit for bit.

 * libref_pclloss.so -- projects/WSL/wsl/layers/csrc/pcl_loss/pcl_loss_cpu.cpp (the code the reference's PCL loss always
   runs, on the CPU) behind plain pointers (oracle/ref_pclloss_driver.cpp).
 * libref_roialign.so -- detectron2/layers/csrc/ROIAlign/ROIAlign_cpu.cpp compiled where it lies against this image's
   ATen (oracle/ref_roialign_driver.cpp): the reference's CPU RoIAlign forward / backward behind plain pointers.

The .so files are built when /root/reference is present (this container) and travels to the GPU box prebuilt.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("JTSM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
WSLPOOL = os.path.join(OUT, "libref_wslpool.so")
ROIALIGN = os.path.join(OUT, "libref_roialign.so")
PCLLOSS = os.path.join(OUT, "libref_pclloss.so")

_wsl = None
_align = None


def build():
    """Build oracle/_ref from the reference sources where they lie; returns the .so path or None when absent."""
    if os.path.isdir(os.path.join(REF, "projects", "WSL")):
        subprocess.run(["make", "-s", "-f", "Makefile.ref", f"REF={REF}"], cwd=HERE, check=True)
    return WSLPOOL if os.path.exists(WSLPOOL) else None


def _build_aten(target, marker, path):
    if os.path.isfile(os.path.join(REF, *marker)):
        import torch
        from torch.utils import cpp_extension as ce

        inc = " ".join(ce.include_paths())
        libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
        subprocess.run(["make", "-s", "-f", "Makefile.ref", f"REF={REF}", f"TORCH_INC={inc}", f"TORCH_LIB={libdir}",
                        "_ref/" + target], cwd=HERE, check=True)
    return path if os.path.exists(path) else None


def build_roi_align():
    return _build_aten("libref_roialign.so", ("detectron2", "layers", "csrc", "ROIAlign", "ROIAlign_cpu.cpp"), ROIALIGN)


def build_pcl_loss():
    return _build_aten("libref_pclloss.so", ("projects", "WSL", "wsl", "layers", "csrc", "pcl_loss", "pcl_loss_cpu.cpp"),
                       PCLLOSS)


_pcl = None


def pcl_loss_available():
    return build_pcl_loss() is not None


def pcl():
    global _pcl
    if _pcl is None:
        path = build_pcl_loss()
        if path is None:
            raise RuntimeError("oracle/_ref/libref_pclloss.so missing and /root/reference not present")
        import torch  # noqa: F401

        _pcl = ctypes.CDLL(path)
    return _pcl


def available():
    return build() is not None


def roi_align_available():
    return build_roi_align() is not None


def align():
    global _align
    if _align is None:
        path = build_roi_align()
        if path is None:
            raise RuntimeError("oracle/_ref/libref_roialign.so missing and /root/reference not present")
        import torch  # noqa: F401  (libtorch_cpu / libc10 must be loaded first)

        _align = ctypes.CDLL(path)
    return _align


def wsl():
    global _wsl
    if _wsl is None:
        path = build()
        if path is None:
            raise RuntimeError("oracle/_ref/libref_wslpool.so missing and /root/reference not present")
        _wsl = ctypes.CDLL(path)
    return _wsl


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def roi_loop_pool_forward(inp, rois, spatial_scale, pooled):
    inp, rois = _f32(inp), _f32(rois)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.empty((3 * r, c, ph, pw), np.float32)
    arg = np.empty((3 * r, c, ph, pw), np.int32)
    wsl().ref_roi_loop_pool_forward(_p(inp), _p(rois), r, c, h, w, ph, pw, ctypes.c_float(spatial_scale),
                                    _p(out), _p(arg))
    return out, arg


def roi_loop_pool_backward(grad_out, argmax, rois, input_shape, spatial_scale=1.0):
    grad_out, argmax, rois = _f32(grad_out), _i32(argmax), _f32(rois)
    n, c, h, w = input_shape
    ph, pw = grad_out.shape[2:]
    gi = np.empty(input_shape, np.float32)
    wsl().ref_roi_loop_pool_backward(_p(grad_out), _p(argmax), _p(rois), rois.shape[0], n, c, h, w, ph, pw,
                                     ctypes.c_float(spatial_scale), _p(gi))
    return gi


def moi_pool_forward(inp, rois, oh_labels, superpixels, spatial_scale, pooled, return_mask=False):
    inp, rois, oh, sp = _f32(inp), _f32(rois), _i32(oh_labels), _i32(superpixels)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.empty((r, c, ph, pw), np.float32)
    arg = np.empty((r, c, ph, pw), np.int32)
    mois = np.empty((r, h, w), np.int32)
    wsl().ref_moi_pool_forward(_p(inp), _p(rois), r, _p(oh), oh.shape[1], _p(sp), sp.shape[1], sp.shape[2],
                               c, h, w, ph, pw, ctypes.c_float(spatial_scale), _p(out), _p(arg), _p(mois))
    return (out, arg, mois) if return_mask else (out, arg)


def moi_pool_backward(grad_out, argmax, rois, input_shape, spatial_scale=1.0):
    grad_out, argmax, rois = _f32(grad_out), _i32(argmax), _f32(rois)
    n, c, h, w = input_shape
    ph, pw = grad_out.shape[2:]
    gi = np.empty(input_shape, np.float32)
    wsl().ref_moi_pool_backward(_p(grad_out), _p(argmax), _p(rois), rois.shape[0], n, c, h, w, ph, pw,
                                ctypes.c_float(spatial_scale), _p(gi))
    return gi


def roi_align_forward(inp, rois, spatial_scale, pooled, sampling_ratio, aligned):
    inp, rois = _f32(inp), _f32(rois)
    n, c, h, w = inp.shape
    r = rois.shape[0]
    ph, pw = pooled
    out = np.empty((r, c, ph, pw), np.float32)
    align().ref_roi_align_forward(_p(inp), n, c, h, w, _p(rois), r, ctypes.c_float(spatial_scale), ph, pw,
                                  int(sampling_ratio), int(bool(aligned)), _p(out))
    return out


def roi_align_backward(grad_out, rois, spatial_scale, input_shape, sampling_ratio, aligned):
    grad_out, rois = _f32(grad_out), _f32(rois)
    n, c, h, w = input_shape
    r, _, ph, pw = grad_out.shape
    gi = np.empty(input_shape, np.float32)
    align().ref_roi_align_backward(_p(grad_out), _p(rois), r, ctypes.c_float(spatial_scale), ph, pw, n, c, h, w,
                                   int(sampling_ratio), int(bool(aligned)), _p(gi))
    return gi


def pcl_loss_forward(probs, labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels):
    a = [_f32(x).reshape(-1) for x in (labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels)]
    probs = _f32(probs)
    r, c = probs.shape
    out = np.zeros(c, np.float32)
    pcl().ref_pcl_loss_forward(_p(probs), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(a[5]), r, c,
                               a[2].shape[0], _p(out))
    return out


def pcl_loss_backward(probs, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                      img_cls_loss_weights, im_labels):
    a = [_f32(x).reshape(-1) for x in (labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                                       img_cls_loss_weights, im_labels)]
    probs = _f32(probs)
    r, c = probs.shape
    g = np.zeros((r, c), np.float32)
    pcl().ref_pcl_loss_backward(_p(probs), *[_p(x) for x in a], r, c, a[3].shape[0], _p(g))
    return g


# ---------------------------------------------------------------- ROIMerge / ROILabel (CMIL / UWSOD heads)
WSLCMIL = os.path.join(OUT, "libref_wslcmil.so")
_cmil = None


def build_wsl_cmil():
    return _build_aten("libref_wslcmil.so", ("projects", "WSL", "wsl", "layers", "csrc", "ROIMerge", "ROIMerge_cpu.cpp"),
                       WSLCMIL)


def wsl_cmil_available():
    return build_wsl_cmil() is not None


def cmil():
    global _cmil
    if _cmil is None:
        path = build_wsl_cmil()
        if path is None:
            raise RuntimeError("oracle/_ref/libref_wslcmil.so missing and /root/reference not present")
        import torch  # noqa: F401

        _cmil = ctypes.CDLL(path)
        _cmil.ref_roi_merge_forward.restype = ctypes.c_int
    return _cmil


def roi_merge_forward(S, J, C, D, P):
    """The reference's ROIMerge_forward_cpu.  P: int32[8] state (copied, returned updated).
    -> (MC [num_id,K], MD, I int32 [R], IC int32 [R], P)."""
    S, J, C, D = _f32(S).reshape(-1), _f32(J), _f32(C), _f32(D)
    P = _i32(P).copy()
    r, k = C.shape
    mc, md = np.zeros((r, k), np.float32), np.zeros((r, k), np.float32)
    i, ic = np.zeros(r, np.int32), np.zeros(r, np.int32)
    num_id = cmil().ref_roi_merge_forward(_p(S), _p(J), _p(C), _p(D), _p(P), r, k, _p(mc), _p(md), _p(i), _p(ic))
    return mc[:num_id].copy(), md[:num_id].copy(), i, ic, P


def roi_merge_backward(C, D, GMC, GMD, I, IC):
    C, D, GMC, GMD = _f32(C), _f32(D), _f32(GMC), _f32(GMD)
    I, IC = _i32(I), _i32(IC)
    r, k = C.shape
    gc, gd = np.zeros((r, k), np.float32), np.zeros((r, k), np.float32)
    cmil().ref_roi_merge_backward(_p(C), _p(D), _p(GMC), _p(GMD), _p(I), _p(IC), r, k, GMC.shape[0], _p(gc), _p(gd))
    return gc, gd


def roi_label_perm(num_roi):
    """The visiting order the reference's ROILabel call draws in this build (fixed seed; see the driver)."""
    perm = np.zeros(num_roi, np.int32)
    cmil().ref_roi_label_perm(int(num_roi), _p(perm))
    return perm


def roi_label_forward(S, U, L, CW, P):
    """The reference's ROILabel_forward_cpu.  P: float32[14] state (copied, returned updated).  CW may be None.
    -> (RL int32 [R], RW float32 [R], P)."""
    S, U, L = _f32(S), _f32(U), _f32(L).reshape(-1)
    P = _f32(P).copy()
    r, ks = S.shape
    rl, rw = np.zeros(r, np.int32), np.zeros(r, np.float32)
    cw = _f32(CW).reshape(-1) if CW is not None else None
    cmil().ref_roi_label_forward(_p(S), _p(U), _p(L), _p(cw) if cw is not None else None, _p(P), r, ks, L.shape[0],
                                 _p(rl), _p(rw))
    return rl, rw, P
