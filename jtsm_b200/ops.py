"""Torch-facing op layer: shape/dtype checks, allocation, and the ctypes call into
libjtsm_b200.so on the current CUDA stream.  Mirrors the role of the reference's
pybind modules `detectron2._C` (detectron2/layers/csrc/vision.cpp:70-117) and
`wsl._C` (projects/WSL/wsl/layers/csrc/vision.cpp:15-33): same op names and
argument meaning, tensors in / tensors out.

No CPU path exists: every op raises on non-CUDA tensors.
"""
import torch

from . import _lib
from ._lib import check, lib

_DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}


def _pair(x):
    return (int(x), int(x)) if isinstance(x, int) else (int(x[0]), int(x[1]))


def _dtype_code(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise RuntimeError("jtsm_b200: unsupported dtype %s (float32/bfloat16/float16 only)" % t.dtype)


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("jtsm_b200 ops are CUDA-only (sm_100a); got a %s tensor" % t.device)


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)


def _rois_f32(rois):
    # The reference casts rois to the feature dtype (roi_align.py:108); coordinates are
    # kept in fp32 here so bin edges do not depend on the feature precision.
    return rois.detach().to(torch.float32).contiguous()


# --------------------------------------------------------------------------- RoIPool
def roi_pool_forward(input, rois, spatial_scale, output_size, row_scale=None, argmax_dtype=None):
    """-> (output [R,C,PH,PW], argmax).  argmax_dtype: torch.int32 (default for API-visible
    argmax) or torch.int16 (uint16 payload; needs H*W < 65535)."""
    _require_cuda(input, rois, row_scale)
    assert rois.dim() == 2 and rois.size(1) == 5
    ph, pw = _pair(output_size)
    n, c, h, w = input.shape
    r = rois.size(0)
    input = input.contiguous()
    rois = _rois_f32(rois)
    if argmax_dtype is None:
        argmax_dtype = torch.int32
    a_code = _lib.ARGMAX_I32 if argmax_dtype == torch.int32 else _lib.ARGMAX_U16
    output = torch.empty((r, c, ph, pw), dtype=input.dtype, device=input.device)
    argmax = torch.empty((r, c, ph, pw), dtype=argmax_dtype, device=input.device)
    if r == 0 or c == 0:
        return output, argmax
    if row_scale is not None:
        row_scale = row_scale.detach().to(torch.float32).contiguous()
        assert row_scale.numel() == r
    L = lib()
    ws_bytes = L.jtsm_roi_pool_workspace_bytes(n, r, ph, pw)
    ws = _workspace(ws_bytes, input.device)
    with torch.cuda.device(input.device):
        check(
            L.jtsm_roi_pool_forward(
                _ptr(input), _ptr(rois), _dtype_code(input), n, c, h, w, r, float(spatial_scale), ph, pw,
                _ptr(row_scale), _ptr(output), _ptr(argmax), a_code, _ptr(ws), ws_bytes, _stream(),
            ),
            "roi_pool_forward",
        )
    return output, argmax


def roi_pool_backward(grad, rois, argmax, input_shape, output_size, row_scale=None):
    _require_cuda(grad, rois, argmax, row_scale)
    ph, pw = _pair(output_size)
    n, c, h, w = (int(v) for v in input_shape)
    r = rois.size(0)
    rois = _rois_f32(rois)
    argmax = argmax.contiguous()
    a_code = _lib.ARGMAX_I32 if argmax.dtype == torch.int32 else _lib.ARGMAX_U16
    grad_input = torch.empty((n, c, h, w), dtype=grad.dtype, device=grad.device)
    if grad_input.numel() == 0:
        return grad_input
    if row_scale is not None:
        row_scale = row_scale.detach().to(torch.float32).contiguous()
    L = lib()
    dt = _dtype_code(grad)
    ws_bytes = L.jtsm_roi_pool_backward_workspace_bytes(dt, n, c, h, w)
    ws = _workspace(ws_bytes, grad.device)
    sn, sc, sh, sw = grad.stride() if grad.numel() else (0, 0, 0, 0)
    with torch.cuda.device(grad.device):
        check(
            L.jtsm_roi_pool_backward(
                _ptr(grad), sn, sc, sh, sw, _ptr(rois), _ptr(argmax), a_code, dt, n, c, h, w, r, ph, pw,
                _ptr(row_scale), _ptr(grad_input), _ptr(ws), ws_bytes, _stream(),
            ),
            "roi_pool_backward",
        )
    return grad_input
