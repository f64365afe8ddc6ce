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
    """RoI dtype rule (one rule for every pooling op, stated in include/jtsm_b200.h): the kernels read box
    coordinates as fp32 and an op uses the VALUES the caller passes, widened losslessly (bf16 / fp16 -> fp32 is exact).
    The reference's `ROIAlign` MODULE rounds the coordinates through the feature dtype first
    (`rois.to(dtype=input.dtype)`, detectron2/layers/roi_align.py:108) before its op sees them; layers/roi_align.py does
    the same cast, so module == reference module and op == reference op on the same tensors."""
    return _aligned16(rois.detach().to(torch.float32).contiguous())


def _aligned16(t):
    """Kernels read boxes / RoI rows with 128-bit loads: a contiguous view at an odd storage offset is cloned."""
    return t.clone() if t.numel() and t.data_ptr() % 16 else t


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


# --------------------------------------------------------------------------- segments
_SEG_CACHE = {}


def segment_offsets(counts, device):
    """int32 device tensor [N+1] of row offsets for per-image row counts (cached per device)."""
    key = (tuple(int(c) for c in counts), str(device))
    t = _SEG_CACHE.get(key)
    if t is None:
        offs = [0]
        for c in key[0]:
            offs.append(offs[-1] + c)
        t = torch.tensor(offs, dtype=torch.int32).to(device)
        if len(_SEG_CACHE) > 256:
            _SEG_CACHE.clear()
        _SEG_CACHE[key] = t
    return t


def _f32c(t):
    return _aligned16(t.detach().to(torch.float32).contiguous())


# --------------------------------------------------------------------------- WSDDN MIL head
def wsddn_forward(cls_logits, det_logits, num_preds_per_image, labels_onehot=None, mean_loss=True):
    """-> dict(scores [R,K], img_scores [N,K] (clamped), loss [1] | None, grad_img, col_stats, seg).
    fast_rcnn_wsddn.py:571-584, 346-379."""
    _require_cuda(cls_logits, det_logits, labels_onehot)
    C, D = _f32c(cls_logits), _f32c(det_logits)
    r, k = C.shape
    n = len(num_preds_per_image)
    assert sum(int(c) for c in num_preds_per_image) == r and D.shape == C.shape
    dev = C.device
    seg = segment_offsets(num_preds_per_image, dev)
    scores = torch.empty_like(C)
    img_scores = torch.empty((n, k), dtype=torch.float32, device=dev)
    col_stats = torch.empty((n, k, 3), dtype=torch.float32, device=dev)
    loss = grad_img = None
    if labels_onehot is not None:
        labels_onehot = _f32c(labels_onehot)
        assert labels_onehot.shape == (n, k)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        grad_img = torch.empty((n, k), dtype=torch.float32, device=dev)
    L = lib()
    ws_bytes = L.jtsm_wsddn_workspace_bytes(n)
    ws = _workspace(ws_bytes, dev)
    with torch.cuda.device(dev):
        check(L.jtsm_wsddn_forward(_ptr(C), _ptr(D), _ptr(seg), n, r, k, _ptr(labels_onehot), int(bool(mean_loss)),
                                   _ptr(scores), _ptr(img_scores), _ptr(col_stats), _ptr(loss), _ptr(grad_img),
                                   _ptr(ws), ws_bytes, _stream()), "wsddn_forward")
    return dict(scores=scores, img_scores=img_scores, loss=loss, grad_img=grad_img, col_stats=col_stats, seg=seg,
                cls=C, det=D, n=n)


def wsddn_backward(fwd, grad_scores=None, grad_loss=None, use_loss=True, grad_img=None):
    """-> (grad_cls, grad_det) for the state returned by wsddn_forward.  Upstream gradient per element:
    grad_scores[r,k] + gi[n(r),k] * grad_loss, with gi = `grad_img` when given (d / d unclamped image sum, e.g. from
    autograd), else the BCE gradient the forward saved (use_loss)."""
    C, D = fwd["cls"], fwd["det"]
    r, k = C.shape
    dC, dD = torch.empty_like(C), torch.empty_like(D)
    if r == 0:
        return dC, dD
    if grad_scores is not None:
        grad_scores = _f32c(grad_scores)
    if grad_loss is not None:
        grad_loss = _f32c(grad_loss).reshape(1)
    gi = _f32c(grad_img) if grad_img is not None else (fwd.get("grad_img") if use_loss else None)
    with torch.cuda.device(C.device):
        check(lib().jtsm_wsddn_backward(_ptr(C), _ptr(D), _ptr(fwd["scores"]), _ptr(fwd["col_stats"]), _ptr(fwd["seg"]),
                                        fwd["n"], r, k, _ptr(grad_scores), _ptr(gi), _ptr(grad_loss), _ptr(dC), _ptr(dD),
                                        _stream()), "wsddn_backward")
    return dC, dD


# --------------------------------------------------------------------------- refinement branch
_CLS_CACHE = {}


def image_class_lists(gt_classes_img_int, device):
    """(offsets int32 [N+1], ids int64 [total], max per image, total) on `device` for ragged per-image class lists.
    A ready (offsets, ids, max per image, capacity) tuple of DEVICE tensors passes through: the lists then live in
    device memory that the caller refreshes per step (CUDA-graph streaming, hotpath.HotPathStep) and `capacity`
    only bounds the pseudo-gt output rows."""
    if isinstance(gt_classes_img_int, tuple) and len(gt_classes_img_int) == 4 and torch.is_tensor(gt_classes_img_int[0]):
        return gt_classes_img_int
    key = (tuple(tuple(int(v) for v in (c.tolist() if torch.is_tensor(c) else c)) for c in gt_classes_img_int), str(device))
    hit = _CLS_CACHE.get(key)
    if hit is None:
        offs, ids = [0], []
        for c in key[0]:
            ids.extend(c)
            offs.append(len(ids))
        hit = (torch.tensor(offs, dtype=torch.int32).to(device), torch.tensor(ids, dtype=torch.int64).to(device),
               max([len(c) for c in key[0]] + [0]), len(ids))
        if len(_CLS_CACHE) > 256:
            _CLS_CACHE.clear()
        _CLS_CACHE[key] = hit
    return hit


def refine_assign(prev_scores, prev_boxes, proposals, num_preds_per_image, gt_classes_img_int, img_scores, num_classes,
                  iou_thresholds=(0.5,), iou_labels=(0, 1), want_iou=False):
    """Pseudo-gt (top-1 per image class) + IoU label assignment (roi_heads_oicr.py:675-811, roi_heads.py:263-370)."""
    _require_cuda(prev_scores, prev_boxes, proposals, img_scores)
    S, P = _f32c(prev_scores), _f32c(proposals)
    Bx = _f32c(prev_boxes)
    dev = S.device
    r = S.shape[0]
    n = len(num_preds_per_image)
    bpr = 1 if Bx.numel() == r * 4 else num_classes
    assert Bx.numel() == r * bpr * 4 and P.shape == (r, 4)
    seg = segment_offsets(num_preds_per_image, dev)
    coff, cids, cmax, ctot = image_class_lists(gt_classes_img_int, dev)
    img_scores = _f32c(img_scores)
    assert img_scores.shape == (n, num_classes)
    thr = list(iou_thresholds)
    lab = list(iou_labels)
    assert len(thr) in (1, 2) and len(lab) == len(thr) + 1
    # one arena per dtype instead of 13 small tensors; the kernel writes every element, so no fill
    fl = torch.empty(ctot * 6 + r * 8 + (r if want_iou else 0), dtype=torch.float32, device=dev)
    il = torch.empty(ctot + 2 * r, dtype=torch.int64, device=dev)
    fo, io = [0], [0]

    def f(*shape):
        n_ = 1
        for v in shape:
            n_ *= v
        t = fl[fo[0]:fo[0] + n_].view(*shape)
        fo[0] += n_
        return t

    def i64(*shape):
        t = il[io[0]:io[0] + shape[0]]
        io[0] += shape[0]
        return t

    out = dict(pgt_boxes=f(ctot, 4), gt_boxes=f(r, 4), pgt_scores=f(ctot), pgt_weights=f(ctot),
               matched_iou=f(r) if want_iou else None, gt_weights=f(r), gt_scores=f(r), loss_weights=f(r),
               pgt_classes=i64(ctot), matched_idx=i64(r), gt_classes=i64(r),
               counters=torch.empty(n * 3 + 1, dtype=torch.int32, device=dev), pgt_offsets=coff)
    with torch.cuda.device(dev):
        check(lib().jtsm_refine_assign(
            _ptr(S), S.shape[1], _ptr(Bx), bpr, _ptr(P), _ptr(seg), n, r, num_classes, _ptr(coff), _ptr(cids), cmax,
            _ptr(img_scores), float(thr[0]), float(thr[1]) if len(thr) > 1 else 0.0, int(lab[0]), int(lab[1]),
            int(lab[2]) if len(lab) > 2 else 0, len(thr), _ptr(out["pgt_boxes"]), _ptr(out["pgt_classes"]),
            _ptr(out["pgt_scores"]), _ptr(out["pgt_weights"]), _ptr(out["matched_idx"]), _ptr(out["matched_iou"]),
            _ptr(out["gt_classes"]), _ptr(out["gt_weights"]), _ptr(out["gt_scores"]), _ptr(out["gt_boxes"]),
            _ptr(out["loss_weights"]), _ptr(out["counters"]), _stream()), "refine_assign")
    out["valid_count"] = out["counters"][n * 3:]
    return out


def refine_match(pgt_boxes, pgt_classes, pgt_scores, pgt_weights, seeds_per_image, proposals, num_preds_per_image,
                 num_classes, iou_thresholds=(0.5,), iou_labels=(0, 1), want_iou=False, seed_counts=None):
    """label_and_sample_proposals for caller-supplied pseudo-gt seeds (roi_heads.py:263-370): seeds of all images
    concatenated ([G,4], [G] int64, [G], [G]) with `seeds_per_image` counts (host ints).  Same outputs as
    refine_assign (without the pgt_* it did not build).
    seed_counts (int32 DEVICE tensor [N], optional): `seeds_per_image` are then SLOT sizes and image n uses the first
    seed_counts[n] entries of its slot -- data-dependent seed sets without a host read."""
    _require_cuda(pgt_boxes, proposals)
    P = _f32c(proposals)
    dev = P.device
    r, n = P.shape[0], len(num_preds_per_image)
    assert len(seeds_per_image) == n
    gb, gs, gw = _f32c(pgt_boxes).view(-1, 4), _f32c(pgt_scores).view(-1), _f32c(pgt_weights).view(-1)
    gc = pgt_classes.to(torch.int64).contiguous().view(-1)
    g_total = int(sum(seeds_per_image))
    assert gb.shape[0] == g_total and gc.numel() == g_total and gs.numel() == g_total and gw.numel() == g_total
    seg = segment_offsets(num_preds_per_image, dev)
    soff = segment_offsets(seeds_per_image, dev)
    thr, lab = list(iou_thresholds), list(iou_labels)
    assert len(thr) in (1, 2) and len(lab) == len(thr) + 1
    fl = torch.empty(r * 8 + (r if want_iou else 0), dtype=torch.float32, device=dev)
    il = torch.empty(2 * r, dtype=torch.int64, device=dev)
    out = dict(gt_boxes=fl[:4 * r].view(r, 4), gt_weights=fl[4 * r:5 * r], gt_scores=fl[5 * r:6 * r],
               loss_weights=fl[6 * r:7 * r], matched_iou=fl[8 * r:9 * r] if want_iou else None,
               matched_idx=il[:r], gt_classes=il[r:], counters=torch.empty(n * 3 + 1, dtype=torch.int32, device=dev))
    if seed_counts is not None:
        seed_counts = seed_counts.to(device=dev, dtype=torch.int32).contiguous()
        assert seed_counts.numel() == n
    with torch.cuda.device(dev):
        check(lib().jtsm_refine_match_counted(
            _ptr(gb), _ptr(gc), _ptr(gs), _ptr(gw), _ptr(soff), _ptr(seed_counts),
            max([0] + [int(v) for v in seeds_per_image]), _ptr(P),
            _ptr(seg), n, r, num_classes, float(thr[0]), float(thr[1]) if len(thr) > 1 else 0.0, int(lab[0]), int(lab[1]),
            int(lab[2]) if len(lab) > 2 else 0, len(thr), _ptr(out["matched_idx"]), _ptr(out["matched_iou"]),
            _ptr(out["gt_classes"]), _ptr(out["gt_weights"]), _ptr(out["gt_scores"]), _ptr(out["gt_boxes"]),
            _ptr(out["loss_weights"]), _ptr(out["counters"]), _stream()), "refine_match")
    out["valid_count"] = out["counters"][n * 3:]
    return out


def refine_ce(logits, gt_classes, loss_weights, valid_count, grad_loss=None, loss_scale=1.0, want_probs=True,
              want_grad=True, want_acc=True):
    """Weighted CE fwd+bwd + softmax (fast_rcnn_oicr.py:282-298, 744-762) -> dict(loss, probs, grad_logits, acc)."""
    _require_cuda(logits, gt_classes, loss_weights, valid_count)
    X = _f32c(logits)
    r, kc = X.shape
    dev = X.device
    gt = gt_classes.to(torch.int64).contiguous()
    w = _f32c(loss_weights)
    vc = valid_count.to(torch.int32).contiguous()
    probs = torch.empty_like(X) if want_probs else None
    grad = torch.empty_like(X) if want_grad else None
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    acc = torch.empty(4, dtype=torch.int32, device=dev) if want_acc else None
    if grad_loss is not None:
        grad_loss = _f32c(grad_loss).reshape(1)
    L = lib()
    wsb = L.jtsm_refine_workspace_bytes()
    ws = _workspace(wsb, dev)
    with torch.cuda.device(dev):
        check(L.jtsm_refine_ce(_ptr(X), r, kc, _ptr(gt), _ptr(w), _ptr(vc), _ptr(grad_loss), float(loss_scale), _ptr(probs),
                               _ptr(grad), _ptr(loss), _ptr(acc), _ptr(ws), wsb, _stream()), "refine_ce")
    return dict(loss=loss, probs=probs, grad_logits=grad, acc=acc)


def softmax_rows(logits, out=None, scale=1.0, accumulate=False):
    _require_cuda(logits)
    X = _f32c(logits)
    if out is None:
        out = torch.empty_like(X)
        accumulate = False
    with torch.cuda.device(X.device):
        check(lib().jtsm_softmax_rows(_ptr(X), X.shape[0], X.shape[1], float(scale), int(bool(accumulate)), _ptr(out),
                                      _stream()), "softmax_rows")
    return out


def refine_box_reg(pred_deltas, proposals, gt_boxes, gt_classes, loss_weights, num_classes, weights=(10.0, 10.0, 5.0, 5.0),
                   smooth_l1_beta=0.0, grad_loss=None, loss_scale=1.0, want_grad=True):
    """fast_rcnn_oicr.py:300-380 -> dict(loss, grad_deltas).  loss_weights=None -> plain "smooth_l1"."""
    _require_cuda(pred_deltas, proposals, gt_boxes, gt_classes, loss_weights)
    Dl, P, G = _f32c(pred_deltas), _f32c(proposals), _f32c(gt_boxes)
    r = P.shape[0]
    bpr = Dl.shape[1] // 4
    dev = Dl.device
    gt = gt_classes.to(torch.int64).contiguous()
    w = _f32c(loss_weights) if loss_weights is not None else None
    grad = torch.empty_like(Dl) if want_grad else None
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    if grad_loss is not None:
        grad_loss = _f32c(grad_loss).reshape(1)
    L = lib()
    wsb = L.jtsm_refine_workspace_bytes()
    ws = _workspace(wsb, dev)
    with torch.cuda.device(dev):
        check(L.jtsm_refine_box_reg(_ptr(Dl), bpr, _ptr(P), _ptr(G), _ptr(gt), _ptr(w), r, num_classes,
                                    *(float(v) for v in weights), float(smooth_l1_beta), _ptr(grad_loss), float(loss_scale),
                                    _ptr(loss), _ptr(grad), _ptr(ws), wsb, _stream()), "refine_box_reg")
    return dict(loss=loss, grad_deltas=grad)


def apply_deltas(deltas, boxes, weights=(10.0, 10.0, 5.0, 5.0), scale_clamp=None, delta_scale=1.0):
    """Box2BoxTransform.apply_deltas (box_regression.py:73-109) -> [R, k*4]."""
    import math

    _require_cuda(deltas, boxes)
    Dl, B = _f32c(deltas), _f32c(boxes)
    r = B.shape[0]
    bpr = Dl.shape[1] // 4 if r else 1
    out = torch.empty_like(Dl)
    if scale_clamp is None:
        scale_clamp = math.log(1000.0 / 16)
    with torch.cuda.device(Dl.device):
        check(lib().jtsm_apply_deltas(_ptr(Dl), _ptr(B), r, bpr, *(float(v) for v in weights), float(scale_clamp),
                                      float(delta_scale), _ptr(out), _stream()), "apply_deltas")
    return out


def pairwise_iou(boxes1, boxes2):
    """detectron2/structures/boxes.py:369-391 on raw [n,4] tensors -> [n1, n2]."""
    _require_cuda(boxes1, boxes2)
    A, B = _f32c(boxes1), _f32c(boxes2)
    out = torch.empty((A.shape[0], B.shape[0]), dtype=torch.float32, device=A.device)
    with torch.cuda.device(A.device):
        check(lib().jtsm_pairwise_iou(_ptr(A), A.shape[0], _ptr(B), B.shape[0], _ptr(out), _stream()), "pairwise_iou")
    return out


# --------------------------------------------------------------------------- NMS
def batched_nms_raw(boxes, scores, idxs, iou_threshold):
    """-> (keep int64 [n] padded with -1, num_keep int64 [1]) without any host sync."""
    _require_cuda(boxes, scores, idxs)
    B, S = _f32c(boxes), _f32c(scores)
    n = B.shape[0]
    dev = B.device
    I = idxs.to(torch.int64).contiguous() if idxs is not None else None
    keep = torch.empty(n, dtype=torch.int64, device=dev)
    num = torch.zeros(1, dtype=torch.int64, device=dev)
    L = lib()
    wsb = L.jtsm_nms_workspace_bytes(n)
    ws = _workspace(wsb, dev)
    with torch.cuda.device(dev):
        check(L.jtsm_batched_nms(_ptr(B), _ptr(S), _ptr(I), n, float(iou_threshold), _ptr(keep), _ptr(num), _ptr(ws), wsb,
                                 _stream()), "batched_nms")
    return keep, num


def batched_nms(boxes, scores, idxs, iou_threshold):
    """detectron2.layers.batched_nms signature (nms.py:10-30): int64 keep indices, descending score.  The returned
    tensor's LENGTH is data dependent, so this signature costs one host read of the count (the reference's returns
    after a device -> host mask copy); device-resident callers use batched_nms_raw / fast_rcnn_inference_raw."""
    assert boxes.shape[-1] == 4
    keep, num = batched_nms_raw(boxes.float(), scores, idxs, iou_threshold)
    return keep[: int(num.item())]


def fast_rcnn_inference_raw(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image):
    """fast_rcnn_inference_single_image (fast_rcnn_oicr.py:95-163) in one C-ABI call, no host sync.
    boxes [R, 4] or [R, 4K], scores [R, K+1] -> dict of PADDED device tensors with `cap` rows (cap = topk, or R*K when
    topk < 0): pred_boxes [cap,4], scores [cap], pred_classes / rows / pred_inds int64 [cap] (-1 padding), num int64 [1]."""
    _require_cuda(boxes, scores)
    B, S = _f32c(boxes), _f32c(scores)
    r, k1 = S.shape
    k = k1 - 1
    nb = B.shape[1] // 4 if r else 1
    dev = S.device
    cap = int(topk_per_image) if topk_per_image >= 0 else r * k
    out = dict(pred_boxes=torch.empty((cap, 4), dtype=torch.float32, device=dev),
               scores=torch.empty(cap, dtype=torch.float32, device=dev),
               pred_classes=torch.empty(cap, dtype=torch.int64, device=dev),
               rows=torch.empty(cap, dtype=torch.int64, device=dev),
               pred_inds=torch.empty(cap, dtype=torch.int64, device=dev),
               num=torch.zeros(1, dtype=torch.int64, device=dev))
    if r == 0 or k <= 0:
        return out
    L = lib()
    wsb = L.jtsm_fast_rcnn_inference_workspace_bytes(r, k)
    if wsb == 0:
        raise RuntimeError("jtsm_b200.fast_rcnn_inference: R x K too large")
    ws = _workspace(wsb, dev)
    h, w = image_shape
    with torch.cuda.device(dev):
        check(L.jtsm_fast_rcnn_inference(_ptr(B), nb, _ptr(S), r, k, float(h), float(w), float(score_thresh),
                                         float(nms_thresh), int(topk_per_image), cap, _ptr(out["pred_boxes"]),
                                         _ptr(out["scores"]), _ptr(out["pred_classes"]), _ptr(out["rows"]),
                                         _ptr(out["pred_inds"]), _ptr(out["num"]), _ptr(ws), wsb, _stream()),
              "fast_rcnn_inference")
    return out


def nms(boxes, scores, iou_threshold):
    """torchvision.ops.nms signature."""
    return batched_nms(boxes, scores, None, iou_threshold)


# --------------------------------------------------------------------------- RoIAlign
def roi_align_forward(input, rois, spatial_scale, pooled_height, pooled_width, sampling_ratio, aligned):
    """Same signature as detectron2._C.roi_align_forward (ROIAlign.h:54-61)."""
    _require_cuda(input, rois)
    assert rois.dim() == 2 and rois.size(1) == 5
    n, c, h, w = input.shape
    r = rois.size(0)
    input = input.contiguous()
    rois = _rois_f32(rois)
    output = torch.empty((r, c, int(pooled_height), int(pooled_width)), dtype=input.dtype, device=input.device)
    if r == 0 or c == 0:
        return output
    L = lib()
    dt = _dtype_code(input)
    wsb = L.jtsm_roi_align_workspace_bytes(dt, n, c, h, w, r, 0)
    ws = _workspace(wsb, input.device)
    with torch.cuda.device(input.device):
        check(L.jtsm_roi_align_forward(_ptr(input), _ptr(rois), dt, n, c, h, w, r, float(spatial_scale),
                                       int(pooled_height), int(pooled_width), int(sampling_ratio), int(bool(aligned)),
                                       _ptr(output), _ptr(ws), wsb, _stream()), "roi_align_forward")
    return output


def roi_align_backward(grad, rois, spatial_scale, pooled_height, pooled_width, batch_size, channels, height, width,
                       sampling_ratio, aligned, deterministic=None):
    """Same signature as detectron2._C.roi_align_backward (ROIAlign.h:86-97).
    deterministic: True = the segmented-reduction backward (no atomics; run-to-run bit-identical and bit-equal to
    the reference's serial CPU backward; sampling_ratio > 0 only); None = follow
    torch.are_deterministic_algorithms_enabled() where it applies; False = the RED kernels."""
    _require_cuda(grad, rois)
    n, c, h, w = int(batch_size), int(channels), int(height), int(width)
    r = rois.size(0)
    rois = _rois_f32(rois)
    grad_input = torch.empty((n, c, h, w), dtype=grad.dtype, device=grad.device)
    if grad_input.numel() == 0:
        return grad_input
    L = lib()
    dt = _dtype_code(grad)
    if deterministic and int(sampling_ratio) <= 0:
        raise RuntimeError("jtsm_b200.roi_align_backward: the deterministic backward needs sampling_ratio > 0")
    if deterministic is None:
        deterministic = torch.are_deterministic_algorithms_enabled() and int(sampling_ratio) > 0
    if deterministic:
        ph, pw, sr = int(pooled_height), int(pooled_width), int(sampling_ratio)
        wsb = L.jtsm_roi_align_backward_det_workspace_bytes(n, c, h, w, r, ph, pw, sr)
        ws = _workspace(wsb, grad.device)
        sn, sc, sh, sw = grad.stride() if grad.numel() else (0, 0, 0, 0)
        with torch.cuda.device(grad.device):
            check(L.jtsm_roi_align_backward_det(_ptr(grad), sn, sc, sh, sw, _ptr(rois), dt, n, c, h, w, r,
                                                float(spatial_scale), ph, pw, sr, int(bool(aligned)), _ptr(grad_input),
                                                _ptr(ws), wsb, _stream()), "roi_align_backward_det")
        return grad_input
    wsb = L.jtsm_roi_align_workspace_bytes(dt, n, c, h, w, r, 1)
    ws = _workspace(wsb, grad.device)
    sn, sc, sh, sw = grad.stride() if grad.numel() else (0, 0, 0, 0)
    with torch.cuda.device(grad.device):
        check(L.jtsm_roi_align_backward(_ptr(grad), sn, sc, sh, sw, _ptr(rois), dt, n, c, h, w, r, float(spatial_scale),
                                        int(pooled_height), int(pooled_width), int(sampling_ratio), int(bool(aligned)),
                                        _ptr(grad_input), _ptr(ws), wsb, _stream()), "roi_align_backward")
    return grad_input


# --------------------------------------------------------------------------- WSL pools
def roi_loop_pool_forward(input, rois, spatial_scale, pooled_height, pooled_width, context_ratio=1.8):
    """wsl._C.roi_loop_pool_forward (ROILoopPool.h:48-53) -> (output [3R,C,PH,PW], argmax int32)."""
    _require_cuda(input, rois)
    assert rois.dim() == 2 and rois.size(1) == 5
    n, c, h, w = input.shape
    r = rois.size(0)
    ph, pw = int(pooled_height), int(pooled_width)
    input = input.contiguous()
    rois = _rois_f32(rois)
    output = torch.empty((3 * r, c, ph, pw), dtype=input.dtype, device=input.device)
    argmax = torch.empty((3 * r, c, ph, pw), dtype=torch.int32, device=input.device)
    if r == 0 or c == 0:
        return output, argmax
    L = lib()
    wsb = L.jtsm_roi_loop_pool_workspace_bytes(n, h, w, r, ph, pw)
    ws = _workspace(wsb, input.device)
    with torch.cuda.device(input.device):
        check(L.jtsm_roi_loop_pool_forward(_ptr(input), _ptr(rois), _dtype_code(input), n, c, h, w, r,
                                           float(spatial_scale), float(context_ratio), ph, pw, _ptr(output),
                                           _ptr(argmax), _ptr(ws), wsb, _stream()), "roi_loop_pool_forward")
    return output, argmax


def roi_loop_pool_backward(grad, rois, argmax, spatial_scale, pooled_height, pooled_width, batch_size, channels,
                           height, width):
    """wsl._C.roi_loop_pool_backward (ROILoopPool.h:67-77): rows [R,2R) and [2R,3R) belong to RoI n % R."""
    return roi_pool_backward(grad, rois.repeat(3, 1), argmax, (batch_size, channels, height, width),
                             (pooled_height, pooled_width))


def moi_pool_forward(input, rois, spatial_scale, pooled_height, pooled_width, oh_labels, superpixels, row_scale=None,
                     mask_scale=False, return_scales=False):
    """wsl._C.moi_pool_forward (MOIPool.h:52-59) -> (output [R,C,PH,PW], argmax int32).
    row_scale [R] / mask_scale: the JTSM heads' per-RoI factors (roi_heads_jtsm.py:615-633) fused into the
    epilogue; return_scales adds the [3,R] tensor (mask scales, row scales, products) for the backward."""
    _require_cuda(input, rois, oh_labels, superpixels)
    assert rois.dim() == 2 and rois.size(1) == 5
    n, c, h, w = input.shape
    r = rois.size(0)
    ph, pw = int(pooled_height), int(pooled_width)
    input = input.contiguous()
    rois = _rois_f32(rois)
    oh = oh_labels.to(torch.int32).contiguous()
    sp = superpixels.to(torch.int32).contiguous()
    assert oh.dim() == 2 and oh.size(0) == r and sp.dim() == 3 and sp.size(0) == n
    output = torch.empty((r, c, ph, pw), dtype=input.dtype, device=input.device)
    argmax = torch.empty((r, c, ph, pw), dtype=torch.int32, device=input.device)
    fused = row_scale is not None or mask_scale
    scales = torch.ones((3, r), dtype=torch.float32, device=input.device) if fused else None
    rs = _f32c(row_scale) if row_scale is not None else None
    if r == 0 or c == 0:
        return (output, argmax, scales) if return_scales else (output, argmax)
    L = lib()
    wsb = L.jtsm_moi_pool_workspace_bytes(n, h, w, r, ph, pw)
    ws = _workspace(wsb, input.device)
    with torch.cuda.device(input.device):
        check(L.jtsm_moi_pool_forward(_ptr(input), _ptr(rois), _ptr(oh), oh.size(1), _ptr(sp), sp.size(1), sp.size(2),
                                      _dtype_code(input), n, c, h, w, r, float(spatial_scale), ph, pw, _ptr(rs),
                                      int(bool(mask_scale)), _ptr(scales), _ptr(output), _ptr(argmax), _ptr(ws), wsb,
                                      _stream()), "moi_pool_forward")
    return (output, argmax, scales) if return_scales else (output, argmax)


def moi_pool_backward(grad, rois, argmax, spatial_scale, pooled_height, pooled_width, batch_size, channels, height,
                      width, row_scale=None):
    """wsl._C.moi_pool_backward (MOIPool.h:85-95)."""
    return roi_pool_backward(grad, rois, argmax, (batch_size, channels, height, width), (pooled_height, pooled_width),
                             row_scale=row_scale)


# --------------------------------------------------------------------------- TTA-AVG merge
TTA_MAX_OPS, TTA_MAX_AUG = 4, 32
TTA_END, TTA_NOOP, TTA_HFLIP, TTA_SCALE = 0, 1, 2, 3


def _tta_op_arrays(op_lists):
    """[[(type, a, b), ...] per augmentation] -> flat ctypes arrays [len * TTA_MAX_OPS] (type, a, b)."""
    import ctypes

    n = len(op_lists) * TTA_MAX_OPS
    ty, aa, bb = (ctypes.c_int * n)(), (ctypes.c_float * n)(), (ctypes.c_float * n)()
    for i, ops_ in enumerate(op_lists):
        if len(ops_) > TTA_MAX_OPS:
            raise RuntimeError("jtsm_b200: a transform list has more than %d stages" % TTA_MAX_OPS)
        for s, (t, a, b) in enumerate(ops_):
            ty[i * TTA_MAX_OPS + s], aa[i * TTA_MAX_OPS + s], bb[i * TTA_MAX_OPS + s] = int(t), float(a), float(b)
    return ty, aa, bb


def tta_merge(all_boxes, all_scores, op_lists):
    """Mean over augmentations of inverse-transformed boxes and of scores
    (GeneralizedRCNNWithTTAAVG._get_augmented_boxes, test_time_augmentation_avg.py:364-390).
    all_boxes: list of A tensors [R, 4k] (or [1, R, 4k]); all_scores: list of A tensors [R, S] (or [1, R, S]) or None;
    op_lists: per augmentation the INVERSE stage list [(type, a, b), ...].  -> (mean_boxes [R, 4k], mean_scores [R, S])."""
    import ctypes

    A = len(all_boxes)
    if A == 0 or A > TTA_MAX_AUG:
        raise RuntimeError("jtsm_b200.tta_merge: 1..%d augmentations, got %d" % (TTA_MAX_AUG, A))
    _require_cuda(*all_boxes)
    B = [_f32c(b.reshape(b.shape[-2], b.shape[-1])) for b in all_boxes]
    r, cols = B[0].shape
    if cols % 4 or any(b.shape != B[0].shape for b in B):
        raise RuntimeError("jtsm_b200.tta_merge: boxes must all be [R, 4k]")
    S = None
    sc = 0
    if all_scores is not None:
        _require_cuda(*all_scores)
        S = [_f32c(s.reshape(s.shape[-2], s.shape[-1])) for s in all_scores]
        if len(S) != A or any(s.shape != S[0].shape for s in S) or S[0].shape[0] != r:
            raise RuntimeError("jtsm_b200.tta_merge: scores must all be [R, S], one per augmentation")
        sc = S[0].shape[1]
    if len(op_lists) != A:
        raise RuntimeError("jtsm_b200.tta_merge: one transform list per augmentation")
    ty, aa, bb = _tta_op_arrays(op_lists)
    dev = B[0].device
    mb = torch.empty((r, cols), dtype=torch.float32, device=dev)
    ms = torch.empty((r, sc), dtype=torch.float32, device=dev) if S is not None else None
    bp = (ctypes.c_void_p * A)(*[b.data_ptr() for b in B])
    sp = (ctypes.c_void_p * A)(*[s.data_ptr() for s in S]) if S is not None else None
    with torch.cuda.device(dev):
        check(lib().jtsm_tta_merge(bp, sp, A, r, cols // 4, sc, ty, aa, bb, _ptr(mb), _ptr(ms), _stream()), "tta_merge")
    return mb, ms


def tta_transform_boxes(boxes, ops_, image_shape, min_box_size=0.0):
    """transform_proposals (test_time_augmentation_avg.py:27-70): stage list forward, Boxes.clip, Boxes.nonempty.
    boxes [n, 4] -> (boxes [n, 4] clipped, keep bool [n])."""
    _require_cuda(boxes)
    B = _f32c(boxes)
    n = B.shape[0]
    out = torch.empty_like(B)
    keep = torch.empty(n, dtype=torch.uint8, device=B.device)
    ty, aa, bb = _tta_op_arrays([ops_])
    h, w = image_shape
    with torch.cuda.device(B.device):
        check(lib().jtsm_tta_transform_boxes(_ptr(B), n, ty, aa, bb, float(h), float(w), float(min_box_size),
                                             _ptr(out), _ptr(keep), _stream()), "tta_transform_boxes")
    return out, keep.bool()


# --------------------------------------------------------------------------- superpixel IoU
def pairwise_iou_sp(oh_labels1, oh_labels2, cnt_sp, thresholds=None, labels=None):
    """pairwise_iou_sp (roi_heads_jtsm.py:49-73) on raw tensors: oh_labels1 [n1, S], oh_labels2 [n2, S] (int32 flags),
    cnt_sp [S] (pixels per superpixel) -> iou [n1, n2] fp32.  With thresholds / labels (Matcher's constructor
    arguments) -> (iou, matched_idxs int64 [n2], matched_labels int8 [n2])."""
    _require_cuda(oh_labels1, oh_labels2, cnt_sp)
    A = oh_labels1.to(torch.int32).contiguous()
    B = oh_labels2.to(torch.int32).contiguous()
    C = cnt_sp.reshape(-1).to(torch.int64).contiguous()
    n1, n2, s = A.shape[0], B.shape[0], B.shape[1]
    if A.shape[1] != s or C.numel() != s:
        raise RuntimeError("jtsm_b200.pairwise_iou_sp: oh_labels1 [n1,S], oh_labels2 [n2,S], cnt_sp [S] expected")
    dev = B.device
    iou = torch.empty((n1, n2), dtype=torch.float32, device=dev)
    mi = ml = None
    t = [0.0, 0.0]
    lb = [0, 0, 0]
    nt = 1
    if thresholds is not None:
        nt = len(thresholds)
        if nt not in (1, 2) or len(labels) != nt + 1:
            raise RuntimeError("jtsm_b200.pairwise_iou_sp: one or two thresholds, len(labels) = len(thresholds) + 1")
        t[:nt] = [float(v) for v in thresholds]
        lb[:nt + 1] = [int(v) for v in labels]
        mi = torch.empty(n2, dtype=torch.int64, device=dev)
        ml = torch.empty(n2, dtype=torch.int8, device=dev)
    with torch.cuda.device(dev):
        check(lib().jtsm_pairwise_iou_sp(_ptr(A), n1, _ptr(B), n2, _ptr(C), s, _ptr(iou), t[0], t[1], lb[0], lb[1], lb[2],
                                         nt, _ptr(mi), _ptr(ml), _stream()), "pairwise_iou_sp")
    return iou if thresholds is None else (iou, mi, ml)


# --------------------------------------------------------------------------- PCL loss
def _pcl_vec(t, dev):
    return t.reshape(-1).to(device=dev, dtype=torch.float32).contiguous()


def pcl_loss_forward(pcl_probs, labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels):
    """wsl._C.pcl_loss_forward (pcl_loss_cpu.cpp:8-57) + the `output.sum() / R` of pcl_loss.py:52, on the device.
    pcl_probs [R, C] CUDA; the small per-proposal / per-cluster vectors may arrive as CPU tensors (the reference builds
    them with numpy, fast_rcnn_oicr.py:925-935) and are moved once.  -> (loss [1], output [C])."""
    _require_cuda(pcl_probs)
    P = _f32c(pcl_probs)
    dev = P.device
    r, c = P.shape
    v = [_pcl_vec(t, dev) for t in (labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels)]
    if v[0].numel() != r or v[1].numel() != r or v[5].numel() != c or v[3].numel() != v[2].numel() \
            or v[4].numel() != v[2].numel():
        raise RuntimeError("jtsm_b200.pcl_loss_forward: inconsistent shapes")
    out = torch.empty(c, dtype=torch.float32, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().jtsm_pcl_loss_forward(_ptr(P), _ptr(v[0]), _ptr(v[1]), _ptr(v[2]), _ptr(v[3]), _ptr(v[4]), _ptr(v[5]),
                                          r, c, v[2].numel(), _ptr(out), _ptr(loss), _stream()), "pcl_loss_forward")
    return loss, out


def pcl_loss_backward(pcl_probs, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                      img_cls_loss_weights, im_labels):
    """wsl._C.pcl_loss_backward (pcl_loss_cpu.cpp:59-115) followed by `/ R` (pcl_loss.py:84) -> grad [R, C]."""
    _require_cuda(pcl_probs)
    P = _f32c(pcl_probs)
    dev = P.device
    r, c = P.shape
    v = [_pcl_vec(t, dev) for t in (labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count,
                                    img_cls_loss_weights, im_labels)]
    grad = torch.empty_like(P)
    with torch.cuda.device(dev):
        check(lib().jtsm_pcl_loss_backward(_ptr(P), *[_ptr(x) for x in v], r, c, v[3].numel(), _ptr(grad), _stream()),
              "pcl_loss_backward")
    return grad


# --------------------------------------------------------------------------- pool -> fc6 on the tensor cores
def fc6_weight_binmajor(weight, channels, pooled):
    """fc6.weight [N, C*P*P] (column = c*P*P + bin, what flatten([R, C, P, P]) feeds) -> bf16 [N, P*P*C] with
    column = bin*C + c: 64 consecutive columns are 64 consecutive channels of one bin.  One-time re-layout."""
    ph, pw = _pair(pooled)
    n = weight.shape[0]
    return weight.detach().view(n, channels, ph * pw).permute(0, 2, 1).reshape(n, ph * pw * channels) \
        .to(torch.bfloat16).contiguous()


def pool_fc6_forward(feat_nhwc, rois, spatial_scale, output_size, weight_binmajor, bias, row_scale=None, relu=True):
    """relu(fc6(flatten(roi_pool(x, rois) * row_scale))) without the pooled tensor: feat_nhwc [N, H, W, C] bf16
    (channels last), weight_binmajor from fc6_weight_binmajor -> [R, out_features] fp32."""
    _require_cuda(feat_nhwc, rois, weight_binmajor, bias, row_scale)
    if feat_nhwc.dtype != torch.bfloat16 or weight_binmajor.dtype != torch.bfloat16:
        raise RuntimeError("jtsm_b200.pool_fc6_forward: bf16 feature map (NHWC) and bf16 bin-major weight expected")
    ph, pw = _pair(output_size)
    n, h, w, c = feat_nhwc.shape
    r = rois.size(0)
    feat = feat_nhwc.contiguous()
    W = weight_binmajor.contiguous()
    nout = W.shape[0]
    if W.shape[1] != ph * pw * c:
        raise RuntimeError("jtsm_b200.pool_fc6_forward: weight has %d columns, expected %d" % (W.shape[1], ph * pw * c))
    rois = _rois_f32(rois)
    out = torch.empty((r, nout), dtype=torch.float32, device=feat.device)
    if r == 0:
        return out
    rs = _f32c(row_scale) if row_scale is not None else None
    b = _f32c(bias)
    L = lib()
    wsb = L.jtsm_pool_fc6_workspace_bytes(n, r, ph, pw)
    ws = _workspace(wsb, feat.device)
    with torch.cuda.device(feat.device):
        check(L.jtsm_pool_fc6_forward(_ptr(feat), _ptr(rois), n, c, h, w, r, float(spatial_scale), ph, pw, _ptr(rs), _ptr(W),
                                      _ptr(b), nout, int(bool(relu)), _ptr(out), _ptr(ws), wsb, _stream()),
              "pool_fc6_forward")
    return out


# --------------------------------------------------------------------------- ROIMerge / ROILabel
def roi_merge_lambda(cur_iter, size_epoch, max_epoch):
    """The merge threshold of an iteration: getlambda(cur_iter / size_epoch, max_epoch), ROIMerge_cpu.cpp:11-17, :102."""
    import math

    import numpy as np

    f = np.float32
    low = f(0.01)
    it = f(f(cur_iter) / f(size_epoch))
    a = math.log(float(f(it + low))) - math.log(float(low))
    b = math.log(float(f(f(max_epoch) + low))) - math.log(float(low))
    return float(f(a / b))


def roi_merge_forward(S, J, C, D, lam, top_k=200, window=40):
    """ROIMerge_forward (wsl/layers/roi_merge.py:13, ROIMerge_cpu.cpp:33-232) on the device, no host read.
    S [R] or [R,1], J [R,R], C / D [R,K] -> dict(MC, MD [R,K] padded with zero rows, I, IC int32 [R],
    stats int32 [4] = num_id, num_top_id, max_clique, min_clique)."""
    _require_cuda(S, J, C, D)
    Cc, Dc, Jc = _f32c(C), _f32c(D), _f32c(J)
    Sc = _f32c(S).reshape(-1)
    r, k = Cc.shape
    assert Sc.numel() == r and Jc.shape == (r, r) and Dc.shape == (r, k)
    dev = Cc.device
    out = dict(MC=torch.empty((r, k), dtype=torch.float32, device=dev), MD=torch.empty((r, k), dtype=torch.float32, device=dev),
               I=torch.empty(r, dtype=torch.int32, device=dev), IC=torch.empty(r, dtype=torch.int32, device=dev),
               stats=torch.empty(4, dtype=torch.int32, device=dev))
    L = lib()
    wsb = L.jtsm_roi_merge_workspace_bytes(r)
    ws = _workspace(wsb, dev)
    with torch.cuda.device(dev):
        check(L.jtsm_roi_merge_forward(_ptr(Sc), _ptr(Jc), _ptr(Cc), _ptr(Dc), r, k, float(lam), int(top_k), int(window),
                                       _ptr(out["MC"]), _ptr(out["MD"]), _ptr(out["I"]), _ptr(out["IC"]),
                                       _ptr(out["stats"]), _ptr(ws), wsb, _stream()), "roi_merge_forward")
    return out


def roi_merge_backward(GMC, GMD, I, IC, num_rois=None):
    """ROIMerge_backward (ROIMerge_cpu.cpp:234-287): GC[n] = GMC[I[n]] / IC[I[n]]."""
    _require_cuda(GMC, GMD, I, IC)
    gmc, gmd = _f32c(GMC), _f32c(GMD)
    r = I.numel() if num_rois is None else int(num_rois)
    k = gmc.shape[1]
    dev = gmc.device
    gc = torch.empty((r, k), dtype=torch.float32, device=dev)
    gd = torch.empty((r, k), dtype=torch.float32, device=dev)
    Ii, ICi = I.to(torch.int32).contiguous(), IC.to(torch.int32).contiguous()
    with torch.cuda.device(dev):
        check(lib().jtsm_roi_merge_backward(_ptr(gmc), _ptr(gmd), _ptr(Ii), _ptr(ICi), r, k, _ptr(gc), _ptr(gd), _stream()),
              "roi_merge_backward")
    return gc, gd


def roi_label_forward(S, U, L, CW, perm, fg_thresh, bg_thresh_hi, bg_thresh_lo, num_pos, num_neg, top_k=1):
    """ROILabel_forward (wsl/layers/roi_label.py:12, ROILabel_cpu.cpp:16-194) on the device with the visiting order as
    an input (int32 [R] permutation, None = index order).  S [R,Ks], U [R,R], L [K] or [1,K], CW [K] / [1,K] / None.
    -> (RL int32 [R], RW float32 [R], stats float32 [4] = fg rois, bg rois, fg weight, bg weight)."""
    _require_cuda(S, U, L)
    Sc, Uc = _f32c(S), _f32c(U)
    Lc = _f32c(L).reshape(-1)
    r, ks = Sc.shape
    k = Lc.numel()
    assert Uc.shape == (r, r)
    dev = Sc.device
    cw = _f32c(CW).reshape(-1) if CW is not None else None
    assert cw is None or cw.numel() >= k
    pm = perm.to(device=dev, dtype=torch.int32).contiguous() if perm is not None else None
    assert pm is None or pm.numel() == r
    rl = torch.empty(r, dtype=torch.int32, device=dev)
    rw = torch.empty(r, dtype=torch.float32, device=dev)
    stats = torch.empty(4, dtype=torch.float32, device=dev)
    Lb = lib()
    wsb = Lb.jtsm_roi_label_workspace_bytes(r)
    ws = _workspace(wsb, dev)
    with torch.cuda.device(dev):
        check(Lb.jtsm_roi_label_forward(_ptr(Sc), ks, _ptr(Uc), _ptr(Lc), _ptr(cw), _ptr(pm), r, k, float(fg_thresh),
                                        float(bg_thresh_hi), float(bg_thresh_lo), int(num_pos), int(num_neg), int(top_k),
                                        _ptr(rl), _ptr(rw), _ptr(stats), _ptr(ws), wsb, _stream()), "roi_label_forward")
    return rl, rw, stats


def pcl_proposal_clusters(rois, centers, center_labels, center_scores, cls_prob, fg_thresh=0.5, bg_thresh=0.1):
    """`_get_proposal_clusters` (third_party/pcl.py:150-200) on the device.  rois [R,4], centers [G,4], center_labels [G]
    (class column ids), center_scores [G], cls_prob [R,C] -> dict of fp32 tensors: labels, cls_loss_weights,
    gt_assignment [R]; pc_labels, pc_probs, pc_count, img_cls_loss_weights [G]."""
    _require_cuda(rois, centers, cls_prob)
    B, Cn, P = _f32c(rois), _f32c(centers), _f32c(cls_prob)
    r, g = B.shape[0], Cn.shape[0]
    dev = B.device
    cl = center_labels.to(device=dev, dtype=torch.float32).contiguous().view(-1)
    cs = center_scores.to(device=dev, dtype=torch.float32).contiguous().view(-1)
    assert cl.numel() == g and cs.numel() == g and P.shape[0] == r and g > 0
    per_roi = torch.empty((3, r), dtype=torch.float32, device=dev)
    per_c = torch.empty((4, g), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().jtsm_pcl_proposal_clusters(_ptr(B), _ptr(Cn), _ptr(cl), _ptr(cs), _ptr(P), r, g, P.shape[1],
                                               float(fg_thresh), float(bg_thresh), _ptr(per_roi[0]), _ptr(per_roi[1]),
                                               _ptr(per_roi[2]), _ptr(per_c[0]), _ptr(per_c[1]), _ptr(per_c[2]),
                                               _ptr(per_c[3]), _stream()), "pcl_proposal_clusters")
    return dict(labels=per_roi[0], cls_loss_weights=per_roi[1], gt_assignment=per_roi[2], pc_labels=per_c[0],
                pc_probs=per_c[1], pc_count=per_c[2], img_cls_loss_weights=per_c[3])
