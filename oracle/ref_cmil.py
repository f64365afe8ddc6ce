"""CPU restatement of the reference's ROIMerge and ROILabel ops (CMIL / UWSOD heads).

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing under jtsm_b200/).  Pinned bit for bit to the reference's own
C++ (oracle/_ref/libref_wslcmil.so = ROIMerge_cpu.cpp + ROILabel_cpu.cpp compiled in place) by tests/test_ref_cmil.py.

Deterministic specification where the reference is not:
 * ROIMerge sorts the scores with std::sort (unstable; ROIMerge_cpu.cpp:20-31): equal scores are visited here lowest
   index first.  The pin uses distinct scores.
 * ROILabel visits the RoIs in std::random_shuffle order seeded by the wall clock (ROILabel_cpu.cpp:108-117): the
   visiting order is an INPUT here (`perm`); the pin feeds the permutation the reference build draws.
"""
import math

import numpy as np

F32 = np.float32


def merge_lambda(cur_iter, size_epoch, max_epoch):
    """getlambda(float(cur_iter) / float(size_epoch), float(max_epoch)), ROIMerge_cpu.cpp:11-17 (+ :102-103).
    float operands; `log` resolves to the double overload (no using-directive in that file), result stored as float."""
    low = F32(0.01)
    it = F32(F32(cur_iter) / F32(size_epoch))
    a = math.log(float(F32(it + low))) - math.log(float(low))
    b = math.log(float(F32(F32(max_epoch) + low))) - math.log(float(low))
    return F32(a / b)


def roi_merge_forward(S, J, C, D, lam, top_k=200, window=40):
    """ROIMerge_forward_cpu (ROIMerge_cpu.cpp:33-232) for a given lambda.
    S [R] scores, J [R,R] IoU, C / D [R,K].  -> dict(MC [num_id,K], MD, I int32 [R], IC int32 [R], num_top_id,
    max_clique, min_clique)."""
    S = np.asarray(S, F32).reshape(-1)
    J, C, D = np.asarray(J, F32), np.asarray(C, F32), np.asarray(D, F32)
    lam = F32(lam)
    r, k = C.shape
    order = np.argsort(-S.astype(np.float64), kind="stable")  # descending, ties: lowest index first
    I = np.full(r, -1, np.int32)
    top = min(top_k, r)
    cur = 0
    for t in range(top):  # :106-143 greedy cliques inside windows of 40 ranks
        n = order[t]
        if I[n] != -1:
            continue
        I[n] = cur
        end = min(t + window, top)
        for tt in range(t, end):
            i = order[tt]
            if I[i] != -1:
                continue
            ok = True
            for ttt in range(t, end):
                j = order[ttt]
                if I[j] == cur and J[i, j] < lam:
                    ok = False
                    break
            if ok:
                I[i] = cur
        cur += 1
    num_top = cur
    for n in range(r):  # :149-158 everything else is its own cluster, in index order
        if I[n] == -1:
            I[n] = cur
            cur += 1
    num_id = cur
    IC = np.zeros(r, np.int32)
    for n in range(r):
        IC[I[n]] += 1
    MC, MD = np.zeros((num_id, k), F32), np.zeros((num_id, k), F32)
    for n in range(r):  # :198-205 running sums of C / count in index order
        d = F32(IC[I[n]])
        MC[I[n]] = (MC[I[n]] + C[n] / d).astype(F32)
        MD[I[n]] = (MD[I[n]] + D[n] / d).astype(F32)
    top_counts = IC[:num_top]
    return dict(MC=MC, MD=MD, I=I, IC=IC, num_top_id=num_top, max_clique=int(max([0] + list(top_counts))),
                min_clique=int(min([top] + list(top_counts))))


def roi_merge_backward(GMC, GMD, I, IC):
    """ROIMerge_backward_cpu (:234-287): every member receives its cluster's gradient / count."""
    I, IC = np.asarray(I), np.asarray(IC)
    d = IC[I].astype(F32)[:, None]
    return (np.asarray(GMC, F32)[I] / d).astype(F32), (np.asarray(GMD, F32)[I] / d).astype(F32)


def roi_label_forward(S, U, L, CW, perm, fg_thresh, bg_thresh_hi, bg_thresh_lo, num_pos, num_neg, top_k):
    """ROILabel_forward_cpu (ROILabel_cpu.cpp:16-194) with the visiting order `perm` given.
    S [R,Ks], U [R,R] IoU, L [K] image labels (== 1 switches a class on), CW [K] class weights or None.
    -> dict(RL int32 [R], RW float32 [R], fg_rois, bg_rois, fg_weight, bg_weight)."""
    S, U = np.asarray(S, F32), np.asarray(U, F32)
    L = np.asarray(L, F32).reshape(-1)
    r, ks = S.shape
    k = L.shape[0]
    hn, hc, hp = [], [], []
    for c in range(k):  # :84-106 seeds: per image class the top_k best rows not chosen before (first maximum)
        if L[c] != 1:
            continue
        for _ in range(top_k):
            best, idx = -np.finfo(F32).max, -1
            for n in range(r):
                if best < S[n, c] and n not in hn:
                    best, idx = S[n, c], n
            hn.append(idx)
            hc.append(c)
            hp.append(F32(best))
    RL, RW = np.zeros(r, np.int32), np.zeros(r, F32)
    npos = nneg = 0
    fg_w = bg_w = F32(0)
    fg_n = bg_n = 0
    fg_thresh, bg_thresh_hi, bg_thresh_lo = F32(fg_thresh), F32(bg_thresh_hi), F32(bg_thresh_lo)
    for n in perm:  # :122-165
        iou = U[n, hn]
        mi = int(np.argmax(iou))  # first maximum (strict <)
        m = iou[mi]
        c = hc[mi]
        w = F32(CW[c]) if CW is not None else hp[mi]
        if m >= fg_thresh and npos <= num_pos:
            npos += 1
            fg_n += 1
            fg_w = F32(fg_w + w)
        elif m >= bg_thresh_lo and m < bg_thresh_hi and nneg <= num_neg:
            c = k
            nneg += 1
            bg_n += 1
            bg_w = F32(bg_w + w)
        else:
            w = F32(0)
        RL[n], RW[n] = c, w
    return dict(RL=RL, RW=RW, fg_rois=fg_n, bg_rois=bg_n, fg_weight=fg_w, bg_weight=bg_w, seeds=hn, seed_classes=hc)
