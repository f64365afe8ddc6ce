"""PCL targets: drop-in for `wsl.modeling.roi_heads.third_party.pcl.PCL` (third_party/pcl.py:26-58).

`PCL(boxes, cls_prob, im_labels, cls_prob_new)` -> the dict `pcl_loss` consumes (same keys, `[1, n]` fp32), as DEVICE
tensors.  Two halves:
 * cluster CENTRES (`_get_graph_centers`, pcl.py:85-147): per image class a 1-D k-means over the class scores
   (sklearn `KMeans(3, random_state=3)`, the reference's own third-party dependency) keeps the high-score cluster, a
   greedy cover of the IoU > 0.4 graph picks centres, the best five per class survive.  A few hundred boxes, RNG-
   dependent third-party code: it stays on the host like the reference's (the IoU graph itself comes from the device
   kernel, bit-equal to `pairwise_iou`).
 * proposal clusters (`_get_proposal_clusters`, pcl.py:150-200): the per-proposal part (R x G IoU, first-maximum
   assignment, thresholds, per-cluster sums) runs on the device: `ops.pcl_proposal_clusters`."""
import numpy as np
import torch

from . import ops

NUM_KMEANS_CLUSTER = 3
RNG_SEED = 3
GRAPH_IOU_THRESHOLD = 0.4
MAX_PC_NUM = 5
FG_THRESH = 0.5
BG_THRESH = 0.1


def _high_score_members(scores):
    """Indices of the k-means cluster with the highest centre (pcl.py:61-72)."""
    from sklearn.cluster import KMeans

    km = KMeans(n_clusters=min(NUM_KMEANS_CLUSTER, scores.shape[0]), random_state=RNG_SEED).fit(scores)
    members = np.where(km.labels_ == np.argmax(km.cluster_centers_))[0]
    return members if len(members) else np.array([np.argmax(scores)])


def _device_iou(boxes):
    if not torch.cuda.is_available():
        raise RuntimeError("jtsm_b200.pcl: CUDA device required (no CPU fallback)")
    t = torch.from_numpy(np.ascontiguousarray(boxes, dtype=np.float32)).cuda()
    return ops.pairwise_iou(t, t).cpu().numpy()


def graph_centers(boxes, cls_prob, im_labels, iou_fn=None):
    """Host half.  boxes [R,4] fp32 numpy, cls_prob [R,K] (clipped), im_labels [1,K] -> (boxes [G,4], classes [G] int32
    1-based, scores [G]).  iou_fn(boxes [n,4]) -> [n,n] numpy; default: the device kernel (tests inject a checker)."""
    iou_fn = iou_fn or _device_iou
    assert im_labels.shape[0] == 1, "batch size shoud be equal to 1"
    boxes, cls_prob = boxes.copy(), cls_prob.copy()
    out_b, out_c, out_s = [np.zeros((0, 4), np.float32)], [np.zeros((0,), np.int32)], [np.zeros((0,), np.float32)]
    for k in np.flatnonzero(im_labels[0] == 1):
        col = cls_prob[:, k].copy()
        pool = np.where(col >= 0)[0]
        pool = pool[_high_score_members(col[pool].reshape(-1, 1))]
        cand_boxes, cand_p = boxes[pool], col[pool]
        graph = (iou_fn(cand_boxes) > GRAPH_IOU_THRESHOLD).astype(np.float32)
        picked, picked_score, left = [], [], cand_p.size
        while True:  # greedy cover: the node of highest degree takes its neighbourhood
            node = graph.sum(axis=1).argsort()[::-1][0]
            nb = np.where(graph[node] > 0)[0]
            picked.append(node)
            picked_score.append(np.max(cand_p[nb]))
            graph[:, nb] = 0
            graph[nb, :] = 0
            left -= len(nb)
            if left <= 5:
                break
        picked_score = np.array(picked_score)
        best = np.argsort(picked_score)[-1:(-1 - min(len(picked_score), MAX_PC_NUM)):-1]
        chosen = pool[np.array(picked)[best]]
        out_b.append(boxes[chosen])
        out_s.append(picked_score[best].astype(np.float32))
        out_c.append(np.full(len(best), k + 1, np.int32))
        cls_prob = np.delete(cls_prob, chosen, axis=0)  # a chosen centre leaves the candidate pool (pcl.py:138-142)
        boxes = np.delete(boxes, chosen, axis=0)
    return np.concatenate(out_b), np.concatenate(out_c), np.concatenate(out_s)


def PCL(boxes, cls_prob, im_labels, cls_prob_new):
    """boxes: numpy [R,4] or tensor; cls_prob / cls_prob_new: tensors [R, K or K+1] / [R, K+1]; im_labels numpy [1,K]."""
    dev = cls_prob_new.device
    eps = 1e-9
    p = cls_prob.detach().float().cpu().numpy()
    if p.shape[1] != im_labels.shape[1]:
        p = p[:, 1:]
    p = np.clip(p, eps, 1 - eps)
    b = boxes.detach().float().cpu().numpy() if torch.is_tensor(boxes) else np.asarray(boxes, np.float32)
    cb, cc, cs = graph_centers(b, p, np.asarray(im_labels))
    pn = cls_prob_new.detach().float().clamp(eps, 1 - eps)
    t = ops.pcl_proposal_clusters(torch.from_numpy(b).to(dev), torch.from_numpy(cb).to(dev),
                                  torch.from_numpy(cc.astype(np.float32)).to(dev), torch.from_numpy(cs).to(dev), pn,
                                  FG_THRESH, BG_THRESH)
    out = {k: v.reshape(1, -1) for k, v in t.items()}
    real = torch.cat([torch.ones(1, 1), torch.as_tensor(np.asarray(im_labels), dtype=torch.float32)], dim=1)
    out["im_labels_real"] = real.to(dev)
    return out
