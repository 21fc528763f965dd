"""Drop-in for the reference's utils/retrieval.py (compute_mAP :23-43, retrieval_dist :46-79,
retrieval_eval :82-126, get_rank :129-136, scan2cad_retrieval_eval_dist :139-167,
scan2cad_retrieval_eval :170-177).

The distance computation and ranking (scipy cdist / pdist + argsort in the reference) run on the
device: exact float64 distances + per-row radix select (cb_retrieval_topk), ties by lower index.
Set intersections and sklearn's average precision stay on the host, as thin as in the reference.
"""
import numpy as np
import torch

from . import _capi
from ._capi import ptr

_METRIC = {"euclidean": 0, "cosine": 1}


def topk_device(q, lib, k, metric="euclidean", exclude_self=0, self_offset=0, ctx=None):
    """k best library rows per query: (idx [Q,k] int32, dist [Q,k] float64) CUDA tensors.
    exclude_self: 0 off, 1 drop column (row + self_offset), 2 force that column's distance to 0
    (scipy squareform's exact-zero diagonal)."""
    ctx = ctx or _capi.get_context()
    qd = ctx.dev(q, torch.float32)
    ld = ctx.dev(lib, torch.float32)
    if qd.dim() != 2 or ld.dim() != 2 or qd.shape[1] != ld.shape[1]:
        raise ValueError("descriptors must be [N,D] with equal D")
    Q, d = qd.shape
    idx = ctx.empty((Q, k), torch.int32)
    dist = ctx.empty((Q, k), torch.float64)
    ctx.call("cb_retrieval_topk", ptr(qd), Q, ptr(ld), ld.shape[0], d, int(k), _METRIC[metric],
             int(exclude_self), int(self_offset), ptr(idx), ptr(dist), ctx.stream)
    return idx, dist


def topk(q, lib, k, metric="euclidean", exclude_self=0):
    ctx = _capi.get_context()
    idx, dist = topk_device(q, lib, k, metric, exclude_self, ctx=ctx)
    return ctx.host(idx).astype(np.int64), ctx.host(dist)


def _cosine_rows(descriptors, k):
    """Sorted cosine neighbours with the diagonal forced to exactly 0 (squareform(pdist))."""
    return topk(descriptors, descriptors, k, "cosine", exclude_self=2)


def _average_precision(positive, score):
    from sklearn.metrics import average_precision_score
    return average_precision_score(positive, score)


def compute_mAP(descriptors, gt_labels):
    """utils/retrieval.py:23-43."""
    n = len(descriptors)
    idx, dist = _cosine_rows(descriptors, n)
    dists = np.empty((n, n))
    np.put_along_axis(dists, idx, dist, 1)
    dists = 2 - dists
    instance_scores, class_scores = [], {}
    for dist, gt_label in zip(dists, gt_labels):
        score = 100.0 * _average_precision(gt_labels == gt_label, dist)
        instance_scores.append(score)
        class_scores.setdefault(gt_label, []).append(score)
    return np.mean(np.array(instance_scores)), [np.mean(class_scores[key]) for key in class_scores.keys()]


def retrieval_dist(dists, threshold, table):
    """utils/retrieval.py:46-79 (takes a precomputed host distance matrix)."""
    scores, percisions, top1_errors = [], [], []
    n_acc = int(len(dists) * threshold)
    rank_predicts = np.argsort(dists, axis=1)
    rank_gts = np.argsort(table, axis=1)
    for dist, rank_predict, rank_gt in zip(dists, rank_predicts, rank_gts):
        rank_predict = rank_predict[1:n_acc + 1]
        rank_gt = rank_gt[:n_acc]
        positive = np.isin(rank_predict, rank_gt).astype(np.int32)
        if table[rank_predict[0], rank_gt[0]] == 200:
            top1_errors.append(0)
        else:
            top1_errors.append(table[rank_predict[0], rank_gt[0]])
        percisions.append(100.0 * positive.sum() / n_acc)
        if positive.sum() == 0:
            scores.append(0)
        else:
            scores.append(100.0 * _average_precision(positive, 2 - dist[rank_predict]))
    return {"mAP": np.mean(scores), "percision": np.mean(percisions), "top1_error": np.mean(top1_errors)}


def retrieval_eval(descriptors, threshold, table):
    """utils/retrieval.py:82-126: cosine ranking on the device, metrics as in the reference."""
    n = len(descriptors)
    n_acc = int(n * threshold)
    idx, dist = _cosine_rows(descriptors, n_acc + 1)
    rank_gts = np.argsort(table, axis=1)
    scores, percisions, top1_errors = [], [], []
    for rank_predict, d, rank_gt in zip(idx, dist, rank_gts):
        rank_predict, d = rank_predict[1:n_acc + 1], d[1:n_acc + 1]
        rank_gt = rank_gt[:n_acc]
        positive = np.isin(rank_predict, rank_gt).astype(np.int32)
        if table[rank_predict[0], rank_gt[0]] == 200:
            top1_errors.append(0)
        else:
            top1_errors.append(table[rank_predict[0], rank_gt[0]])
        percisions.append(100.0 * positive.sum() / n_acc)
        if positive.sum() == 0:
            scores.append(0)
        else:
            scores.append(100.0 * _average_precision(positive, 2 - d))
    return {"mAP": np.mean(scores), "percision": np.mean(percisions), "top1_error": np.mean(top1_errors)}


def get_rank(descriptors, top_n=10):
    """utils/retrieval.py:129-136 -> [n, top_n] int64 (first column, the query itself, dropped)."""
    idx, _ = _cosine_rows(descriptors, top_n + 1)
    return idx[:, 1:top_n + 1]


def _scan2cad_stats(rank_predict, table, best_match, pos_n):
    rank_gt = np.argsort(table[best_match, :], axis=1, kind="stable")
    precision, top1_error, top1_predict, gt = [], [], [], []
    for g, p in zip(rank_gt, rank_predict):
        positive = np.isin(p[:pos_n], g[:pos_n]).astype(np.int32)
        precision.append(100.0 * positive.sum() / pos_n)
        top1_error.append(table[p[0], g[0]])
        top1_predict.append(p[0])
        gt.append(g[0])
    return {"precision": sum(precision) / len(precision), "top1_error": sum(top1_error) / len(top1_error),
            "top1_predict": top1_predict, "gt": gt}


def scan2cad_retrieval_eval_dist(dists, table, best_match, pos_n):
    """utils/retrieval.py:139-167 (takes a precomputed host distance matrix)."""
    return _scan2cad_stats(np.argsort(dists, axis=1, kind="stable"), table, best_match, pos_n)


def scan2cad_retrieval_eval(scan_feats, lib_feats, best_match, table, pos_n):
    """utils/retrieval.py:170-177: Euclidean ranking of the scans against the CAD library."""
    idx, _ = topk(scan_feats, lib_feats, max(1, int(pos_n)))
    return _scan2cad_stats(idx, table, best_match, pos_n)
