"""Restatement of utils/retrieval.py, utils/pc_dist.py and utils/_transform_estimation.py.
TEST INFRASTRUCTURE (see oracle/port.py)."""
import numpy as np
import torch
from scipy.spatial.distance import cdist, pdist as sp_pdist, squareform

from . import port

# ------------------------------------------------------------------------- retrieval.py


def stable_argsort_rows(d):
    """np.argsort(d, 1) with ties made canonical (lower index first)."""
    return np.argsort(d, axis=1, kind="stable")


def get_rank(descriptors, top_n=10):
    """utils/retrieval.py:129-136 (cosine; self excluded by position)."""
    d = squareform(sp_pdist(descriptors, "cosine"))
    return stable_argsort_rows(d)[:, 1:top_n + 1]


def scan2cad_retrieval_eval_dist(dists, table, best_match, pos_n):
    """utils/retrieval.py:139-167."""
    pred = stable_argsort_rows(dists)
    gt = stable_argsort_rows(table[best_match, :])
    prec, err, p0, g0 = [], [], [], []
    for g, p in zip(gt, pred):
        prec.append(100.0 * np.isin(p[:pos_n], g[:pos_n]).sum() / pos_n)
        err.append(table[p[0], g[0]])
        p0.append(p[0])
        g0.append(g[0])
    return {"precision": sum(prec) / len(prec), "top1_error": sum(err) / len(err),
            "top1_predict": p0, "gt": g0}


def scan2cad_retrieval_eval(scan_feats, lib_feats, best_match, table, pos_n):
    """utils/retrieval.py:170-177."""
    return scan2cad_retrieval_eval_dist(cdist(scan_feats, lib_feats), table, best_match, pos_n)


def _ap(positive, score):
    from sklearn.metrics import average_precision_score
    return average_precision_score(positive, score)


def retrieval_dist(dists, threshold, table):
    """utils/retrieval.py:46-79."""
    n_acc = int(len(dists) * threshold)
    rp, rg = np.argsort(dists, axis=1), np.argsort(table, axis=1)
    sc, pr, te = [], [], []
    for d, p, g in zip(dists, rp, rg):
        p, g = p[1:n_acc + 1], g[:n_acc]
        pos = np.isin(p, g).astype(np.int32)
        te.append(0 if table[p[0], g[0]] == 200 else table[p[0], g[0]])
        pr.append(100.0 * pos.sum() / n_acc)
        sc.append(0 if pos.sum() == 0 else 100.0 * _ap(pos, 2 - d[p]))
    return {"mAP": np.mean(sc), "percision": np.mean(pr), "top1_error": np.mean(te)}


def retrieval_eval(descriptors, threshold, table):
    """utils/retrieval.py:82-126."""
    return retrieval_dist(squareform(sp_pdist(descriptors, "cosine")), threshold, table)


def compute_mAP(descriptors, gt_labels):
    """utils/retrieval.py:23-43."""
    d = squareform(sp_pdist(descriptors, "cosine"))
    inst, cls = [], {}
    for row, c in zip(d, gt_labels):
        s = 100.0 * _ap(gt_labels == c, 2 - row)
        inst.append(s)
        cls.setdefault(c, []).append(s)
    return np.mean(np.array(inst)), [np.mean(v) for v in cls.values()]


# --------------------------------------------------------------------------- pc_dist.py


def compute_dist(pcs):
    """utils/pc_dist.py:61-99 without the thread fan-out: upper triangle, += T (diagonal 200)."""
    L = len(pcs)
    table = np.eye(L) * 100
    for i in range(L):
        for j in range(i + 1, L):
            table[i, j] = port.chamfer(pcs[i], pcs[j])
    table += table.T
    return table


# ------------------------------------------------------------- _transform_estimation.py


def rot_x(x):
    c, s = torch.cos(x), torch.sin(x)
    o = torch.zeros((3, 3))
    o[0, 0], o[1, 1], o[1, 2], o[2, 1], o[2, 2] = 1, c, -s, s, c
    return o


def rot_y(x):
    c, s = torch.cos(x), torch.sin(x)
    o = torch.zeros((3, 3))
    o[0, 0], o[0, 2], o[1, 1], o[2, 0], o[2, 2] = c, s, 1, -s, c
    return o


def rot_z(x):
    c, s = torch.cos(x), torch.sin(x)
    o = torch.zeros((3, 3))
    o[0, 0], o[0, 1], o[1, 0], o[1, 1], o[2, 2] = c, -s, s, c, 1
    return o


def get_trans(x):
    """utils/_transform_estimation.py:57-61."""
    T = torch.eye(4)
    T[:3, :3] = rot_z(x[2]).mm(rot_y(x[1])).mm(rot_x(x[0]))
    T[:3, 3] = x[3:, 0]
    return T


def update_pcd(pts, trans):
    """:64-71."""
    trans = trans.float()
    return torch.t(trans[:3, :3] @ torch.t(pts.float())) + trans[:3, 3]


def build_linear_system(pts0, pts1, weight):
    """:74-95 -> A [3N,6], b [3N,1] (x rows, then y rows, then z rows)."""
    n = pts0.shape[0]
    x, y, z = pts0[:, 0], pts0[:, 1], pts0[:, 2]
    zero, one = torch.zeros(n), torch.ones(n)
    A = torch.cat((torch.stack((zero, z, -y, one, zero, zero), 1),
                   torch.stack((-z, zero, x, zero, one, zero), 1),
                   torch.stack((y, -x, zero, zero, zero, one), 1)), 0)
    b = torch.cat((pts1[:, 0] - x, pts1[:, 1] - y, pts1[:, 2] - z), 0).unsqueeze(1)
    return weight.repeat(3, 6) * A, weight.repeat(3, 1) * b


def solve_linear_system(A, b):
    """:98-103."""
    A, b = A.float(), b.float()
    return torch.inverse(A.t().mm(A)).mm(A.t()).mm(b)


def compute_weights(pts0, pts1, par):
    """:106-107."""
    return par / (torch.norm(pts0 - pts1, dim=1).unsqueeze(1) + par)


def est_quad_linear_robust(pts0, pts1, weight=None):
    """:110-137."""
    cur, trans, par = pts0, torch.eye(4), 1.0
    if weight is None:
        weight = torch.ones(pts0.size()[0], 1)
    for i in range(20):
        if i > 0 and i % 5 == 0:
            par /= 2.0
        A, b = build_linear_system(cur, pts1, weight)
        tc = get_trans(solve_linear_system(A, b))
        cur = update_pcd(cur, tc)
        weight = compute_weights(cur, pts1, par)
        trans = tc.mm(trans)
    return trans


def pose_from_features(xyz0, xyz1, F0, F1, return_corr=False):
    """utils/_transform_estimation.py:149-159 after the two model forwards (:143-147)."""
    corr = F0.mm(F1.t())
    weight, inds = corr.max(dim=1)
    weight = weight.unsqueeze(1).cpu()
    trans = est_quad_linear_robust(xyz0, xyz1[inds, :], weight)
    return (trans, weight, corr) if return_corr else (trans, weight)
