"""Restatement of utils/symmetry.py's hot functions.  TEST INFRASTRUCTURE (see oracle/port.py)."""
import numpy as np
import torch
from sklearn.cluster import KMeans

from . import port


def split_corr(pcsA, pcsB, featsA, featsB, knn, subsample_size=1000, find_kcorr=None):
    """utils/symmetry.py:145-179."""
    fk = find_kcorr or port.find_kcorr
    A, B = [], []
    for pcA, pcB, fA, fB in zip(pcsA, pcsB, featsA, featsB):
        i0, i1 = fk(fA, fB, k=knn, subsample_size=subsample_size)
        A.append(pcA[i0])
        B.append(pcB[i1])
    return np.concatenate(A, 0), np.concatenate(B, 0)


def cut_candidates(feat, raw_pc, K, max_sample=100, n_nn=50):
    """The per-sample quantities of symmetric_cut4 (utils/symmetry.py:193-236), exposed so
    tests can compare stage by stage.  Consumes the global NumPy RNG exactly once like the
    reference (:193).  Returns list of dicts (index, nn_ids, centers, min_center_dist,
    ratios, errors, accepted) and the winning position (or -1)."""
    picks = np.random.choice(len(raw_pc), max_sample, replace=False)
    out, cur, win = [], 100, -1
    for pos, index in enumerate(picks):
        ld = np.linalg.norm(feat[index:index + 1, :] - feat, axis=1)
        srt = np.argsort(ld)
        rank = np.zeros(len(srt))
        rank[srt] = np.arange(len(srt))
        sel = rank < n_nn
        nns = raw_pc[sel, :]
        km = KMeans(n_clusters=K, random_state=0, n_init=10).fit(nns)
        c = km.cluster_centers_
        dist = np.linalg.norm(c[None, :, :] - c[:, None, :], 2, 2)
        dist[np.arange(K), np.arange(K)] = 100
        labels = km.predict(raw_pc)
        ratios = [(labels == i).sum() / len(labels) for i in range(K)]
        nl = km.predict(nns)
        err = [np.linalg.norm(nns[nl == l] - c[l], axis=1).mean() for l in range(K)]
        sd = np.sqrt(np.var(ratios))
        acc = bool(dist.min() > 0.15 > max(err) and cur > sd)
        if acc:
            cur, win = sd, pos
        out.append(dict(index=int(index), nn_ids=np.nonzero(sel)[0], centers=c, labels=labels,
                        min_center_dist=dist.min(), ratios=ratios, errors=err, accepted=acc))
    return out, win


def masks_from_model(centers, labels, K):
    """utils/symmetry.py:238-259: K=4 order is [0, nearest, farthest, middle] of centre 0."""
    dist = np.linalg.norm(centers[None, :, :] - centers[:, None, :], 2, 2)
    dist[np.arange(K), np.arange(K)] = 100
    rc = np.argsort(dist[0, 1:])
    if K == 2:
        return [labels == 0, labels == 1]
    if K == 4:
        return [labels == 0, labels == rc[0] + 1, labels == rc[2] + 1, labels == rc[1] + 1]
    raise ValueError("not defined")


def symmetric_cut4(feat, raw_pc, K, max_sample=100):
    """utils/symmetry.py:182-259.  No accepted model -> AttributeError like the reference
    (``None.cluster_centers_`` at :238)."""
    if isinstance(raw_pc, torch.Tensor):
        raw_pc = raw_pc.detach().cpu()
    cands, win = cut_candidates(feat, raw_pc, K, max_sample)
    if win < 0:
        raise AttributeError("'NoneType' object has no attribute 'cluster_centers_'")
    return masks_from_model(cands[win]["centers"], cands[win]["labels"], K)


def sym_pose(baseF, xyz0, posF, xyz1, pos_sym, k_nn=5, max_corr=0.20, seed=0, *,
             registration=None, cut=None, find_kcorr=None):
    """utils/symmetry.py:262-358.  ``registration(src, tgt, max_corr, seed)`` defaults to the
    restated RANSAC; hooks exist so tests can share minimal sets / masks with the CUDA path."""
    reg = registration or port.registration_based_on_corr
    cutf = cut or symmetric_cut4
    fk = find_kcorr or port.find_kcorr
    i0, i1 = fk(baseF, posF, k=k_nn, subsample_size=-1)
    T_r = torch.from_numpy(np.asarray(reg(xyz0[i0], xyz1[i1], max_corr, seed)).astype(np.float32))
    ch_r = port.chamfer_kdtree_1direction(port.apply_transform(xyz0, T_r), xyz1)
    ch_b, T_b = ch_r, T_r
    try:
        K = 4 if pos_sym >= 2 else 2
        bm = cutf(baseF, xyz0, K, max_sample=100)
        pm = cutf(posF, xyz1, K, max_sample=100)
    except Exception:
        return T_b, ch_b, T_r, ch_r, False

    def sweep(bm, pm, T_b, ch_b):
        for _ in range(len(bm)):
            A, B = split_corr([xyz0[m] for m in bm], [xyz1[m] for m in pm],
                              [baseF[m] for m in bm], [posF[m] for m in pm], k_nn,
                              subsample_size=-1, find_kcorr=fk)
            T = torch.from_numpy(np.asarray(reg(A, B, max_corr, seed)).astype(np.float32))
            ch = port.chamfer_kdtree_1direction(port.apply_transform(xyz0, T), xyz1)
            pm.append(pm.pop(0))
            if ch_b > ch:
                ch_b, T_b = ch, T
        return T_b, ch_b

    T_b, ch_b = sweep(bm, pm, T_b, ch_b)
    if pos_sym >= 2:
        pm = [pm[0], pm[3], pm[2], pm[1]]
        T_b, ch_b = sweep(bm, pm, T_b, ch_b)
    return T_b, ch_b, T_r, ch_r, True
