"""NumPy/scipy/sklearn restatement of the reference's hot-path Python.  TEST INFRASTRUCTURE.

Every function cites the reference lines it follows (paths relative to /root/reference).
It travels to the GPU box (the reference does not) and is itself pinned against the
reference's own code by tests/test_oracle_vs_reference.py (run where /root/reference exists)
and against tests/golden/*.npz.  The single call the reference delegates to Open3D is
restated in oracle/ransac_np.py (parity unpinned, see there).
"""
import numpy as np
import torch
from scipy.spatial import KDTree
from scipy.spatial.distance import cdist, pdist as sp_pdist, squareform

from . import ransac_np

# --------------------------------------------------------------------------- find_nn.py


def pdist(A, B, dist_type="L2"):
    """utils/find_nn.py:24-31 (difference form; L2 adds 1e-7 under the root)."""
    d2 = ((A[:, None, :] - B[None, :, :]) ** 2).sum(2)
    if dist_type == "L2":
        return torch.sqrt(d2 + 1e-7)
    if dist_type == "SquareL2":
        return d2
    raise NotImplementedError("Not implemented")


def find_knn_exact(feat0, feat1, k):
    """Canonical kNN: float64 difference-form distances, ties by (distance, index).

    Same neighbours as find_knn_cpu (utils/find_nn.py:43-49, scipy KDTree in float64) with
    scipy's unspecified tie order made canonical.  Returns (idx [N0,k] int64, d2 [N0,k] f64).
    """
    f0 = np.asarray(feat0, np.float64)
    f1 = np.asarray(feat1, np.float64)
    idx = np.empty((len(f0), k), np.int64)
    d2o = np.empty((len(f0), k), np.float64)
    step = max(1, (1 << 24) // max(1, len(f1)))
    n1 = len(f1)
    kk = min(n1, k + 8)
    for s in range(0, len(f0), step):
        blk = f0[s:s + step]
        d2 = np.zeros((len(blk), n1))
        for c in range(f0.shape[1]):
            df = blk[:, c:c + 1] - f1[None, :, c]
            d2 += df * df
        if kk < n1 and n1 > 4096:
            # large problems: order only the kk smallest of every row, and fall back to the full
            # ordering for rows whose k-th value ties with the cut (same result, far less sorting)
            part = np.argpartition(d2, kk - 1, axis=1)[:, :kk]
            pv = np.take_along_axis(d2, part, 1)
            o = np.lexsort((part, pv), axis=1)
            part, pv = np.take_along_axis(part, o, 1), np.take_along_axis(pv, o, 1)
            order = part[:, :k]
            for row in np.nonzero(pv[:, k - 1] >= pv[:, kk - 1])[0]:
                order[row] = np.lexsort((np.arange(n1), d2[row]))[:k]
        else:
            order = np.lexsort((np.broadcast_to(np.arange(n1), d2.shape), d2), axis=1)[:, :k]
        idx[s:s + step] = order
        d2o[s:s + step] = np.take_along_axis(d2, order, 1)
    return idx, d2o


def find_nn_cpu(feat0, feat1, return_distance=False):
    """utils/find_nn.py:34-40."""
    dists, inds = KDTree(feat1).query(feat0, k=1, workers=-1)
    return (inds, dists) if return_distance else inds


def find_knn_cpu(feat0, feat1, k, return_distance=False):
    """utils/find_nn.py:43-49."""
    dists, inds = KDTree(feat1).query(feat0, k=k, workers=-1)
    return (inds, dists) if return_distance else inds


def _chunks(n, step, floor_mode):
    """Row chunking of find_nn_gpu (ceil, :55-56) / find_knn_gpu (floor + tail, :90-108)."""
    if step <= 1:
        return [(0, n)]
    c = n // step if floor_mode else -(-n // step)
    out = [(i * step, min(n, (i + 1) * step)) for i in range(c)]
    if c * step < n:
        out.append((c * step, n))
    return out


def find_nn_gpu(F0, F1, nn_max_n=-1, return_distance=False, dist_type="SquareL2"):
    """utils/find_nn.py:52-82 -> inds [N0] int64 CPU, dists [N0,1] CPU."""
    ds, ix = [], []
    for a, b in _chunks(len(F0), nn_max_n, False):
        m, i = pdist(F0[a:b], F1, dist_type).min(dim=1)
        ds.append(m.detach().unsqueeze(1).cpu())
        ix.append(i.cpu())
    ds, ix = torch.cat(ds), torch.cat(ix)
    return (ix, ds) if return_distance else ix


def find_knn_gpu(F0, F1, k, nn_max_n=-1, return_distance=False, dist_type="SquareL2"):
    """utils/find_nn.py:86-119 -> inds [N0*k] int64 CPU, dists [N0*k,1] CPU."""
    ds, ix = [], []
    for a, b in _chunks(len(F0), nn_max_n, True):
        m, i = pdist(F0[a:b], F1, dist_type).topk(k=k, dim=1, largest=False)
        ds.append(m.reshape(-1).detach().unsqueeze(1).cpu())
        ix.append(i.reshape(-1).cpu())
    ds, ix = torch.cat(ds), torch.cat(ix)
    assert len(ix) == len(F0) * k
    return (ix, ds) if return_distance else ix


# ------------------------------------------------------------------------- eval_pose.py


def find_kcorr(F0, F1, k=1, nn_max_n=500, subsample_size=-1):
    """utils/eval_pose.py:48-79 (global np.random for the optional subsample)."""
    sub = subsample_size > 0 and len(F0) > subsample_size
    if sub:
        i0 = np.random.choice(len(F0), min(len(F0), subsample_size), replace=False)
        i1 = np.random.choice(len(F1), min(len(F1), subsample_size), replace=False)
        F0, F1 = F0[i0], F1[i1]
    nn = (find_nn_cpu(F0, F1) if k == 1 else find_knn_cpu(F0, F1, k)).flatten()
    if sub:
        return np.repeat(i0, k), i1[nn]
    return np.repeat(np.arange(len(F0)), k), nn


def find_corr(xyz0, xyz1, F0, F1, nn_max_n, subsample_size=-1):
    """utils/eval_pose.py:28-45."""
    sub = subsample_size > 0 and len(F0) > subsample_size
    if sub:
        i0 = np.random.choice(len(F0), min(len(F0), subsample_size), replace=False)
        i1 = np.random.choice(len(F1), min(len(F1), subsample_size), replace=False)
        F0, F1 = F0[i0], F1[i1]
    nn = find_nn_gpu(F0, F1, nn_max_n=nn_max_n)
    if sub:
        return xyz0[i0], xyz1[i1[nn]]
    return xyz0, xyz1[nn]


def registration_based_on_corr(source_pcd, target_pcd, max_corr_dist=0.03, seed=0, *,
                               minimal_sets=None, n_hypotheses=100000, adaptive=True):
    """utils/eval_pose.py:82-100 with the Open3D call restated (oracle/ransac_np.py).

    Without explicit ``minimal_sets`` they are drawn with ransac_np.draw_minimal_sets(M, H,
    10, seed) -- the repo-wide convention replacing Open3D's unreproducible RNG stream.
    """
    src = np.asarray(source_pcd).astype(np.float64)
    tgt = np.asarray(target_pcd).astype(np.float64)
    if minimal_sets is None:
        minimal_sets = ransac_np.draw_minimal_sets(len(src), n_hypotheses, 10, seed)
    return ransac_np.ransac(src, tgt, max_corr_dist, minimal_sets, adaptive=adaptive)["T"]


def rot_y_mat(theta):
    """transforms3d.euler.euler2mat(0, theta, 0) as used at utils/eval_pose.py:114."""
    c, s = np.cos(theta), np.sin(theta)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def eval_pose(T_est, T0, T1, axis_symmetry=1):
    """utils/eval_pose.py:103-128; T_est is a torch tensor."""
    t_best, r_best = np.inf, np.inf
    for i in range(axis_symmetry):
        trans = np.eye(4)
        trans[:3, :3] = rot_y_mat(i * (2 * np.pi / axis_symmetry))
        T_gt = (T1 @ (np.linalg.inv(trans) @ np.linalg.inv(T0))).astype(np.float32)
        # float32 argument, arccos in float64: the dtype / values of the r_losses_* caches the reference's own
        # evaluation.py wrote (data/cache_ret: float64 whose cosines are float32 values); today's NumPy
        # would keep float32 here
        arg = np.float32(np.clip((np.trace(np.asarray(T_est[:3, :3].t() @ T_gt[:3, :3])) - 1) / 2, -1, 1))
        r = np.arccos(np.float64(arg))
        t = np.linalg.norm(T_est[:3, 3] - T_gt[:3, 3])
        if r_best > r:
            r_best, t_best = r, t
    return t_best, r_best


# ------------------------------------------------------------------------ preprocess.py


def apply_transform(pointcloud, T):
    """utils/preprocess.py:39-48."""
    homo = np.concatenate([pointcloud, np.ones([len(pointcloud), 1])], 1)
    return np.matmul(T, homo.T).T[:, :3]


def chamfer_gpu(pc0, pc1):
    """utils/preprocess.py:51-56 == utils/pc_dist.py:45-52."""
    nrm = (pc0[None, :, :] - pc1[:, None, :]).norm(dim=2)
    return nrm.min(0)[0].mean() + nrm.min(1)[0].mean()


chamfer = chamfer_gpu


def chamfer_gpu_1direction(pc0, pc1):
    """utils/preprocess.py:59-64."""
    return (pc0[None, :, :] - pc1[:, None, :]).norm(dim=2).min(0)[0].mean().item()


def chamfer_kdtree_1direction(pc0, pc1):
    """utils/preprocess.py:67-70."""
    dd, _ = KDTree(pc1).query(pc0, k=1)
    return dd.mean()


def load_norm_pc_array(pc, samples):
    """utils/preprocess.py:32-36 on an in-memory array."""
    pc = np.array(pc[:samples], copy=True)
    pc -= pc.mean(0)
    return pc / np.max(np.linalg.norm(pc, 2, 1))
