"""Drop-in for the reference's utils/eval_pose.py (find_corr :28-45, find_kcorr :48-79,
registration_based_on_corr :82-100, eval_pose :103-128)."""
import numpy as np
import torch

from . import _capi
from ._capi import ptr
from .find_nn import find_knn_cpu, find_nn_cpu, find_nn_gpu  # noqa: F401  (re-exported like the reference)

RANSAC_N = 10            # utils/eval_pose.py:96
MAX_ITERATION = 100000   # Open3D RANSACConvergenceCriteria default
CONFIDENCE = 0.999       # Open3D RANSACConvergenceCriteria default


def find_corr(xyz0, xyz1, F0, F1, nn_max_n, subsample_size=-1):
    """utils/eval_pose.py:28-45 (global np.random for the optional subsample)."""
    subsample = len(F0) > subsample_size
    if subsample_size > 0 and subsample:
        N0 = min(len(F0), subsample_size)
        N1 = min(len(F1), subsample_size)
        inds0 = np.random.choice(len(F0), N0, replace=False)
        inds1 = np.random.choice(len(F1), N1, replace=False)
        F0, F1 = F0[inds0], F1[inds1]
    nn_inds = find_nn_gpu(F0, F1, nn_max_n=nn_max_n)
    if subsample_size > 0 and subsample:
        return xyz0[inds0], xyz1[inds1[nn_inds]]
    return xyz0, xyz1[nn_inds]


def find_kcorr(F0, F1, k=1, nn_max_n=500, subsample_size=-1):
    """utils/eval_pose.py:48-79 -> (inds0 [N0*k], inds1 [N0*k]) int64."""
    subsample = len(F0) > subsample_size
    if subsample_size > 0 and subsample:
        N0 = min(len(F0), subsample_size)
        N1 = min(len(F1), subsample_size)
        inds0 = np.random.choice(len(F0), N0, replace=False)
        inds1 = np.random.choice(len(F1), N1, replace=False)
        F0, F1 = F0[inds0], F1[inds1]
    if k == 1:
        nn_inds = find_nn_cpu(F0, F1).flatten()
    else:
        nn_inds = find_knn_cpu(F0, F1, k=k).flatten()
    if subsample_size > 0 and subsample:
        return np.repeat(inds0, k), inds1[nn_inds]
    return np.repeat(np.arange(len(F0)), k), nn_inds


def ransac_device(src_xyz, tgt_xyz, max_corr_dist, seed=0, *, src_idx=None, tgt_idx=None,
                  minimal_sets=None, n_hypotheses=MAX_ITERATION, adaptive=True, confidence=CONFIDENCE,
                  return_all=False, ctx=None):
    """Device-resident RANSAC.  Correspondence i pairs src_xyz[src_idx[i]] with
    tgt_xyz[tgt_idx[i]] (identity when the index arrays are None).  Returns a dict of CUDA
    tensors: T [4,4] f64, stats [4] f64 = (best_h, inliers, sum_d2, n_evaluated) and, with
    return_all, counts [H] i32 / sums [H] f64."""
    ctx = ctx or _capi.get_context()
    s = ctx.dev(src_xyz, torch.float32)
    t = ctx.dev(tgt_xyz, torch.float32)
    si = ctx.dev(src_idx, torch.int64) if src_idx is not None else None
    ti = ctx.dev(tgt_idx, torch.int64) if tgt_idx is not None else None
    m = len(si) if si is not None else (len(ti) if ti is not None else s.shape[0])
    if si is None and ti is None and s.shape[0] != t.shape[0]:
        raise ValueError("source and target must pair row by row")
    ms = None
    H = int(n_hypotheses)
    n = RANSAC_N
    if minimal_sets is not None:
        ms = ctx.dev(minimal_sets, torch.int32)
        H, n = ms.shape
    T = ctx.empty((4, 4), torch.float64)
    stats = ctx.empty((4,), torch.float64)
    counts = torch.zeros((H,), dtype=torch.int32, device=ctx.device) if return_all else None
    sums = torch.zeros((H,), dtype=torch.float64, device=ctx.device) if return_all else None
    ctx.call("cb_ransac", ptr(s), s.shape[0], ptr(t), t.shape[0], ptr(si), ptr(ti), m, ptr(ms),
             int(seed) & 0xFFFFFFFFFFFFFFFF, H, n, float(max_corr_dist), int(bool(adaptive)),
             float(confidence), ptr(T), ptr(stats), ptr(counts), ptr(sums), ctx.stream)
    out = {"T": T, "stats": stats}
    if return_all:
        out["counts"], out["sums"] = counts, sums
    return out


def draw_minimal_sets(M, H, ransac_n=RANSAC_N, seed=0, ctx=None):
    """The library's seeded minimal-set rule (cb_draw_minimal_sets) -> [H, ransac_n] int32 NumPy."""
    ctx = ctx or _capi.get_context()
    out = ctx.empty((H, ransac_n), torch.int32)
    ctx.call("cb_draw_minimal_sets", int(seed) & 0xFFFFFFFFFFFFFFFF, int(M), int(H), int(ransac_n),
             ptr(out), ctx.stream)
    return out.cpu().numpy()


def registration_based_on_corr(source_pcd, target_pcd, max_corr_dist=0.03, seed=0, *,
                               minimal_sets=None, n_hypotheses=MAX_ITERATION, adaptive=True):
    """utils/eval_pose.py:82-100 -> T_est [4,4] float64 NumPy.

    Open3D's RNG stream cannot be reproduced (shared engine inside an OpenMP loop,
    README.md:260 of the reference); minimal sets are explicit (``minimal_sets``) or drawn by the
    library's counter-based rule from ``seed``."""
    res = ransac_device(np.asarray(source_pcd), np.asarray(target_pcd), max_corr_dist, seed,
                        minimal_sets=minimal_sets, n_hypotheses=n_hypotheses, adaptive=adaptive)
    return res["T"].cpu().numpy()


def _rot_y(theta):
    # transforms3d.euler.euler2mat(0, theta, 0) (static xyz): rotation about +y
    c, s = np.cos(theta), np.sin(theta)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def eval_pose(T_est, T0, T1, axis_symmetry=1):
    """utils/eval_pose.py:103-128 (host NumPy, negligible cost).  T_est is a torch tensor."""
    t_loss_best = np.inf
    r_loss_best = np.inf
    for i in range(axis_symmetry):
        trans = np.eye(4)
        trans[:3, :3] = _rot_y(i * (2 * np.pi / axis_symmetry))
        T_gt = np.matmul(T1, np.matmul(np.linalg.inv(trans), np.linalg.inv(T0))).astype(np.float32)
        # float32 argument, arccos evaluated in float64: what the reference's pinned NumPy / torch did — the
        # r_losses_* files it wrote (data/cache_ret) are float64 whose cosines are float32 values, the
        # t_losses_* files float32 (tests/golden/cache_format.json).  Today's NumPy would return float32.
        arg = np.float32(np.clip((np.trace(np.asarray(T_est[:3, :3].t() @ T_gt[:3, :3])) - 1) / 2, -1, 1))
        r_loss = np.arccos(np.float64(arg))
        t_loss = np.linalg.norm(T_est[:3, 3] - T_gt[:3, 3])
        if r_loss_best > r_loss:
            r_loss_best = r_loss
            t_loss_best = t_loss
    return t_loss_best, r_loss_best
