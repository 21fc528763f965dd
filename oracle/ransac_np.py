"""NumPy float64 restatement of Open3D's correspondence RANSAC.  TEST INFRASTRUCTURE.

PARITY UNPINNED: restates ``open3d.pipelines.registration.
registration_ransac_based_on_correspondence`` (Open3D commit 160209d0, pinned by the
reference's docker/Dockerfile:77-80; called at utils/eval_pose.py:95-97 with ransac_n=10,
TransformationEstimationPointToPoint(with_scaling=False), no checkers,
RANSACConvergenceCriteria(100000, 0.999)).  Open3D is absent from the reference tree and
cannot be installed here, and the reference holds no golden vectors for this call, so the
restatement follows the published algorithm:

  fit    Eigen::umeyama without scaling: R = U diag(1,1,s) V^T of
         Sigma = 1/n sum (t_i-mu_t)(s_i-mu_s)^T, s=-1 iff det(U)det(V)<0; t = mu_t - R mu_s
  score  d2_i = |R s_i + t - t_i|^2; inlier iff d2_i < max_dist^2 (strict);
         fitness = good/M; inlier_rmse = sqrt(sum_inlier d2 / good) (0 if good==0)
  keep   better iff fitness greater, or equal fitness and rmse smaller (strict: the lowest
         hypothesis index wins exact ties -- the reference's own order is thread dependent)
  stop   after an improvement: k = log(1-conf)/log(1-(good/M)^n); budget=min(budget,ceil(k))

Open3D draws the minimal sets from a shared RNG inside an OpenMP loop (non-reproducible,
README.md:260); here they are an explicit input ``minimal_sets [H, n]`` (uniform with
replacement from [0, M)), exactly what the CUDA path consumes.
"""
import math

import numpy as np


def umeyama_noscale(src, tgt):
    """src,tgt [n,3] float64 -> 4x4 float64 (a10)."""
    src = np.asarray(src, np.float64)
    tgt = np.asarray(tgt, np.float64)
    mu_s = src.mean(0)
    mu_t = tgt.mean(0)
    sigma = (tgt - mu_t).T @ (src - mu_s) / src.shape[0]
    U, _, Vt = np.linalg.svd(sigma)
    S = np.ones(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2] = -1.0
    R = (U * S) @ Vt
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = mu_t - R @ mu_s
    return T


def score(T, src, tgt, max_dist):
    """(a11) -> (good:int, sum_d2:float)."""
    R, t = T[:3, :3], T[:3, 3]
    y = src[:, 0:1] * R[:, 0] + src[:, 1:2] * R[:, 1]
    y = y + src[:, 2:3] * R[:, 2]
    y = y + t
    e = y - tgt
    d2 = e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1] + e[:, 2] * e[:, 2]
    inl = d2 < max_dist * max_dist
    return int(inl.sum()), float(d2[inl].sum())


def _rmse(good, s):
    return math.sqrt(s / good) if good > 0 else 0.0


def budget_from_count(good, M, ransac_n, confidence, cap):
    """Open3D's exit-condition update; returns the new iteration budget (<= cap)."""
    if confidence >= 1.0 or good <= 0:
        return cap
    w = good / M
    denom = math.log(1.0 - w ** ransac_n) if w < 1.0 else -math.inf
    if denom == 0.0:
        return cap
    k = math.log(1.0 - confidence) / denom
    return int(math.ceil(k)) if k < cap else cap


def ransac(src, tgt, max_dist, minimal_sets, adaptive=False, confidence=0.999,
           return_all=False):
    """Sequential restatement.  src,tgt [M,3]; minimal_sets [H,n] int.

    Returns dict(T [4,4] f64, best_h, inliers, sum_d2, n_evaluated[, counts [H], sums [H]]).
    """
    src = np.asarray(src, np.float64)
    tgt = np.asarray(tgt, np.float64)
    ms = np.asarray(minimal_sets)
    H, n = ms.shape
    M = src.shape[0]
    out = dict(T=np.eye(4), best_h=-1, inliers=0, sum_d2=0.0, n_evaluated=0)
    if n < 3 or M < n or max_dist <= 0.0:
        return out
    best = (0.0, 0.0)  # (fitness, rmse) of RegistrationResult() default
    budget = H
    counts = np.zeros(H, np.int64)
    sums = np.zeros(H, np.float64)
    h = 0
    while h < budget:
        idx = ms[h]
        T = umeyama_noscale(src[idx], tgt[idx])
        good, s = score(T, src, tgt, max_dist)
        counts[h], sums[h] = good, s
        fit, rm = good / M, _rmse(good, s)
        if fit > best[0] or (fit == best[0] and rm < best[1]):
            best = (fit, rm)
            out.update(T=T, best_h=h, inliers=good, sum_d2=s)
            if adaptive:
                budget = min(budget, budget_from_count(good, M, n, confidence, budget))
        h += 1
    out["n_evaluated"] = h
    if return_all:
        out["counts"], out["sums"] = counts[:h], sums[:h]
    return out


def splitmix64(x):
    """splitmix64 finaliser on a uint64 array (wrapping arithmetic)."""
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def draw_minimal_sets(M, H, ransac_n=10, seed=0):
    """The repo-wide convention for seeded minimal sets (include/corsair_b200.h,
    cb_draw_minimal_sets): counter-based, uniform with replacement over [0, M):
        set[h][j] = (hi32(splitmix64(seed*0x9E3779B97F4A7C15 + h*ransac_n + j)) * M) >> 32
    It replaces Open3D's shared-RNG-inside-OpenMP stream, which is not reproducible."""
    with np.errstate(over="ignore"):
        c = np.arange(H * ransac_n, dtype=np.uint64)
        base = np.array([seed & 0xFFFFFFFFFFFFFFFF], dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
        x = splitmix64(base + c)
        idx = ((x >> np.uint64(32)) * np.uint64(M)) >> np.uint64(32)
    return idx.astype(np.int32).reshape(H, ransac_n)
