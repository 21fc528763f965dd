"""Drop-in for the hot functions of the reference's utils/symmetry.py: split_corr (:145-179),
symmetric_cut4 (:182-259) and sym_pose (:262-358).  The four visualisation helpers of that file
need a graphics stack and are out of scope.

Everything numeric runs in the C-ABI library: feature kNN (cb_knn / cb_knn_ragged),
correspondence RANSAC (cb_ransac), one-direction Chamfer (cb_chamfer) and the three device pieces
of the symmetry segmentation (cb_feature_rows_topk, cb_kmeans_batch, cb_assign_labels).  Host code
only draws the 100 seed points from the global NumPy RNG exactly like the reference, replays the
(data-independent) RandomState(0) stream sklearn's k-means++ would consume, and applies the
sequential acceptance rule to 100 rows of statistics.
"""
import ctypes
import functools

import numpy as np
import torch
from tqdm import tqdm

from . import _capi
from ._capi import ptr
from .eval_pose import *  # noqa: F401,F403  (the reference re-exports these: utils/symmetry.py:33)
from .eval_pose import CONFIDENCE, find_kcorr, ransac_device
from .preprocess import *  # noqa: F401,F403

N_NN = 50        # utils/symmetry.py:214
N_INIT = 10      # KMeans(..., n_init=10), utils/symmetry.py:216
MIN_CENTER = 0.15  # utils/symmetry.py:234


@functools.lru_cache(maxsize=None)
def _kmeanspp_draws(n_nn, K, n_init=N_INIT):
    """The draws sklearn's KMeans(random_state=0, n_init) k-means++ takes from RandomState(0) for
    n_nn equally weighted samples: per init one ``choice`` (first centre) and K-1 ``uniform``
    vectors of 2+int(log(K)) local trials.  They do not depend on the data."""
    rs = np.random.RandomState(0)
    sw = np.ones(n_nn, dtype=np.float32)
    p = sw / sw.sum()
    n_trials = 2 + int(np.log(K))
    first = np.zeros(n_init, np.int32)
    unif = np.zeros((n_init, max(K - 1, 1), n_trials), np.float64)
    for i in range(n_init):
        first[i] = rs.choice(n_nn, p=p)
        for c in range(1, K):
            unif[i, c - 1] = rs.uniform(size=n_trials)
    return first, unif, n_trials


def split_corr(pcsA, pcsB, featsA, featsB, knn, subsample_size=1000):
    """utils/symmetry.py:145-179."""
    xyzA_corrs, xyzB_corrs = [], []
    for pcA, pcB, featA, featB in zip(pcsA, pcsB, featsA, featsB):
        idx_0, idx_1 = find_kcorr(featA, featB, k=knn, subsample_size=subsample_size)
        xyzA_corrs.append(pcA[idx_0])
        xyzB_corrs.append(pcB[idx_1])
    return np.concatenate(xyzA_corrs, axis=0), np.concatenate(xyzB_corrs, axis=0)


def cut_model_device(feat_d, xyz_d, K, random_points, ctx):
    """Device part of symmetric_cut4 for the sampled seed rows.  Returns (centers [S,K,3] f32,
    stats [S,4] f64, nn_idx [S,n_nn] i32) as CUDA tensors."""
    n, d = feat_d.shape
    S = len(random_points)
    n_nn = min(N_NN, n)
    first, unif, n_trials = _kmeanspp_draws(n_nn, K)
    samp = ctx.dev(np.asarray(random_points, dtype=np.int32), torch.int32)
    nn_idx = ctx.empty((S, n_nn), torch.int32)
    ctx.call("cb_feature_rows_topk", ptr(feat_d), n, d, ptr(samp), S, n_nn, ptr(nn_idx), ctx.stream)
    cache = ctx.__dict__.setdefault("_km_draws", {})  # data-independent: uploaded once per context
    if (n_nn, K) not in cache:
        cache[(n_nn, K)] = (ctx.dev(first, torch.int32), ctx.dev(unif, torch.float64))
    first_d, unif_d = cache[(n_nn, K)]
    centers = ctx.empty((S, K, 3), torch.float32)
    stats = ctx.empty((S, 4), torch.float64)
    ctx.call("cb_kmeans_batch", ptr(xyz_d), n, ptr(nn_idx), S, n_nn, K, ptr(first_d), ptr(unif_d), N_INIT,
             n_trials, ptr(centers), ptr(stats), ctx.stream)
    return centers, stats, nn_idx


def select_cut(stats):
    """The sequential acceptance rule of utils/symmetry.py:234-236 over the per-sample statistics
    (min centre distance, max intra-cluster error, std of label ratios).  -1: no model."""
    curr_var, win = 100, -1
    for i, (dmin, emax, sd, _) in enumerate(stats):
        if dmin > MIN_CENTER > emax and curr_var > sd:
            curr_var, win = sd, i
    return win


def _masks_device(feat_d, xyz_d, K, max_sample, ctx, rng=None):
    if K not in (2, 4):
        raise ValueError("not defined")
    n = xyz_d.shape[0]
    # global RNG like utils/symmetry.py:193 unless the caller hands in its own RandomState
    random_points = (rng or np.random).choice(n, max_sample, replace=False)
    centers, stats, _ = cut_model_device(feat_d, xyz_d, K, random_points, ctx)
    win = select_cut(stats.cpu().numpy())
    if win < 0:
        # the reference dereferences curr_model = None here (utils/symmetry.py:238)
        raise AttributeError("'NoneType' object has no attribute 'cluster_centers_'")
    c = centers[win].contiguous()
    labels = ctx.empty((n,), torch.int32)
    ctx.call("cb_assign_labels", ptr(xyz_d), n, ptr(c), K, ptr(labels), None, ctx.stream)
    if K == 2:
        order = [0, 1]
    else:
        ch = c.cpu().numpy()
        dist = np.linalg.norm(ch[None, :, :] - ch[:, None, :], 2, 2)
        dist[np.arange(K), np.arange(K)] = 100
        rank_centers = np.argsort(dist[0, 1:])
        order = [0, rank_centers[0] + 1, rank_centers[2] + 1, rank_centers[1] + 1]
    return [labels == int(o) for o in order]


def _segments_device(feat_d, xyz_d, K, max_sample, ctx, rng=None):
    """symmetric_cut4 without a host round trip: (perm [n] int64, seg_off [K+2] int32) on the device
    — the concatenated ascending point indices of the K masks, their offsets and the accepted
    sample (-1: none).  Nothing is synchronised here."""
    if K not in (2, 4):
        raise ValueError("not defined")
    n = xyz_d.shape[0]
    random_points = (rng or np.random).choice(n, max_sample, replace=False)
    centers, stats, _ = cut_model_device(feat_d, xyz_d, K, random_points, ctx)
    perm = ctx.empty((n,), torch.int64)
    seg_off = ctx.empty((K + 2,), torch.int32)
    ctx.call("cb_symmetry_segments", ptr(xyz_d), n, ptr(centers), ptr(stats), len(random_points), K, MIN_CENTER,
             ptr(perm), None, ptr(seg_off), ctx.stream)
    return perm, seg_off


def symmetric_cut4(feat, raw_pc, K, max_sample=100):
    """utils/symmetry.py:182-259 -> list of K boolean NumPy masks over the N points."""
    ctx = _capi.get_context()
    feat_d = ctx.dev(feat, torch.float32)
    xyz_d = ctx.dev(raw_pc, torch.float32)
    return [m.cpu().numpy() for m in _masks_device(feat_d, xyz_d, K, max_sample, ctx)]


def _pair_maps(K, second_sweep):
    """pos segment matched with base segment s in every pairing of utils/symmetry.py:303-356: the K
    cyclic rotations of the pos segments, then (pos_sym >= 2) those of [B0, B3, B2, B1]."""
    maps = [[(s + r) % K for s in range(K)] for r in range(K)]
    if second_sweep:
        order = [0, 3, 2, 1]
        maps += [[order[(s + r) % K] for s in range(K)] for r in range(K)]
    return np.asarray(maps, dtype=np.int32)


def sym_pose(baseF, xyz0, posF, xyz1, pos_sym, k_nn=5, max_corr=0.20, seed=0, *, n_hypotheses=100000,
             adaptive=True, rng=None, details=False):
    """utils/symmetry.py:262-358: pose with and without the symmetry-aided pairing search.

    ``rng``: where the segmentation's seed points come from instead of the global NumPy RNG the
    reference draws from — a ``np.random.RandomState``, or an int ``s``: the points
    ``np.random.seed(s)`` followed by the two ``choice`` calls would give, replayed inside the C call
    (no interpreter time); one of the two is needed when pairs are registered concurrently
    (sym_pose_many).

    Everything after the seed-row draws — both segmentations, the vanilla registration and every
    pairing (kNN -> RANSAC -> Chamfer each) — is one cb_sym_pose call.

    Returns (T_est_best [4,4] f32 tensor, chamf_best float, T_est_ransac, chamf_ransac, success);
    ``details=True`` appends a dict: ``best_index`` (0 = the vanilla registration, 1 + p = pairing p in the
    order of :303-356), ``n_pairings`` evaluated and the call's ``status``."""
    ctx = _capi.get_context()
    baseF_d = ctx.dev(baseF, torch.float32)
    posF_d = ctx.dev(posF, torch.float32)
    xyz0_d = ctx.dev(xyz0, torch.float32)
    xyz1_d = ctx.dev(xyz1, torch.float32)
    n0, n1, d = baseF_d.shape[0], posF_d.shape[0], baseF_d.shape[1]
    if n1 < k_nn:
        raise ValueError(f"k_nn={k_nn} exceeds the {n1} target points")

    # host side of the pair: the two seed-row draws (global NumPy RNG like utils/symmetry.py:193
    # unless the caller hands in a RandomState), then ONE C call that releases the GIL until the
    # poses are back on the host.  The reference segments pos only when base succeeded (:292-301):
    # if base accepts no model, the stream is rewound to where the reference would have left it.
    K = 4 if pos_sym >= 2 else 2
    why, rewind, n_s, np_seed = None, None, 100, -1
    s_base = s_pos = None
    if isinstance(rng, (int, np.integer)):
        np_seed = int(rng)  # drawn inside the call, as after np.random.seed(rng)
        if not 0 <= np_seed < 2 ** 32:
            raise ValueError("Seed must be between 0 and 2**32 - 1")
        if min(n0, n1) < n_s:
            why = "Cannot take a larger sample than population when 'replace=False'"
    else:
        src = rng or np.random
        try:  # same swallow-and-fallback as utils/symmetry.py:292-301
            s_base = src.choice(n0, n_s, replace=False).astype(np.int32)
            rewind = src.get_state()
            s_pos = src.choice(n1, n_s, replace=False).astype(np.int32)
        except Exception as e:
            why, n_s = str(e), 0
            s_base = s_pos = None
    n_nn = min(N_NN, n0, n1)
    first, unif, n_trials = _kmeanspp_draws(n_nn, K)
    out = np.empty(40, np.float64)
    vp = _capi.c_void_p
    ctx.call("cb_sym_pose", ptr(baseF_d), ptr(xyz0_d), n0, ptr(posF_d), ptr(xyz1_d), n1, d, int(pos_sym), int(k_nn),
             s_base.ctypes.data_as(vp) if s_base is not None else None,
             s_pos.ctypes.data_as(vp) if s_pos is not None else None, n_s, np_seed, n_nn, first.ctypes.data_as(vp),
             unif.ctypes.data_as(vp), N_INIT, n_trials, MIN_CENTER, float(max_corr), int(seed) & 0xFFFFFFFFFFFFFFFF,
             int(n_hypotheses), int(bool(adaptive)), CONFIDENCE, out.ctypes.data_as(vp), ctx.stream)
    status = int(out[34])
    if status in (1, 2):
        if status == 1 and rewind is not None:
            src.set_state(rewind)
        # the reference dereferences curr_model = None here (utils/symmetry.py:237)
        why = "'NoneType' object has no attribute 'cluster_centers_'"
    if why is not None:
        tqdm.write(f"symmetry failed use ransac: {why}")
    T = torch.from_numpy(out[:34].astype(np.float32))
    res = (T[:16].reshape(4, 4), float(out[16]), T[17:33].reshape(4, 4), float(out[33]), why is None)
    if details:
        res += (dict(best_index=int(out[35]), n_pairings=int(out[36]), status=status),)
    return res


# ---------------------------------------------------------------------------------------------
# Concurrent driver.  The reference registers its pairs one after the other (evaluation.py:297)
# or under tqdm's thread_map (evaluation-shapenet.py:341); pairs are independent, and one pair's
# chain of small dependent kernels (fit, fold, k-means: a handful of blocks each) cannot fill 148
# SMs.  cb_sym_pose_many keeps `workers` pairs in flight from the calling thread alone (each pair on
# its own internal streams; the thread polls their events and enqueues whichever pair can go on), so
# the device overlaps one pair's latency-bound kernels with another pair's scoring and no host thread
# spins in the driver.  (Round 1 ran this loop on a pool of Python threads.)
DEFAULT_WORKERS = 8


def driver_name():
    return "cb_sym_pose_many: one host thread, pairs in flight on internal streams, event polling"


def _job_record(ctx, args, kwargs, dev, keep):
    baseF, xyz0, posF, xyz1, pos_sym = args[:5]
    kwargs = dict(zip(("k_nn", "max_corr", "seed"), args[5:]), **kwargs)  # sym_pose's positional order
    k_nn, max_corr, seed = kwargs.get("k_nn", 5), kwargs.get("max_corr", 0.20), kwargs.get("seed", 0)
    baseF_d, posF_d, xyz0_d, xyz1_d = dev(baseF), dev(posF), dev(xyz0), dev(xyz1)
    n0, n1, d = baseF_d.shape[0], posF_d.shape[0], baseF_d.shape[1]
    if n1 < k_nn:
        raise ValueError(f"k_nn={k_nn} exceeds the {n1} target points")
    np_seed = int(kwargs["rng"])
    if not 0 <= np_seed < 2 ** 32:
        raise ValueError("Seed must be between 0 and 2**32 - 1")
    K = 4 if pos_sym >= 2 else 2
    n_nn = min(N_NN, n0, n1)
    first, unif, n_trials = _kmeanspp_draws(n_nn, K)
    out = np.empty(40, np.float64)
    keep += [baseF_d, posF_d, xyz0_d, xyz1_d, out]
    vp = _capi.c_void_p
    rec = _capi.SymJob(baseF_d.data_ptr(), xyz0_d.data_ptr(), posF_d.data_ptr(), xyz1_d.data_ptr(), n0, n1, d,
                       int(pos_sym), int(k_nn), 100, n_nn, N_INIT, n_trials, int(kwargs.get("n_hypotheses", 100000)),
                       int(bool(kwargs.get("adaptive", True))), 0, np_seed, None, None,
                       first.ctypes.data_as(vp), unif.ctypes.data_as(vp), MIN_CENTER, float(max_corr), CONFIDENCE,
                       int(seed) & 0xFFFFFFFFFFFFFFFF, out.ctypes.data_as(vp))
    return rec, out, min(n0, n1) < 100


def sym_pose_many(jobs, workers=DEFAULT_WORKERS):
    """Register independent pairs concurrently.  ``jobs``: iterable of ``(args, kwargs)`` for
    sym_pose; returns sym_pose's result tuples in job order.  Jobs whose ``rng`` is an int seed (the
    seed rows are then replayed inside the C call) go through ONE cb_sym_pose_many call that keeps
    ``workers`` pairs in flight; a list containing any other kind of ``rng`` (a RandomState, or the
    global NumPy stream the reference draws from) keeps its sequential draw order and is registered
    one pair after the other.  Host arrays shared between jobs (one query against ten candidates) are
    uploaded once per call."""
    jobs = list(jobs)
    if not jobs:
        return []
    if not all(isinstance(k.get("rng"), (int, np.integer)) for _, k in jobs):
        return [sym_pose(*a, **k) for a, k in jobs]
    ctx = _capi.get_context()
    cache = {}

    def dev(x):
        if isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.float32 and x.is_contiguous():
            return x
        key = id(x)
        if key not in cache:
            cache[key] = ctx.dev(x, torch.float32)
        return cache[key]

    keep, outs, small = [], [], []
    recs = (_capi.SymJob * len(jobs))()
    # longest first: a pair with pos_sym >= 2 runs 9 registrations, the others 3 (utils/symmetry.py:303-356)
    order = sorted(range(len(jobs)), key=lambda i: 0 if jobs[i][0][4] >= 2 else 1)
    slot_of = {}
    for pos, i in enumerate(order):
        recs[pos], out, sm = _job_record(ctx, jobs[i][0], jobs[i][1], dev, keep)
        slot_of[i] = pos
        outs.append(out)
        small.append(sm)
    status = (ctypes.c_int32 * len(jobs))()
    ctx.call("cb_sym_pose_many", ctypes.cast(recs, _capi.c_void_p), len(jobs), int(workers),
             ctypes.cast(status, _capi.c_void_p), ctx.stream)
    results = []
    for i in range(len(jobs)):
        out = outs[slot_of[i]]
        st = int(out[34])
        why = None
        if small[slot_of[i]]:
            why = "Cannot take a larger sample than population when 'replace=False'"
        elif st in (1, 2):
            why = "'NoneType' object has no attribute 'cluster_centers_'"
        if why is not None:
            tqdm.write(f"symmetry failed use ransac: {why}")
        T = torch.from_numpy(out[:34].astype(np.float32))
        results.append((T[:16].reshape(4, 4), float(out[16]), T[17:33].reshape(4, 4), float(out[33]), why is None))
    return results
