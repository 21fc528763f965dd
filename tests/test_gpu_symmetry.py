"""GPU parity: symmetric_cut4 pieces, sym_pose end to end, IRLS pose."""
import numpy as np
import pytest
import torch

from corsair_b200 import synth
from oracle import port, port_retrieval, port_symmetry, ransac_c, ransac_np

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("d", [16, 32])
def test_feature_rows_top50(ctx, rng, d):
    from corsair_b200 import _capi
    p = synth.make_pair(5, d, "chair")
    feat = p["posF"]
    n = len(feat)
    samp = rng.choice(n, 100, replace=False).astype(np.int32)
    f = ctx.dev(feat, torch.float32)
    s = ctx.dev(samp, torch.int32)
    out = ctx.empty((100, 50), torch.int32)
    ctx.call("cb_feature_rows_topk", _capi.ptr(f), n, d, _capi.ptr(s), 100, 50, _capi.ptr(out), ctx.stream)
    out = out.cpu().numpy()
    for r, index in enumerate(samp):
        d2 = ((feat[index].astype(np.float64) - feat.astype(np.float64)) ** 2).sum(1)
        ref = np.sort(np.lexsort((np.arange(n), d2))[:50])
        assert np.array_equal(out[r], ref)


@pytest.mark.parametrize("category,K", [("chair", 2), ("table", 4)])
def test_symmetric_cut4_masks(ctx, category, K):
    """Masks of the device segmentation vs the reference algorithm (sklearn KMeans) under the same
    global NumPy seed on three full-size clouds: EQUAL masks (round 1 asserted an agreement rate; the
    device k-means replays sklearn's float32 operations step by step and no cloud tried since has
    differed — tests/test_gpu_golden.py::test_symmetric_cut4_ten_seeded_clouds lists per-model
    differences if one ever does)."""
    from corsair_b200 import symmetry
    for seed in range(3):
        p = synth.make_pair(seed, 32, category)
        np.random.seed(31)
        ref_c, ref_win = port_symmetry.cut_candidates(p["posF"], p["xyz1"], K)
        ref = port_symmetry.masks_from_model(ref_c[ref_win]["centers"], ref_c[ref_win]["labels"], K)
        np.random.seed(31)
        got = symmetry.symmetric_cut4(p["posF"], p["xyz1"], K)
        assert len(got) == K and all(m.dtype == bool and m.shape == (len(p["xyz1"]),) for m in got)
        assert np.stack(got).sum(0).min() == 1 and np.stack(got).sum(0).max() == 1  # a partition
        assert np.array_equal(np.stack(got), np.stack(ref)), f"seed {seed}: masks differ from sklearn's"


@pytest.mark.parametrize("category,K", [("chair", 2), ("table", 4), ("chair", 4), ("table", 2)])
def test_segments_device_equals_masks(ctx, category, K):
    """cb_symmetry_segments (selection rule + labels + centre ranking + permutation on the device)
    against the host-side selection of symmetric_cut4 on the same draws: bit-exact."""
    from corsair_b200 import symmetry
    for seed in range(3):
        p = synth.make_pair(seed, 32, category)
        feat_d, xyz_d = ctx.dev(p["posF"], torch.float32), ctx.dev(p["xyz1"], torch.float32)
        try:
            masks = symmetry._masks_device(feat_d, xyz_d, K, 100, ctx, np.random.RandomState(5))
        except AttributeError:
            masks = None
        perm, off = symmetry._segments_device(feat_d, xyz_d, K, 100, ctx, np.random.RandomState(5))
        perm, off = perm.cpu().numpy(), off.cpu().numpy()
        if masks is None:
            assert off[K + 1] == -1 and not off[:K + 1].any()
            continue
        assert off[K + 1] >= 0 and off[0] == 0 and off[K] == len(p["xyz1"])
        want = [np.nonzero(m.cpu().numpy())[0] for m in masks]
        assert np.array_equal(np.diff(off[:K + 1]), [len(w) for w in want])
        assert np.array_equal(perm, np.concatenate(want))


def test_segments_device_no_model(ctx, rng):
    """No sample passes the acceptance rule -> winner -1, zero offsets (the reference raises, :237)."""
    from corsair_b200 import symmetry
    n, K, S = 3000, 2, 7
    xyz = ctx.dev(rng.standard_normal((n, 3)).astype(np.float32), torch.float32)
    centers = ctx.dev(rng.standard_normal((S, K, 3)).astype(np.float32), torch.float32)
    stats = np.tile(np.array([0.1, 0.05, 0.2, 0.0]), (S, 1))  # min centre distance below 0.15
    stats[3, 1] = np.nan
    perm, off = ctx.empty((n,), torch.int64), ctx.empty((K + 2,), torch.int32)
    from corsair_b200._capi import ptr
    ctx.call("cb_symmetry_segments", ptr(xyz), n, ptr(centers), ptr(ctx.dev(stats, torch.float64)), S, K, 0.15, ptr(perm),
             None, ptr(off), ctx.stream)
    assert off.cpu().tolist() == [0, 0, 0, -1]
    # one acceptable row, a better (lower ratio std) one later, and a NaN error row in between
    stats[1] = [0.3, 0.1, 0.4, 0.0]
    stats[3] = [0.3, np.nan, 0.0, 0.0]
    stats[5] = [0.3, 0.1, 0.2, 0.0]
    stats[6] = [0.3, 0.1, 0.2, 0.0]  # equal std: the earlier row stays
    labels = ctx.empty((n,), torch.int32)
    ctx.call("cb_symmetry_segments", ptr(xyz), n, ptr(centers), ptr(ctx.dev(stats, torch.float64)), S, K, 0.15, ptr(perm),
             ptr(labels), ptr(off), ctx.stream)
    off, perm, labels = off.cpu().numpy(), perm.cpu().numpy(), labels.cpu().numpy()
    assert off[K + 1] == 5 == symmetry.select_cut(stats)
    ref = ctx.empty((n,), torch.int32)
    ctx.call("cb_assign_labels", ptr(xyz), n, ptr(centers[5].contiguous()), K, ptr(ref), None, ctx.stream)
    assert np.array_equal(labels, ref.cpu().numpy())
    assert np.array_equal(perm, np.concatenate([np.nonzero(labels == 0)[0], np.nonzero(labels == 1)[0]]))


def test_kmeans_batch_stats_vs_sklearn(ctx):
    """Per-sample statistics (min centre distance, max error, ratio std) against sklearn on the
    same 50-point neighbourhoods."""
    from corsair_b200 import symmetry
    p = synth.make_pair(1, 32, "chair")
    np.random.seed(7)
    ref_c, _ = port_symmetry.cut_candidates(p["posF"], p["xyz1"], 2)
    picks = np.array([c["index"] for c in ref_c])
    feat_d, xyz_d = ctx.dev(p["posF"], torch.float32), ctx.dev(p["xyz1"], torch.float32)
    centers, stats, nn_idx = symmetry.cut_model_device(feat_d, xyz_d, 2, picks, ctx)
    stats, nn_idx = stats.cpu().numpy(), nn_idx.cpu().numpy()
    ok = 0
    for i, c in enumerate(ref_c):
        assert np.array_equal(nn_idx[i], c["nn_ids"])
        sd = np.sqrt(np.var(c["ratios"]))
        if (abs(stats[i, 0] - c["min_center_dist"]) < 1e-4 and abs(stats[i, 1] - max(c["errors"])) < 1e-4
                and abs(stats[i, 2] - sd) < 1e-3):
            ok += 1
    assert ok >= 95, f"only {ok}/100 neighbourhood models match sklearn"


@pytest.mark.parametrize("category", ["chair", "table"])
def test_sym_pose_matches_oracle(ctx, category):
    """sym_pose against the oracle's restatement on a full-size pair (chair: K = 2, three registrations;
    table: pos_sym = 4, K = 4, nine).  The oracle is handed the DEVICE segmentation (masks of
    symmetry.symmetric_cut4 under the same NumPy seed), so the pairing logic — both sweeps, the mirrored
    order, the strictly-lower-Chamfer selection — is compared unconditionally and independently of k-means
    rounding ties; the masks themselves are pinned against the reference in test_gpu_golden.py."""
    from corsair_b200 import symmetry
    p = synth.make_pair(11, 32, category)
    K = 4 if p["pos_sym"] >= 2 else 2
    calls = []

    def reg(src, tgt, max_corr, seed):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        calls.append(ransac_c.ransac(src, tgt, max_corr, ms, adaptive=True)["T"])
        return calls[-1]

    np.random.seed(31)
    dev_masks = [symmetry.symmetric_cut4(p["baseF"], p["xyz0"], K), symmetry.symmetric_cut4(p["posF"], p["xyz1"], K)]
    it = iter(dev_masks)
    ref = port_symmetry.sym_pose(p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"], k_nn=5,
                                 max_corr=0.2, seed=31, registration=reg, cut=lambda f, x, k, max_sample=100: next(it),
                                 find_kcorr=lambda a, b, k, subsample_size: (np.repeat(np.arange(len(a)), k), port.find_knn_exact(a, b, k)[0].reshape(-1)))
    chs = [port.chamfer_kdtree_1direction(port.apply_transform(p["xyz0"], torch.from_numpy(T.astype(np.float32))), p["xyz1"]) for T in calls]
    best = 0
    for i in range(1, len(chs)):
        if chs[best] > chs[i]:
            best = i
    np.random.seed(31)
    got = symmetry.sym_pose(p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"], k_nn=5, max_corr=0.2, seed=31, details=True)
    assert isinstance(got[0], torch.Tensor) and got[0].dtype == torch.float32 and got[0].shape == (4, 4)
    assert isinstance(got[1], float) and isinstance(got[3], float) and got[4] is True and ref[4] is True
    assert len(calls) == (1 + 2 * K if p["pos_sym"] >= 2 else 1 + K) == got[5]["n_pairings"] + 1
    # vanilla branch: same correspondences and minimal sets -> same pose, same Chamfer
    assert torch.allclose(got[2], ref[2], atol=1e-5)
    assert abs(got[3] - ref[3]) <= 1e-5 * ref[3]
    # symmetry-aided branch, unconditionally: the same winner, its pose, its Chamfer distance
    assert got[5]["best_index"] == best
    assert abs(got[1] - ref[1]) <= 1e-5 * ref[1]
    assert torch.allclose(got[0], ref[0], atol=1e-5)


def test_sym_pose_fallback_when_segmentation_fails(ctx, rng):
    from corsair_b200 import symmetry
    # fewer points than the 100 seed samples: np.random.choice raises, sym_pose returns the
    # vanilla result with success=False (utils/symmetry.py:299-301)
    F0 = rng.standard_normal((60, 16)).astype(np.float32)
    x0 = rng.standard_normal((60, 3)).astype(np.float32)
    out = symmetry.sym_pose(F0, x0, F0.copy(), x0.copy(), 1, k_nn=5, max_corr=0.2, seed=0)
    assert out[4] is False and torch.equal(out[0], out[2]) and out[1] == out[3]


def test_irls_pose(ctx, rng):
    from corsair_b200 import _transform_estimation as te
    n = 5000
    p0 = torch.from_numpy(rng.uniform(-1, 1, (n, 3)).astype(np.float32))
    ang = torch.tensor([[0.05], [-0.03], [0.08], [0.02], [-0.01], [0.03]])
    T = port_retrieval.get_trans(ang)
    p1 = port_retrieval.update_pcd(p0, T) + 0.002 * torch.from_numpy(rng.standard_normal((n, 3)).astype(np.float32))
    p1[::17] += 0.3  # outliers for the robust weights
    w = torch.from_numpy(rng.uniform(0.5, 1.0, (n, 1)).astype(np.float32))
    for weight in (None, w):
        ref = port_retrieval.est_quad_linear_robust(p0, p1, weight)
        got = te.est_quad_linear_robust(p0, p1, weight)
        assert got.shape == (4, 4) and got.dtype == torch.float32 and got.device.type == "cpu"
        Rd = got[:3, :3].double().T @ ref[:3, :3].double()
        ang_err = np.arccos(np.clip((np.trace(Rd.numpy()) - 1) / 2, -1, 1))
        assert ang_err < 1e-4 and torch.norm(got[:3, 3] - ref[:3, 3]).item() < 1e-5
    F0 = torch.from_numpy(rng.standard_normal((300, 16)).astype(np.float32))
    F1 = torch.from_numpy(rng.standard_normal((400, 16)).astype(np.float32))
    x0 = torch.from_numpy(rng.uniform(-1, 1, (300, 3)).astype(np.float32))
    x1 = torch.from_numpy(rng.uniform(-1, 1, (400, 3)).astype(np.float32))
    tr, wt, corr = te.pose_from_features(x0, x1, F0, F1, return_corr=True)
    rtr, rwt, rcorr = port_retrieval.pose_from_features(x0, x1, F0, F1, return_corr=True)
    torch.testing.assert_close(corr.cpu(), rcorr, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(wt, rwt, rtol=1e-5, atol=1e-6)
    assert torch.allclose(tr, rtr, atol=1e-4)


def test_sym_pose_many_matches_sequential(ctx):
    """cb_sym_pose_many (one host thread, several pairs in flight on internal streams) returns exactly
    what one cb_sym_pose call per pair returns, in job order, whatever the concurrency; a chair
    (three registrations) and a table (nine) are mixed so that the longest-first reordering is used."""
    from corsair_b200 import symmetry, synth
    pairs = [synth.make_pair(200 + i, 32, "table" if i == 2 else "chair", n_raw=3000) for i in range(5)]
    def jobs():
        return [((p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"]),
                 dict(k_nn=5, max_corr=0.2, seed=3, rng=3 + i, n_hypotheses=20000)) for i, p in enumerate(pairs)]
    one = [symmetry.sym_pose(*a, **k) for a, k in jobs()]
    seq = symmetry.sym_pose_many(jobs(), workers=1)
    par = symmetry.sym_pose_many(jobs(), workers=3)
    par2 = symmetry.sym_pose_many(jobs(), workers=8)
    assert len({id(r) for r in par}) == len(pairs)
    for a, b, c, d in zip(one, seq, par, par2):
        for x in (b, c, d):
            assert np.array_equal(a[0].numpy(), x[0].numpy()) and np.array_equal(a[2].numpy(), x[2].numpy())
            assert a[1] == x[1] and a[3] == x[3] and a[4] == x[4]
    # positional k_nn / max_corr / seed (evaluation_registration passes them that way) and a RandomState
    # (sequential draw order is kept: one pair after the other)
    pos = symmetry.sym_pose_many([(a + (5, 0.2, 3), dict(rng=k["rng"], n_hypotheses=20000)) for a, k in jobs()], workers=3)
    rs = symmetry.sym_pose_many([(a, dict(k, rng=np.random.RandomState(k["rng"]))) for a, k in jobs()], workers=3)
    for a, b, c in zip(one, pos, rs):
        assert np.array_equal(a[0].numpy(), b[0].numpy()) and a[1] == b[1]
        assert np.array_equal(a[0].numpy(), c[0].numpy()) and a[1] == c[1]


def test_sym_pose_many_reports_a_failing_pair(ctx):
    """A pair the C call rejects (k_nn above the target cloud's size is caught in Python; here a
    feature width the engine does not take) surfaces as an exception, not as a silent default."""
    from corsair_b200 import symmetry, synth
    p = synth.make_pair(7, 32, "chair", n_raw=2000)
    bad = ((p["baseF"], p["xyz0"], p["posF"][:3], p["xyz1"][:3], 1), dict(k_nn=5, rng=1))
    with pytest.raises(ValueError):
        symmetry.sym_pose_many([bad, bad], workers=2)


@pytest.mark.parametrize("concurrency", [1, 2])
def test_sym_pose_many_status_tells_ok_from_never_run(ctx, concurrency):
    """h_status of cb_sym_pose_many: 0 for a pair that finished, the error of the pair that failed and
    CB_ENOTRUN (-6) for every pair that was never started because of it, whose h_out is all zero with
    h_out[34] = -1 (round-2 advice: 'ok' and 'never ran' must not look alike)."""
    import ctypes
    from corsair_b200 import _capi, symmetry, synth
    p = synth.make_pair(11, 32, "chair", n_raw=2000)
    good = ((p["baseF"], p["xyz0"], p["posF"], p["xyz1"], 1), dict(k_nn=5, rng=1, n_hypotheses=5000))
    bad = ((p["baseF"], p["xyz0"], p["posF"][:3], p["xyz1"][:3], 1), dict(k_nn=3, rng=1))  # k_nn 5 > 3 points is caught earlier
    keep, outs = [], []
    order = [good, bad, good, good, good, good]
    recs = (_capi.SymJob * len(order))()
    for i, (a, k) in enumerate(order):
        recs[i], out, _ = symmetry._job_record(ctx, a, k, lambda x: ctx.dev(x, torch.float32), keep)
        out[:] = 123.0
        outs.append(out)
    recs[1].k_nn = 5  # more neighbours than target points: the C call rejects this pair
    status = (ctypes.c_int32 * len(order))(*([77] * len(order)))
    rc = ctx.lib.cb_sym_pose_many(ctx.handle, ctypes.cast(recs, _capi.c_void_p), len(order), concurrency,
                                  ctypes.cast(status, _capi.c_void_p), ctx.stream)
    status = list(status)
    assert rc < 0 and rc != -6 and status[1] == rc, (rc, status)
    assert status[0] == 0 and outs[0][34] in (0.0, 1.0, 2.0, 3.0) and outs[0][16] != 123.0
    assert set(status) <= {0, rc, -6} and status.count(rc) == 1 and -6 in status, status
    for st, out in zip(status, outs):
        if st == -6:
            assert out[34] == -1.0 and not out[:34].any() and not out[35:].any()
        elif st == 0:
            assert out[34] >= 0.0 and np.isfinite(out[:34]).all()
    if concurrency == 1:
        assert status == [0, rc, -6, -6, -6, -6]


@pytest.mark.parametrize("category", ["chair", "table"])
def test_sym_pose_int_seed_and_streams(ctx, category, monkeypatch):
    """rng=<int> (seed rows replayed in C) equals rng=RandomState(<int>) equals np.random.seed(<int>),
    and cb_sym_pose gives the same result from cb_register_pairings' single-stream composition
    (segments -> pairings, as tests above pin against the oracle)."""
    from corsair_b200 import symmetry, synth
    from corsair_b200._capi import ptr
    p = synth.make_pair(41, 32, category, n_raw=3000)
    args = (p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"])
    kw = dict(k_nn=5, max_corr=0.2, seed=9, n_hypotheses=30000)
    a = symmetry.sym_pose(*args, rng=17, **kw)
    b = symmetry.sym_pose(*args, rng=np.random.RandomState(17), **kw)
    np.random.seed(17)
    c = symmetry.sym_pose(*args, **kw)
    for x in (b, c):
        assert torch.equal(a[0], x[0]) and torch.equal(a[2], x[2]) and a[1] == x[1] and a[3] == x[3] and a[4] == x[4]
    assert a[4] is True

    # the same pair composed from the separately tested pieces on ONE stream
    K = 4 if p["pos_sym"] >= 2 else 2
    fd = [ctx.dev(v, torch.float32) for v in args[:4]]
    rs = np.random.RandomState(17)
    A_perm, offA = symmetry._segments_device(fd[0], fd[1], K, 100, ctx, rs)
    B_perm, offB = symmetry._segments_device(fd[2], fd[3], K, 100, ctx, rs)
    off = torch.stack([offA, offB]).cpu().numpy()
    assert off[0, K + 1] >= 0 and off[1, K + 1] >= 0
    maps = symmetry._pair_maps(K, p["pos_sym"] >= 2)
    P = len(maps)
    T_all, ch = ctx.empty((1 + P, 4, 4), torch.float64), ctx.empty((1 + P,), torch.float64)
    q_off, r_off = np.ascontiguousarray(off[0, :K + 1]), np.ascontiguousarray(off[1, :K + 1])
    ctx.call("cb_register_pairings", ptr(fd[0]), ptr(fd[1]), len(p["xyz0"]), ptr(fd[2]), ptr(fd[3]), len(p["xyz1"]), 32, 5,
             ptr(A_perm), q_off.ctypes.data, ptr(B_perm), r_off.ctypes.data, K, maps.ctypes.data, P, 0.2, 9, 30000, 1,
             symmetry.CONFIDENCE, ptr(T_all), ptr(ch), None, ctx.stream)
    T_all, ch = T_all.cpu().numpy(), ch.cpu().numpy()
    best = 0
    for i in range(1, 1 + P):
        if ch[best] > ch[i]:
            best = i
    assert np.array_equal(T_all[best].astype(np.float32), a[0].numpy()) and ch[best] == a[1]
    assert np.array_equal(T_all[0].astype(np.float32), a[2].numpy()) and ch[0] == a[3]


def test_sym_pose_int_seed_small_cloud_falls_back(ctx, rng):
    from corsair_b200 import symmetry
    F0 = rng.standard_normal((60, 16)).astype(np.float32)
    x0 = rng.standard_normal((60, 3)).astype(np.float32)
    out = symmetry.sym_pose(F0, x0, F0.copy(), x0.copy(), 1, k_nn=5, max_corr=0.2, seed=0, rng=5)
    assert out[4] is False and torch.equal(out[0], out[2]) and out[1] == out[3]
