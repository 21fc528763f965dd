"""GPU parity: correspondence RANSAC (cb_ransac) vs the restated Open3D algorithm."""
import numpy as np
import pytest

from oracle import ransac_c, ransac_np

pytestmark = pytest.mark.gpu


def make_problem(rng, M, inlier=0.5, noise=0.01):
    src = rng.uniform(-1, 1, (M, 3)).astype(np.float32)
    a = rng.uniform(-np.pi, np.pi, 3)
    Rx = np.array([[1, 0, 0], [0, np.cos(a[0]), -np.sin(a[0])], [0, np.sin(a[0]), np.cos(a[0])]])
    Ry = np.array([[np.cos(a[1]), 0, np.sin(a[1])], [0, 1, 0], [-np.sin(a[1]), 0, np.cos(a[1])]])
    Rz = np.array([[np.cos(a[2]), -np.sin(a[2]), 0], [np.sin(a[2]), np.cos(a[2]), 0], [0, 0, 1]])
    R, t = Rz @ Ry @ Rx, rng.uniform(-1, 1, 3)
    tgt = src @ R.T + t + noise * rng.standard_normal((M, 3))
    out = rng.random(M) > inlier
    tgt[out] = rng.uniform(-2, 2, (int(out.sum()), 3))
    return src, tgt.astype(np.float32), R, t


def pose_err(Ta, Tb):
    Rd = Ta[:3, :3].T @ Tb[:3, :3]
    ang = np.arccos(np.clip((np.trace(Rd) - 1) / 2, -1, 1))
    return ang, np.linalg.norm(Ta[:3, 3] - Tb[:3, 3])


@pytest.mark.parametrize("M,H,thr", [(13000, 64, 0.2), (5003, 4096, 0.2), (50000, 4096, 0.4), (999, 1, 0.2)])
def test_counts_bit_exact_fixed_sets(ctx, rng, M, H, thr):
    from corsair_b200 import eval_pose
    src, tgt, _, _ = make_problem(rng, M)
    ms = ransac_np.draw_minimal_sets(M, H, 10, seed=7)
    got = eval_pose.ransac_device(src, tgt, thr, minimal_sets=ms, adaptive=False, return_all=True)
    ref = ransac_c.ransac(src, tgt, thr, ms, adaptive=False, return_all=True)
    counts = got["counts"].cpu().numpy()
    assert np.array_equal(counts, ref["counts"]), "per-hypothesis inlier counts differ"
    # sums come from the fp32 evaluation (counts are the exact part of the contract)
    np.testing.assert_allclose(got["sums"].cpu().numpy(), ref["sums"], rtol=5e-5, atol=1e-9)
    st = got["stats"].cpu().numpy()
    assert int(st[0]) == ref["best_h"] and int(st[1]) == ref["inliers"] and int(st[3]) == ref["n_evaluated"]
    ang, tr = pose_err(got["T"].cpu().numpy(), ref["T"])
    assert ang < 1e-4 and tr < 1e-5


def test_device_draw_matches_oracle_rule(ctx):
    from corsair_b200 import eval_pose
    for M, H, seed in [(34145, 1000, 31), (11, 7, 0), (250000, 4096, 2**40 + 5)]:
        assert np.array_equal(eval_pose.draw_minimal_sets(M, H, 10, seed), ransac_np.draw_minimal_sets(M, H, 10, seed))


@pytest.mark.parametrize("inlier", [0.3, 0.6, 0.9])
def test_adaptive_termination_matches_sequential(ctx, rng, inlier):
    from corsair_b200 import eval_pose
    M, H = 20000, 100000
    src, tgt, R, t = make_problem(rng, M, inlier=inlier)
    got = eval_pose.ransac_device(src, tgt, 0.2, seed=5, n_hypotheses=H, adaptive=True)
    ms = ransac_np.draw_minimal_sets(M, H, 10, seed=5)
    ref = ransac_c.ransac(src, tgt, 0.2, ms, adaptive=True)
    st = got["stats"].cpu().numpy()
    assert int(st[0]) == ref["best_h"] and int(st[1]) == ref["inliers"]
    assert int(st[3]) == ref["n_evaluated"]
    ang, tr = pose_err(got["T"].cpu().numpy(), ref["T"])
    assert ang < 1e-4 and tr < 1e-5
    Tgt = np.eye(4)
    Tgt[:3, :3], Tgt[:3, 3] = R, t
    ang, tr = pose_err(got["T"].cpu().numpy(), Tgt)
    if inlier >= 0.6:  # and it actually registers the clouds (no refit on inliers, like Open3D)
        assert ang < 0.05 and tr < 0.05


def test_numpy_oracle_small(ctx, rng):
    from corsair_b200 import eval_pose
    src, tgt, _, _ = make_problem(rng, 3000)
    ms = ransac_np.draw_minimal_sets(3000, 50, 10, seed=1)
    got = eval_pose.ransac_device(src, tgt, 0.2, minimal_sets=ms, adaptive=False, return_all=True)
    ref = ransac_np.ransac(src, tgt, 0.2, ms, adaptive=False, return_all=True)
    assert np.array_equal(got["counts"].cpu().numpy(), ref["counts"])
    assert int(got["stats"][0].item()) == ref["best_h"]


def test_degenerate_sets_and_requests(ctx, rng):
    from corsair_b200 import eval_pose
    src, tgt, _, _ = make_problem(rng, 2000)
    ms = ransac_np.draw_minimal_sets(2000, 32, 10, seed=3)
    ms[0, :] = 5                      # one repeated point: rank-0 covariance
    ms[1, :] = [1, 1, 1, 1, 1, 2, 2, 2, 2, 2]  # two distinct points: rank 1
    src2 = src.copy()
    src2[ms[2]] = np.linspace(0, 1, 10)[:, None] * np.array([[1.0, 2.0, 3.0]], np.float32)  # collinear
    got = eval_pose.ransac_device(src2, tgt, 0.2, minimal_sets=ms, adaptive=False, return_all=True)
    ref = ransac_c.ransac(src2, tgt, 0.2, ms, adaptive=False, return_all=True)
    c = got["counts"].cpu().numpy()
    assert np.array_equal(c[3:], ref["counts"][3:])
    T = got["T"].cpu().numpy()
    assert np.allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-9) and np.isclose(np.linalg.det(T[:3, :3]), 1)
    # fewer correspondences than ransac_n, or a non-positive threshold: identity, nothing evaluated
    for kwargs in (dict(src=src[:5], tgt=tgt[:5], thr=0.2), dict(src=src, tgt=tgt, thr=0.0)):
        g = eval_pose.ransac_device(kwargs["src"], kwargs["tgt"], kwargs["thr"], n_hypotheses=64, adaptive=False)
        assert np.array_equal(g["T"].cpu().numpy(), np.eye(4)) and g["stats"][3].item() == 0


def test_registration_based_on_corr_signature(ctx, rng):
    from corsair_b200 import eval_pose
    src, tgt, R, t = make_problem(rng, 8000, inlier=0.7)
    T = eval_pose.registration_based_on_corr(src, tgt, 0.2, seed=0)
    assert T.shape == (4, 4) and T.dtype == np.float64
    Tref = __import__("oracle.port", fromlist=["x"]).registration_based_on_corr(src, tgt, 0.2, seed=0)
    ang, tr = pose_err(T, Tref)
    assert ang < 1e-4 and tr < 1e-5


def test_fit_minimal_matches_umeyama(ctx, rng):
    import torch
    from corsair_b200 import _capi
    src, tgt, _, _ = make_problem(rng, 500, inlier=1.0)
    ms = ransac_np.draw_minimal_sets(500, 200, 10, seed=9)
    c = _capi.get_context()
    s, t, m = c.dev(src, torch.float32), c.dev(tgt, torch.float32), c.dev(ms, torch.int32)
    out = c.empty((200, 4, 4), torch.float64)
    c.call("cb_fit_minimal", _capi.ptr(s), _capi.ptr(t), 500, _capi.ptr(m), 200, 10, _capi.ptr(out), c.stream)
    out = out.cpu().numpy()
    for h in range(200):
        ref = ransac_np.umeyama_noscale(src[ms[h]].astype(np.float64), tgt[ms[h]].astype(np.float64))
        np.testing.assert_allclose(out[h], ref, atol=1e-10)
