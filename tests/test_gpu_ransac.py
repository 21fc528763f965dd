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


@pytest.fixture(params=[2, 1], ids=["tcgen05", "simt"])
def variant(request, ctx):
    ctx.set_ransac_variant(request.param)
    assert ctx.ransac_variant() == ("simt" if request.param == 1 else "tcgen05")
    yield request.param
    ctx.set_ransac_variant(0)


@pytest.mark.parametrize("M,H,thr", [(13000, 64, 0.2), (5003, 4096, 0.2), (50000, 4096, 0.4), (999, 1, 0.2),
                                     (34145, 9000, 0.2), (127, 300, 0.2), (128, 256, 0.3), (250000, 512, 0.2),
                                     (13000, 100000, 0.2), (250000, 4096, 0.2)])  # the reference budget; C5
def test_counts_bit_exact_fixed_sets(ctx, rng, variant, M, H, thr):
    from corsair_b200 import eval_pose
    src, tgt, _, _ = make_problem(rng, M)
    ms = ransac_np.draw_minimal_sets(M, H, 10, seed=7)
    got = eval_pose.ransac_device(src, tgt, thr, minimal_sets=ms, adaptive=False, return_all=True)
    ref = ransac_c.ransac(src, tgt, thr, ms, adaptive=False, return_all=True)
    counts = got["counts"].cpu().numpy()
    assert np.array_equal(counts, ref["counts"]), "per-hypothesis inlier counts differ"
    # streamed sums are approximate (fp32 / split-bf16 tensor-core values): counts are the exact
    # part of the contract; the winner's sum is recomputed in float64 and must be exact
    if variant == 1:
        np.testing.assert_allclose(got["sums"].cpu().numpy(), ref["sums"], rtol=5e-5, atol=1e-9)
    else:  # every tensor-core value is within its band (< 1e-3 here) of the float64 one
        assert np.all(np.abs(got["sums"].cpu().numpy() - ref["sums"]) <= 1e-4 * ref["counts"] + 1e-9)
    st = got["stats"].cpu().numpy()
    assert abs(st[2] - ref["sum_d2"]) <= 1e-11 * max(1.0, ref["sum_d2"])
    assert int(st[0]) == ref["best_h"] and int(st[1]) == ref["inliers"] and int(st[3]) == ref["n_evaluated"]
    ang, tr = pose_err(got["T"].cpu().numpy(), ref["T"])
    assert ang < 1e-4 and tr < 1e-5


@pytest.mark.parametrize("offset", [0.0, 7.5])
def test_tensor_core_values_within_band(ctx, rng, offset):
    """The raw tcgen05 value U = d^2 - thr^2 of every evaluation is within the per-hypothesis
    rounding band of the float64 value (the premise of the exact in-band re-decision), also for
    clouds far from the origin (the features are centred on the correspondence means)."""
    import torch
    from corsair_b200 import eval_pose
    M, H, thr = 3001, 700, 0.2
    src, tgt, _, _ = make_problem(rng, M)
    src = (src + offset).astype(np.float32)
    tgt = (tgt - 2 * offset).astype(np.float32)
    ms = ransac_np.draw_minimal_sets(M, H, 10, seed=11)
    U = torch.full((H, M), float("nan"), dtype=torch.float32, device="cuda")
    ctx.set_ransac_variant(2)
    ctx.ransac_debug_dump(U)
    try:
        got = eval_pose.ransac_device(src, tgt, thr, minimal_sets=ms, adaptive=False, return_all=True)
        torch.cuda.synchronize()
    finally:
        ctx.ransac_debug_dump(None)
        ctx.set_ransac_variant(0)
    T = ransac_c.fit_all(src, tgt, ms)
    s64, q64 = src.astype(np.float64), tgt.astype(np.float64)
    cs, cq = s64.mean(0), q64.mean(0)
    smax, tmax = np.linalg.norm(s64 - cs, axis=1).max(), np.linalg.norm(q64 - cq, axis=1).max()
    U = U.cpu().numpy().astype(np.float64)
    worst = 0.0
    for h in range(H):
        e = s64 @ T[h, :3, :3].T + T[h, :3, 3] - q64
        u = (e * e).sum(1) - thr * thr
        tp = T[h, :3, 3] + T[h, :3, :3] @ cs - cq
        tn = np.linalg.norm(tp)
        bmag = 2 * np.sqrt(3) * smax * tmax + 2 * tn * (smax + tmax)
        mag = bmag + smax ** 2 + tmax ** 2 + abs(tn * tn - thr * thr)
        band = 1.1 * 2.0 ** -25 * bmag + 12 * 2.0 ** -23 * mag
        worst = max(worst, np.abs(U[h] - u).max() / band)
    print(f"max |U - u| / band = {worst:.3f}")
    assert worst < 1.0
    # diagnostic: the tensor core's fp32 accumulation error alone, against the float64 sum of the
    # same split-bf16 products (in units of 2^-23 * sum of term magnitudes)
    def bf16(x):
        b = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
        b = ((b + 0x7FFF + ((b >> 16) & 1)) >> 16) << 16
        return b.astype(np.uint32).view(np.float32).astype(np.float64)
    def split3(x):
        a = bf16(x); b = bf16(x - a)
        return a, b, bf16(x - a - b)
    sc, qc = s64 - cs, q64 - cq
    phi = np.concatenate([(qc[:, :, None] * sc[:, None, :]).reshape(M, 9), sc, qc], 1)
    pa, pb, pc = split3(phi)
    n3 = split3((sc * sc).sum(1) + (qc * qc).sum(1))
    acc_worst = 0.0
    for h in range(0, H, 7):
        R, tp = T[h, :3, :3], T[h, :3, 3] + T[h, :3, :3] @ cs - cq
        wv = np.concatenate([(-2 * R).reshape(9), 2 * (R.T @ tp), -2 * tp])
        wa, wb, wc = split3(wv)
        c3 = split3(tp @ tp - thr * thr)
        emu = pa @ wa + pb @ wa + pa @ wb + pc @ wa + pa @ wc + pb @ wb + sum(n3) + sum(c3)
        mag = np.abs(pa) @ np.abs(wa) + sum(np.abs(n3)) + sum(np.abs(c3))
        acc_worst = max(acc_worst, (np.abs(U[h] - emu) / (2.0 ** -23 * mag)).max())
    print(f"tensor-core accumulation error: max {acc_worst:.2f} x 2^-23 x sum|terms|")
    assert acc_worst < 6.0  # the band allows 12
    ref = ransac_c.ransac(src, tgt, thr, ms, adaptive=False, return_all=True)
    assert np.array_equal(got["counts"].cpu().numpy(), ref["counts"])


def test_device_draw_matches_oracle_rule(ctx):
    from corsair_b200 import eval_pose
    for M, H, seed in [(34145, 1000, 31), (11, 7, 0), (250000, 4096, 2**40 + 5)]:
        assert np.array_equal(eval_pose.draw_minimal_sets(M, H, 10, seed), ransac_np.draw_minimal_sets(M, H, 10, seed))


@pytest.mark.parametrize("inlier", [0.3, 0.6, 0.9])
def test_adaptive_termination_matches_sequential(ctx, rng, variant, inlier):
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


def test_degenerate_sets_and_requests(ctx, rng, variant):
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
