"""GPU parity: Chamfer (cb_chamfer, cb_chamfer_matrix) and retrieval top-k (cb_retrieval_topk)."""
import numpy as np
import pytest
import torch

from oracle import port, port_retrieval

pytestmark = pytest.mark.gpu


def cloud(rng, n):
    p = rng.standard_normal((n, 3)).astype(np.float32)
    if n > 1:
        p -= p.mean(0)
        p /= np.linalg.norm(p, axis=1).max()
    return p.astype(np.float32)


def rigid(rng):
    q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] *= -1
    T = np.eye(4)
    T[:3, :3], T[:3, 3] = q, rng.uniform(-1, 1, 3)
    return T


@pytest.mark.parametrize("n0,n1", [(2000, 2048), (10000, 9000), (15000, 15000), (1, 7), (300, 1)])
def test_chamfer_one_direction(ctx, rng, n0, n1):
    from corsair_b200 import preprocess
    a, b = cloud(rng, n0), cloud(rng, n1)
    T = torch.from_numpy(rigid(rng).astype(np.float32))
    ref = port.chamfer_kdtree_1direction(port.apply_transform(a, T), b)
    got = preprocess.chamfer_device(a, b, T).item()
    assert abs(got - ref) <= 1e-5 * ref
    got2 = preprocess.chamfer_kdtree_1direction(preprocess.apply_transform(a, T), b)
    assert isinstance(got2, float) and abs(got2 - ref) <= 1e-5 * ref
    got3 = preprocess.chamfer_gpu_1direction(torch.from_numpy(a), torch.from_numpy(b))
    ref3 = port.chamfer_gpu_1direction(torch.from_numpy(a), torch.from_numpy(b))
    assert abs(got3 - ref3) <= 1e-5 * ref3


def test_chamfer_two_direction_and_matrix(ctx, rng):
    from corsair_b200 import pc_dist, preprocess
    pcs = [torch.from_numpy(cloud(rng, 2000)) for _ in range(6)]
    ref = port_retrieval.compute_dist(pcs)
    got = preprocess.chamfer_gpu(pcs[0], pcs[1])
    assert abs(got.item() - ref[0, 1]) <= 1e-5 * ref[0, 1]
    assert abs(pc_dist.chamfer(pcs[2], pcs[3]).item() - ref[2, 3]) <= 1e-5 * ref[2, 3]
    table = pc_dist.compute_dist(pcs)
    assert table.shape == (6, 6) and table.dtype == np.float64
    assert np.array_equal(table, table.T) and np.all(np.diag(table) == 200)
    np.testing.assert_allclose(table, ref, rtol=1e-5)


def test_chamfer_errors(ctx):
    from corsair_b200 import preprocess
    with pytest.raises(ValueError):
        preprocess.chamfer_gpu_1direction(np.zeros((0, 3), np.float32), np.zeros((4, 3), np.float32))


def descriptors(rng, n, d=256):
    x = rng.standard_normal((n, d)).astype(np.float32)
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


@pytest.mark.parametrize("Q,L", [(1, 6778), (993, 652), (64, 830), (1000, 6778), (1000, 8436), (291, 830)])
def test_retrieval_matches_cdist_argsort(ctx, rng, Q, L):
    from corsair_b200 import retrieval
    from scipy.spatial.distance import cdist
    lib = descriptors(rng, L)
    g = rng.integers(0, L, Q)
    q = descriptors(rng, Q) * 0.3 + lib[g]
    q = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    k = max(10, int(0.1 * L))
    idx, dist = retrieval.topk(q, lib, k)
    d = cdist(q, lib)
    ref = port_retrieval.stable_argsort_rows(d)[:, :k]
    assert np.array_equal(idx, ref)
    np.testing.assert_allclose(dist, np.take_along_axis(d, ref, 1), rtol=1e-12, atol=1e-15)


def test_retrieval_eval_and_get_rank(ctx, rng):
    from corsair_b200 import retrieval
    L, Q = 652, 200
    lib, q = descriptors(rng, L), descriptors(rng, Q)
    table = np.abs(rng.standard_normal((L, L)))
    table = table + table.T
    np.fill_diagonal(table, 0)
    best = rng.integers(0, L, Q)
    pos_n = int(0.1 * L)
    got = retrieval.scan2cad_retrieval_eval(q, lib, best, table, pos_n)
    ref = port_retrieval.scan2cad_retrieval_eval(q, lib, best, table, pos_n)
    assert got["top1_predict"] == ref["top1_predict"] and got["gt"] == ref["gt"]
    assert got["precision"] == pytest.approx(ref["precision"], rel=1e-12)
    assert got["top1_error"] == pytest.approx(ref["top1_error"], rel=1e-12)
    assert np.array_equal(retrieval.get_rank(lib, top_n=10), port_retrieval.get_rank(lib, top_n=10))
    r1, r2 = retrieval.retrieval_eval(lib, 0.1, table), port_retrieval.retrieval_eval(lib, 0.1, table)
    for key in r2:
        assert r1[key] == pytest.approx(r2[key], rel=1e-9)


@pytest.mark.parametrize("L,P", [(7, 300), (8, 777), (3, 2048), (5, 1), (2, 4266)])
def test_chamfer_table_half_and_ragged_sizes(ctx, rng, L, P):
    """The packed-fp32 table kernel at point counts that are not multiples of its tiling, and the
    circulant half used by row-sharded runs: H + H^T (+ diagonal) is the full table bit for bit, every
    unordered pair appears exactly once, every row owns the same number of pairs (+-1 for even L)."""
    from corsair_b200 import pc_dist
    pts = np.stack([cloud(rng, P) if P > 1 else rng.standard_normal((1, 3)).astype(np.float32) for _ in range(L)])
    full = pc_dist.chamfer_rows_device(pts).cpu().numpy()
    ref = port_retrieval.compute_dist([torch.from_numpy(p) for p in pts])
    np.testing.assert_allclose(full, ref, rtol=1e-5)
    assert np.array_equal(full, full.T) and np.all(np.diag(full) == 200)
    half = pc_dist.chamfer_rows_device(pts, half=True).cpu().numpy()
    nz = half != 0
    assert not (nz & nz.T).any() and (nz | nz.T | np.eye(L, dtype=bool)).all()
    per_row = nz.sum(1)
    assert per_row.max() - per_row.min() <= (1 if L % 2 == 0 else 0) and per_row.sum() == L * (L - 1) // 2
    rebuilt = half + half.T
    rebuilt[np.arange(L), np.arange(L)] = 200.0
    assert np.array_equal(rebuilt, full)
    # a row block of the half equals the same rows of the whole half; a row block of the full table
    # (mirror entries inside the block written by one CTA) equals the same rows of the full table
    if L >= 4:
        assert np.array_equal(pc_dist.chamfer_rows_device(pts, 1, 3, half=True).cpu().numpy(), half[1:3])
        assert np.array_equal(pc_dist.chamfer_rows_device(pts, 1, 4).cpu().numpy(), full[1:4])


@pytest.mark.parametrize("as_numpy", [True, False])
def test_compute_dist_clouds_of_different_sizes(ctx, rng, as_numpy):
    """utils/pc_dist.py:61-99 on a list of clouds that do NOT share a point count (the all-pairs kernel
    needs equal sizes; the drop-in then runs one pair at a time): NumPy clouds and CPU tensors, against the
    reference's loop restated on the KD-tree."""
    from corsair_b200 import pc_dist
    sizes = [300, 457, 300, 1200]
    pcs = [cloud(rng, n) for n in sizes]
    got = pc_dist.compute_dist(pcs if as_numpy else [torch.from_numpy(p) for p in pcs])
    ref = port_retrieval.compute_dist([torch.from_numpy(p) for p in pcs])
    assert isinstance(got, np.ndarray) and got.shape == (4, 4)
    np.testing.assert_allclose(got, ref, rtol=1e-5)
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 200)


@pytest.mark.parametrize("L,P", [(7, 100), (5, 2049), (9, 31), (4, 4266), (6, 1)])
def test_chamfer_table_one_tile_equals_two_sweeps(ctx, rng, L, P, monkeypatch):
    """The table kernel takes both directions from ONE distance tile (like the reference's single [m, n]
    matrix with min(0) and min(1), utils/pc_dist.py:45-52: cloud i in registers, cloud j streamed, the
    streamed points' minima through warp minima and shared-memory atomicMin).  The two-sweep kernel of
    rounds 1-2 is kept behind CORSAIR_B200_CHAMFER_TWO_SWEEPS=1; both evaluate the same fp32 distances and
    add them in the same order, so full tables, half tables and row blocks agree bit for bit."""
    from corsair_b200 import pc_dist
    pts = rng.standard_normal((L, P, 3)).astype(np.float32)
    got = {}
    for two in ("0", "1"):
        monkeypatch.setenv("CORSAIR_B200_CHAMFER_TWO_SWEEPS", two)
        got[two] = [pc_dist.chamfer_rows_device(pts).cpu().numpy(),
                    pc_dist.chamfer_rows_device(pts, half=True).cpu().numpy(),
                    pc_dist.chamfer_rows_device(pts, 1, 3).cpu().numpy()]
    for a, b in zip(got["0"], got["1"]):
        assert np.array_equal(a, b)
    full = got["0"][0]
    assert np.array_equal(full, full.T) and np.all(np.diag(full) == 200)
    # ... and against the pairwise entry point (float64 re-evaluated neighbour distances): 1e-6 relative
    ref01 = pc_dist.chamfer(pts[0], pts[1]).item()
    assert abs(full[0, 1] - ref01) <= 1e-6 * ref01
