"""GPU parity: kNN engine (cb_knn) through the drop-in find_nn module vs the oracle."""
import numpy as np
import pytest
import torch

from conftest import unit_features
from oracle import port

pytestmark = pytest.mark.gpu


def _check(F0, F1, k, variant=None, ctx=None):
    from corsair_b200 import find_nn
    if variant is not None:
        ctx.set_knn_variant(variant)
    try:
        idx, d = find_nn.knn_device(F0, F1, k, return_distance=True)
    finally:
        if variant is not None:
            ctx.set_knn_variant(0)
    ref_idx, ref_d2 = port.find_knn_exact(F0, F1, k)
    idx = idx.cpu().numpy()
    assert np.array_equal(idx, ref_idx), f"{(idx != ref_idx).sum()} index mismatches"
    np.testing.assert_allclose(d.cpu().numpy(), ref_d2, rtol=1e-6, atol=1e-12)


@pytest.mark.parametrize("d", [16, 32])
@pytest.mark.parametrize("n0,n1,k", [(257, 301, 5), (4513, 3999, 5), (1000, 2000, 1), (777, 64, 12)])
def test_knn_index_exact_simt(ctx, rng, d, n0, n1, k):
    _check(unit_features(rng, n0, d), unit_features(rng, n1, d), k, variant=1, ctx=ctx)


@pytest.mark.parametrize("d", [16, 32])
def test_knn_index_exact_default_variant(ctx, rng, d):
    _check(unit_features(rng, 3000, d, clusters=40), unit_features(rng, 5000, d, clusters=40), 5)


def test_knn_duplicates_and_ties(ctx, rng):
    # exact duplicates: ties must resolve to the lower index, and the proof obligation must
    # send saturated rows to the exact fallback
    base = unit_features(rng, 40, 16)
    F1 = np.repeat(base, 30, axis=0)  # every reference appears 30 times
    F0 = base[:17].copy()
    _check(F0, F1, 5, variant=1, ctx=ctx)
    _check(F0, F1, 12, variant=1, ctx=ctx)


@pytest.mark.parametrize("d,k", [(7, 3), (24, 5), (64, 5), (32, 50), (16, 20)])
def test_knn_other_shapes(ctx, rng, d, k):
    # padded widths (d<32), wide features and large k go through padding / engine E2
    _check(unit_features(rng, 300, d), unit_features(rng, 1500, d), k)


def _l2_report(ctx, F0, F1, got, oracle):
    """Which side of a find_nn_gpu(dist_type="L2") mismatch is wrong: both against the float64 difference
    form computed here with NumPy (the round-2 driver run failed on this comparison with right indices
    and 29 distances off by 2e-4; 800 repetitions on another box — scripts/repro_find_nn_l2.py — were
    bit-equal to the float64 form, so a failure must name the side)."""
    d2 = ((F0[:, None, :].astype(np.float64) - F1[None].astype(np.float64)) ** 2).sum(2)
    two = np.sort(d2, axis=1)[:, :2]
    truth = np.sqrt(two[:, 0].astype(np.float32) + np.float32(1e-7))
    second = np.sqrt(two[:, 1].astype(np.float32) + np.float32(1e-7))
    got, oracle = got.numpy().reshape(-1), oracle.numpy().reshape(-1)
    eg, eo = np.abs(got - truth) / truth, np.abs(oracle - truth) / truth
    rows = np.nonzero((eg > 2e-6) | (eo > 2e-6))[0]
    lines = [f"fallback_rows {ctx.knn_last_fallback_rows()}, device max rel err vs float64 {eg.max():.3e} "
             f"({(eg > 2e-6).sum()} rows), torch-CPU oracle max rel err vs float64 {eo.max():.3e} ({(eo > 2e-6).sum()} rows)"]
    for r in rows[:10]:
        lines.append(f"row {r}: device {got[r]:.8f} oracle {oracle[r]:.8f} float64 {truth[r]:.8f} second-NN {second[r]:.8f}")
    return eg, eo, "\n".join(lines)


def test_reference_signatures(ctx, rng):
    from corsair_b200 import find_nn
    F0, F1 = unit_features(rng, 1203, 32), unit_features(rng, 997, 32)
    ref = port.find_knn_exact(F0, F1, 5)[0]
    inds = find_nn.find_knn_cpu(F0, F1, 5)
    assert inds.dtype == np.int64 and inds.shape == (1203, 5) and np.array_equal(inds, ref)
    inds, dists = find_nn.find_knn_cpu(F0, F1, 5, return_distance=True)
    np.testing.assert_allclose(dists, port.find_knn_cpu(F0, F1, 5, True)[1], rtol=2e-6)
    nn = find_nn.find_nn_cpu(F0, F1)
    assert nn.shape == (1203,) and np.array_equal(nn, ref[:, 0])
    # *_gpu variants: CPU tensors, flattened, [N*k,1] distances (utils/find_nn.py:86-119)
    t0, t1 = torch.from_numpy(F0), torch.from_numpy(F1)
    gi, gd = find_nn.find_knn_gpu(t0, t1, 5, nn_max_n=500, return_distance=True)
    oi, od = port.find_knn_gpu(t0, t1, 5, nn_max_n=500, return_distance=True)
    assert gi.device.type == "cpu" and gi.dtype == torch.int64 and gi.shape == (1203 * 5,)
    assert gd.shape == (1203 * 5, 1)
    assert torch.equal(gi, oi)
    torch.testing.assert_close(gd, od, rtol=1e-5, atol=1e-7)
    gi, gd = find_nn.find_nn_gpu(t0, t1, nn_max_n=500, return_distance=True, dist_type="L2")
    oi, od = port.find_nn_gpu(t0, t1, nn_max_n=500, return_distance=True, dist_type="L2")
    assert torch.equal(gi, oi) and gd.shape == (1203, 1)
    eg, eo, report = _l2_report(ctx, F0, F1, gd, od)
    assert eg.max() <= 2e-6, "device distances differ from the float64 difference form\n" + report
    assert eo.max() <= 2e-6, "the torch-CPU oracle differs from the float64 difference form on this host\n" + report
    torch.testing.assert_close(gd, od, rtol=1e-5, atol=1e-7, msg=lambda m: m + "\n" + report)
    with pytest.raises(NotImplementedError):
        find_nn.find_nn_gpu(t0, t1, dist_type="cosine")
    with pytest.raises(NotImplementedError):
        find_nn.pdist(t0, t1, dist_type="cosine")
    for dt in ("L2", "SquareL2"):
        torch.testing.assert_close(find_nn.pdist(t0[:100], t1[:90], dt), port.pdist(t0[:100], t1[:90], dt),
                                   rtol=1e-5, atol=1e-6)


def test_knn_empty_and_errors(ctx, rng):
    from corsair_b200 import find_nn
    F1 = unit_features(rng, 10, 16)
    out = find_nn.find_knn_cpu(np.zeros((0, 16), np.float32), F1, 5)
    assert out.shape == (0, 5)
    with pytest.raises(ValueError):
        find_nn.find_knn_cpu(F1, F1[:3], 5)  # k > number of references


def test_mutual_flag(ctx, rng):
    from corsair_b200 import find_nn
    F0, F1 = unit_features(rng, 500, 16), unit_features(rng, 400, 16)
    idx, mut = find_nn.knn_device(F0, F1, 1, mutual=True)
    fwd = port.find_knn_exact(F0, F1, 1)[0][:, 0]
    rev = port.find_knn_exact(F1, F0, 1)[0][:, 0]
    assert np.array_equal(mut.cpu().numpy().astype(bool), rev[fwd] == np.arange(500))
