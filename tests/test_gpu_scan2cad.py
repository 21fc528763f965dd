"""GPU parity: shape-estimation branch of evaluation-scan2cad.py and the precompute cross table
 against reference-generated golden vectors and the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import port_scan2cad as sc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "scan2cad.npz")


def test_golden(ctx):
    from corsair_b200 import scan2cad_eval as se
    g = np.load(GOLDEN)
    table = se.chamfer_cross_table(g["gt"], g["recon"])
    assert table.shape == (5, 4) and table.dtype == np.float64
    np.testing.assert_allclose(table, g["table"], rtol=1e-5)  # Chamfer contract: 1e-5 relative
    top, ret = se.rerank_by_chamfer(g["desc"], g["lib"], g["ctab"], g["best"], int(g["use_best"]))
    assert np.array_equal(top, g["topN"]) and np.array_equal(ret, g["retrieved"])
    one = se.evaluate_retrieval(g["gt"][1], g["recon"][3])
    assert abs(one - g["table"][1, 3]) <= 1e-5 * g["table"][1, 3]
    many = se.evaluate_retrieval_many([(g["gt"][i], g["recon"][j]) for i, j in ((0, 0), (4, 2), (2, 1))])
    np.testing.assert_allclose(many, [g["table"][0, 0], g["table"][4, 2], g["table"][2, 1]], rtol=1e-5)
    assert se.evaluate_retrieval_many([]) == []


def test_full_size_clouds_and_rerank_vs_oracle(ctx, rng):
    """15 000-point clouds (the app's size) and a Scan2CAD-sized re-rank (use_best = 30)."""
    from corsair_b200 import scan2cad_eval as se
    a = rng.standard_normal((2, 15000, 3)).astype(np.float32)
    b = (a[::-1] * 0.97 + 0.01 * rng.standard_normal(a.shape)).astype(np.float32)
    np.testing.assert_allclose(se.chamfer_cross_table(a, b), sc.chamfer_cross_table(a.astype(np.float64), b.astype(np.float64)),
                               rtol=1e-5)
    lib = rng.standard_normal((700, 256)).astype(np.float32)
    lib /= np.linalg.norm(lib, axis=1, keepdims=True)
    best = rng.integers(0, 700, 90)
    desc = (lib[best] + 0.05 * rng.standard_normal((90, 256))).astype(np.float32)
    ctab = rng.random((700, 700))
    top, ret = se.rerank_by_chamfer(desc, lib, ctab, best, 30)
    rtop, rret = sc.rerank_by_chamfer(desc.astype(np.float64), lib.astype(np.float64), ctab, best, 30)
    assert np.array_equal(top, rtop) and np.array_equal(ret, rret)


def test_cross_table_errors(ctx):
    from corsair_b200 import scan2cad_eval as se
    with pytest.raises(ValueError):
        se.chamfer_cross_table(np.zeros((3, 4), np.float32), np.zeros((1, 4, 3), np.float32))
