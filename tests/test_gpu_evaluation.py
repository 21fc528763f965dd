"""GPU parity: the registration sweep of evaluation.py:297-334 through evaluation_registration
 against the explicit sym_pose + eval_pose loop."""
import numpy as np
import pytest
import torch

from corsair_b200 import synth

pytestmark = pytest.mark.gpu


def _scene(n_q=5, n_lib=4):
    lib = [synth.make_pair(700 + j, 32, "table" if j % 2 else "chair", n_raw=3000) for j in range(n_lib)]
    lib_outputs = [torch.from_numpy(p["posF"]) for p in lib]
    lib_origins = [torch.from_numpy(p["xyz1"]) for p in lib]
    lib_Ts = [torch.eye(4) for _ in lib]
    sym_label = [p["pos_sym"] for p in lib]
    pos_idx = [i % n_lib for i in range(n_q)]
    base_outputs = [torch.from_numpy(lib[j]["baseF"]) for j in pos_idx]
    base_origins = [torch.from_numpy(lib[j]["xyz0"]) for j in pos_idx]
    base_Ts = [torch.from_numpy(np.linalg.inv(lib[j]["T_gt"]).astype(np.float32)) for j in pos_idx]
    return base_outputs, base_origins, base_Ts, lib_outputs, lib_origins, lib_Ts, sym_label, pos_idx


def _same(a, b):
    for k in a:
        for x, y in zip(a[k], b[k]):
            assert np.array_equal(np.asarray(x), np.asarray(y)), k


def test_register_queries(ctx, tmp_path):
    import json
    import os
    from corsair_b200 import eval_pose, evaluation_registration as er, symmetry
    sc = _scene()
    bo, bx, bT, lo, lx, lT, sym, pos = sc
    # the reference's loop: global NumPy stream, one pair after the other
    np.random.seed(31)
    got = er.register_queries(*sc, k_nn=5, max_corr=0.2)
    np.random.seed(31)
    for i in range(len(bo)):
        p = pos[i]
        r = symmetry.sym_pose(bo[i].numpy(), bx[i].numpy(), lo[p].numpy(), lx[p].numpy(), sym[p], 5, 0.2)
        assert np.array_equal(r[0].numpy(), got["Ts_est_best"][i]) and np.array_equal(r[2].numpy(), got["Ts_est_ransac"][i])
        assert r[1] == got["chamfer_dist_sym"][i] and r[3] == got["chamfer_dist_ransac"][i] and r[4] == got["sym_ransac_success"][i]
        t, rr = eval_pose.eval_pose(r[0], bT[i].numpy(), lT[p].numpy(), axis_symmetry=sym[p])
        assert t == got["t_losses_sym"][i] and rr == got["r_losses_sym"][i]
    assert set(got) == set(er.KEYS) and all(len(v) == len(bo) for v in got.values())
    # per-query seeds: threaded == sequential == device-resident inputs
    seeds = [11 + i for i in range(len(bo))]
    a = er.register_queries(*sc, k_nn=5, max_corr=0.2, seeds=seeds)
    b = er.register_queries(*sc, k_nn=5, max_corr=0.2, seeds=seeds, workers=3)
    dev = [[t.cuda() for t in sc[0]], [t.cuda() for t in sc[1]], sc[2], [t.cuda() for t in sc[3]],
           [t.cuda() for t in sc[4]], sc[5], sc[6], sc[7]]
    c = er.register_queries(*dev, k_nn=5, max_corr=0.2, seeds=seeds, workers=3, on_device=True)
    _same(a, b)
    _same(a, c)
    with pytest.raises(ValueError):
        er.register_queries(*sc, workers=2)
    s, text = er.summarize(a)
    assert 0.0 <= s["rte_015_sym"] <= 1.0 and "rte 0.02" in text
    # the files this sweep writes have the dtype / rank / trailing shape of the caches the reference's own
    # evaluation.py wrote (data/cache_ret, recorded in tests/golden/cache_format.json)
    fmt = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cache_format.json")))["files"]
    er.save_results(str(tmp_path), "chair", False, a)
    for k in er.KEYS:
        arr = np.load(tmp_path / f"{k}_chair_gt.npy")
        dtype, ndim, trailing, _ = fmt[f"{k}_chair_gt.npy"]
        assert (str(arr.dtype), arr.ndim, list(arr.shape[1:])) == (dtype, ndim, trailing), k
