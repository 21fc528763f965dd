"""CPU: the oracle (oracle/port*.py) against the golden vectors produced by the reference's own
code (tests/golden/make_golden.py).  This is what pins the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import port, port_retrieval, port_symmetry, ransac_c, ransac_np

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name + ".npz"), allow_pickle=False)


def f32_close(x, y):
    """float32 results of torch's CPU broadcast-and-sum (the reference's pdist) agree with the golden
    file to the last bits only on the host the file was made on: the reduction order over D follows
    the vector width of the CPU (AVX2 / AVX-512 hosts differ by 1 ulp in a few entries).  Indices are
    still compared exactly; values within 4 ulp of float32."""
    return np.allclose(x, y, rtol=4 * np.finfo(np.float32).eps, atol=0)


@pytest.mark.parametrize("d", [16, 32])
def test_find_nn(d):
    g = load("find_nn")
    F0, F1 = g[f"F0_{d}"], g[f"F1_{d}"]
    idx, d2 = port.find_knn_exact(F0, F1, 5)
    assert np.array_equal(idx, g[f"knn_cpu_idx_{d}"])          # scipy KD-tree == canonical exact kNN
    np.testing.assert_allclose(np.sqrt(d2), g[f"knn_cpu_dist_{d}"], rtol=1e-12)
    assert np.array_equal(port.find_knn_cpu(F0, F1, 5), g[f"knn_cpu_idx_{d}"])
    assert np.array_equal(port.find_nn_cpu(F0, F1), g[f"nn_cpu_idx_{d}"])
    t0, t1 = torch.from_numpy(F0), torch.from_numpy(F1)
    gi, gd = port.find_knn_gpu(t0, t1, 5, nn_max_n=300, return_distance=True)
    assert np.array_equal(gi.numpy(), g[f"knn_gpu_idx_{d}"]) and f32_close(gd.numpy(), g[f"knn_gpu_dist_{d}"])
    ni, nd = port.find_nn_gpu(t0, t1, nn_max_n=300, return_distance=True, dist_type="L2")
    assert np.array_equal(ni.numpy(), g[f"nn_gpu_idx_{d}"]) and f32_close(nd.numpy(), g[f"nn_gpu_dist_{d}"])
    assert f32_close(port.pdist(t0[:50], t1[:60], "L2").numpy(), g[f"pdist_L2_{d}"])
    # the fp32 broadcast top-k of find_knn_gpu and the float64 KD-tree agree on these inputs
    assert np.array_equal(g[f"knn_gpu_idx_{d}"].reshape(-1, 5), g[f"knn_cpu_idx_{d}"])
    with pytest.raises(NotImplementedError):
        port.pdist(t0, t1, "cosine")


def test_find_kcorr():
    g = load("find_nn")
    i0, i1 = port.find_kcorr(g["F0_16"], g["F1_16"], k=5)
    assert np.array_equal(i0, g["kcorr_idx0"]) and np.array_equal(i1, g["kcorr_idx1"])


def test_chamfer():
    g = load("chamfer")
    a, b, c, T = g["a"], g["b"], g["c"], torch.from_numpy(g["T"])
    ta = port.apply_transform(a, T)
    np.testing.assert_allclose(ta, g["apply_transform"], rtol=0, atol=1e-15)
    assert port.chamfer_kdtree_1direction(ta, b) == pytest.approx(float(g["kdtree_1dir"]), rel=1e-14)
    pa, pb = torch.from_numpy(a), torch.from_numpy(b)
    assert port.chamfer_gpu_1direction(pa, pb) == pytest.approx(float(g["gpu_1dir"]), rel=1e-6)
    assert port.chamfer_gpu(pa, pb).item() == pytest.approx(float(g["gpu_2dir"]), rel=1e-6)
    table = port_retrieval.compute_dist([torch.from_numpy(x) for x in (a, b, c)])
    np.testing.assert_allclose(table, g["table"], rtol=1e-6)
    assert np.all(np.diag(g["table"]) == 200)  # eye*100 doubled by table += table.T


def test_retrieval():
    g = load("retrieval")
    st = port_retrieval.scan2cad_retrieval_eval(g["q"], g["lib"], g["best_match"], g["table"], 30)
    assert np.array_equal(np.array(st["top1_predict"]), g["top1_predict"]) and np.array_equal(np.array(st["gt"]), g["gt"])
    assert st["precision"] == pytest.approx(float(g["precision"]), rel=1e-12)
    assert st["top1_error"] == pytest.approx(float(g["top1_error"]), rel=1e-12)
    assert np.array_equal(port_retrieval.get_rank(g["lib"], 10), g["get_rank"])
    ev = port_retrieval.retrieval_eval(g["lib"], 0.1, g["table"])
    assert ev["mAP"] == pytest.approx(float(g["eval_mAP"]), rel=1e-9)
    assert ev["percision"] == pytest.approx(float(g["eval_percision"]), rel=1e-12)
    assert ev["top1_error"] == pytest.approx(float(g["eval_top1"]), rel=1e-12)
    m = port_retrieval.compute_mAP(g["lib"], g["labels"])
    assert m[0] == pytest.approx(float(g["mAP_instance"]), rel=1e-9)
    np.testing.assert_allclose(np.array(m[1]), g["mAP_class"], rtol=1e-9)


def test_irls():
    g = load("irls")
    p0, p1, w = (torch.from_numpy(g[k]) for k in ("p0", "p1", "w"))
    x = torch.tensor([[0.04], [-0.02], [0.06], [0.02], [-0.03], [0.01]])
    assert np.allclose(port_retrieval.get_trans(x).numpy(), g["get_trans"], atol=1e-7)
    assert np.allclose(port_retrieval.est_quad_linear_robust(p0, p1).numpy(), g["T_unweighted"], atol=1e-5)
    assert np.allclose(port_retrieval.est_quad_linear_robust(p0, p1, w).numpy(), g["T_weighted"], atol=1e-5)


@pytest.mark.parametrize("cat,K", [("chair", 2), ("table", 4)])
def test_symmetric_cut4(cat, K):
    g = load("symmetry")
    np.random.seed(31)
    masks = port_symmetry.symmetric_cut4(g[f"{cat}_F"], g[f"{cat}_xyz"], K)
    assert np.array_equal(np.stack(masks), g[f"{cat}_masks"])


def test_eval_pose_and_sym_pose():
    g = load("symmetry")
    Te = torch.from_numpy(g["T_est"])
    np.testing.assert_allclose(np.array(port.eval_pose(Te, g["T0"], g["T1"], 1)), g["eval_pose_sym1"], rtol=1e-6)
    np.testing.assert_allclose(np.array(port.eval_pose(Te, g["T0"], g["T1"], 4)), g["eval_pose_sym4"], rtol=1e-6)
    assert port.eval_pose(Te, g["T0"], g["T1"], 0) == (np.inf, np.inf)

    def reg(src, tgt, max_corr, seed):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        return ransac_c.ransac(src, tgt, max_corr, ms, adaptive=True)["T"]

    np.random.seed(31)
    Tb, cb, Tr, cr, ok = port_symmetry.sym_pose(g["sp_baseF"], g["sp_xyz0"], g["sp_posF"], g["sp_xyz1"], 1, k_nn=5,
                                                max_corr=0.2, seed=31, registration=reg)
    assert ok == bool(g["sp_ok"])
    assert np.allclose(Tr.numpy(), g["sp_T_ransac"], atol=1e-6) and cr == pytest.approx(float(g["sp_ch_ransac"]), rel=1e-9)
    assert np.allclose(Tb.numpy(), g["sp_T_best"], atol=1e-6) and cb == pytest.approx(float(g["sp_ch_best"]), rel=1e-9)


def test_ransac_restatements_agree():
    g = load("ransac_restated")
    src, tgt, ms = g["src"], g["tgt"], g["minimal_sets"]
    assert np.array_equal(ransac_np.draw_minimal_sets(2000, 256, 10, seed=3), ms)
    for impl in (ransac_np, ransac_c):
        r = impl.ransac(src, tgt, 0.2, ms, adaptive=False, return_all=True)
        assert np.array_equal(np.asarray(r["counts"]), g["counts"])
        np.testing.assert_allclose(r["sums"], g["sums"], rtol=1e-12)
        assert r["best_h"] == int(g["best_h"]) and r["inliers"] == int(g["inliers"])
        np.testing.assert_allclose(r["T"], g["T"], atol=1e-12)
    a = ransac_np.ransac(src, tgt, 0.2, ms, adaptive=True)
    b = ransac_c.ransac(src, tgt, 0.2, ms, adaptive=True)
    assert a["best_h"] == b["best_h"] and a["n_evaluated"] == b["n_evaluated"] and a["inliers"] == b["inliers"]
    # known-answer property: a noiseless rigid motion is recovered exactly and every point is an inlier
    T = np.eye(4)
    T[:3, :3] = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]])
    T[:3, 3] = [0.3, -0.2, 0.5]
    s = src.astype(np.float64)
    t = s @ T[:3, :3].T + T[:3, 3]
    r = ransac_np.ransac(s, t, 0.05, ms[:4], adaptive=False)
    assert r["inliers"] == len(s) and np.allclose(r["T"], T, atol=1e-12)
    # degenerate requests return Open3D's default result
    r = ransac_c.ransac(s[:5], t[:5], 0.2, ms[:4] % 5, adaptive=False)
    assert r["best_h"] == -1 and np.array_equal(r["T"], np.eye(4))


def test_shapenet_symmetry_label():
    """oracle/port_shapenet.py against vectors produced by the reference's own source
    (evaluation-shapenet.py:122-154, tests/golden/make_shapenet_golden.py)."""
    from oracle import port_shapenet as ps
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "shapenet.npz"))
    n = len(g["labels"])
    assert sorted(set(g["labels"].tolist())) == [1, 2, 3, 4, 6, 8, 12]
    for c in range(n):
        pc = g[f"pc{c}"]
        assert ps.get_symmetry_label(pc, float(g["threshold"])) == g["labels"][c]
        errs = [e for sym in ps.SYM_ORDERS for e in ps.rotation_errors(sym, pc)]
        np.testing.assert_allclose(errs, g["errors"][c], rtol=0, atol=1e-12)
        assert ps.test_symmetry_label(4, pc, 0.05) == bool(g["test_label4_thr005"][c])


def test_scan2cad_shape_branch():
    """oracle/port_scan2cad.py against vectors made with the reference's chamfer_kdtree_1direction and
    the app's own cdist / argsort / argmin statements (tests/golden/make_scan2cad_golden.py)."""
    from oracle import port_scan2cad as sc
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "scan2cad.npz"))
    np.testing.assert_allclose(sc.chamfer_cross_table(g["gt"], g["recon"]), g["table"], rtol=0, atol=1e-12)
    top, ret = sc.rerank_by_chamfer(g["desc"], g["lib"], g["ctab"], g["best"], int(g["use_best"]))
    assert np.array_equal(top, g["topN"]) and np.array_equal(ret, g["retrieved"])


def test_evaluation_cache_and_metric_tables(tmp_path):
    """evaluation_registration: the nine cache files (evaluation.py:390-441 names and shapes) round-trip,
    and the RTE / RRE tables equal the restated closures of :336-348 (and hand-computed values)."""
    from corsair_b200 import evaluation_registration as er
    from oracle import port_evaluation as pe
    rng = np.random.default_rng(3)
    n = 23
    res = {
        "Ts_est_ransac": [rng.standard_normal((4, 4)).astype(np.float32) for _ in range(n)],
        "Ts_est_best": [rng.standard_normal((4, 4)).astype(np.float32) for _ in range(n)],
        "t_losses_ransac": list(rng.random(n) * 0.2), "t_losses_sym": list(rng.random(n) * 0.1),
        "r_losses_ransac": list(rng.random(n) * 1.0), "r_losses_sym": list(rng.random(n) * 0.3),
        "sym_ransac_success": [bool(b) for b in rng.random(n) < 0.8],
        "chamfer_dist_ransac": list(rng.random(n) * 0.05), "chamfer_dist_sym": list(rng.random(n) * 0.04),
    }
    assert er.load_results(str(tmp_path), "chair", True) is None
    er.save_results(str(tmp_path), "chair", True, res)
    assert sorted(os.listdir(tmp_path)) == sorted(f"{k}_chair_top1.npy" for k in er.KEYS)
    assert np.load(tmp_path / "Ts_est_best_chair_top1.npy").shape == (n, 16)
    assert er.load_results(str(tmp_path), "chair", False) is None  # the _gt files are a different cache
    back = er.load_results(str(tmp_path), "chair", True)
    for k in er.KEYS:
        assert np.array_equal(np.asarray(back[k]), np.asarray(res[k])), k
    assert back["Ts_est_ransac"][0].shape == (4, 4)
    s, text = er.summarize(back)
    assert (s["rte_002_ransac"], s["rte_005_ransac"], s["rte_010_ransac"], s["rte_015_ransac"]) == pe.compute_rte(res["t_losses_ransac"])
    assert (s["rre_005_sym"], s["rre_015_sym"], s["rre_045_sym"]) == pe.compute_rre(res["r_losses_sym"])
    assert s["sym_success_rate"] == np.mean(res["sym_ransac_success"]) and s["t_loss_sym"] == np.mean(res["t_losses_sym"])
    assert "vanilla ransac:" in text and "sym ransac:" in text and f"sym success rate: {s['sym_success_rate']}" in text
    # hand-computed: thresholds are inclusive
    assert pe.compute_rte([0.02, 0.05, 0.2, 0.01]) == (0.5, 0.75, 0.75, 0.75)
    assert pe.compute_rre(np.deg2rad([4.0, 20.0, 50.0, 10.0])) == (0.25, 0.5, 0.75)


def test_cache_files_match_the_reference_held_cache(tmp_path):
    """the reference tree holds caches its own evaluation.py:420-441 wrote on the real
    data (data/cache_ret).  tests/golden/cache_format.json records every file's dtype / rank / trailing shape
    and tests/golden/cache_ret_head the first rows of the chair files (make_cache_format_golden.py): the
    drop-in's loader must read them, and what it saves must have the same names, dtypes and shapes."""
    import json
    from corsair_b200 import evaluation_registration as er
    here = os.path.join(os.path.dirname(__file__), "golden")
    fmt = json.load(open(os.path.join(here, "cache_format.json")))
    names = {f"{k}_{cat}{sfx}" for k in er.KEYS for cat in ("chair", "table") for sfx in ("_gt.npy", "_top1.npy")}
    assert names == set(fmt["files"]), "the nine keys x category x suffix are exactly the reference's file names"
    head = er.load_results(os.path.join(here, "cache_ret_head"), "chair", False)
    assert head is not None and er.load_results(os.path.join(here, "cache_ret_head"), "chair", True) is None
    assert len(head["Ts_est_best"]) == 12 and head["Ts_est_best"][0].shape == (4, 4)
    for T in head["Ts_est_ransac"] + head["Ts_est_best"]:  # rigid poses: what eval_pose consumes
        np.testing.assert_allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-5)
        assert abs(np.linalg.det(T[:3, :3]) - 1) < 1e-5 and np.array_equal(T[3], [0, 0, 0, 1])
    s, text = er.summarize(head)
    assert 0 <= s["rre_005_sym"] <= 1 and "sym success rate" in text
    # write them back through save_results: same names, dtypes, ranks and trailing shapes as the reference's files
    er.save_results(str(tmp_path), "chair", False, head)
    for k in er.KEYS:
        name = f"{k}_chair_gt.npy"
        a = np.load(tmp_path / name)
        dtype, ndim, trailing, _ = fmt["files"][name]
        assert (str(a.dtype), a.ndim, list(a.shape[1:])) == (dtype, ndim, trailing), name
        assert np.array_equal(a, np.load(os.path.join(here, "cache_ret_head", name)))
    # the means the reference prints for its chair cache (0.0613 / 0.0559, success 100 %)
    m = fmt["chair_gt_means"]
    assert abs(m["chamfer_dist_ransac"] - 0.0613) < 5e-4 and abs(m["chamfer_dist_sym"] - 0.0559) < 5e-4
    assert m["sym_ransac_success"] == 1.0


def test_find_corr_golden():
    """utils/eval_pose.py:28-45: full clouds, and the seeded np.random.choice subsample."""
    g = load("find_corr")
    F0, F1 = torch.from_numpy(g["F0"]), torch.from_numpy(g["F1"])
    a, b = port.find_corr(g["xyz0"], g["xyz1"], F0, F1, nn_max_n=250, subsample_size=-1)
    assert np.array_equal(a, g["full_0"]) and np.array_equal(b, g["full_1"])
    np.random.seed(5)
    c, d = port.find_corr(g["xyz0"], g["xyz1"], F0, F1, nn_max_n=250, subsample_size=200)
    assert np.array_equal(c, g["sub200_0"]) and np.array_equal(d, g["sub200_1"])
    np.random.seed(6)
    e, f = port.find_corr(g["xyz0"], g["xyz1"], F0, F1, nn_max_n=-1, subsample_size=700)
    assert np.array_equal(e, g["sub700_0"]) and np.array_equal(f, g["sub700_1"])


@pytest.mark.parametrize("case", [0, 3])
def test_sym_pose_k4_golden(case):
    """utils/symmetry.py:262-358 with both sweeps (case 0: table, pos_sym 4, nine registrations, the mirrored
    order [0,3,2,1] of :326-356) and a chair (case 3): the oracle's restatement reproduces the reference's
    masks, its winner and every registration's Chamfer distance."""
    g = load("sympose_k4")
    pre = f"c{case}_"
    pos_sym = int(g[pre + "pos_sym"])
    K = 4 if pos_sym >= 2 else 2
    calls = []

    def reg(src, tgt, max_corr, seed):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        calls.append(ransac_c.ransac(src, tgt, max_corr, ms, adaptive=True)["T"])
        return calls[-1]

    np.random.seed(31)
    bm = port_symmetry.symmetric_cut4(g[pre + "baseF"], g[pre + "xyz0"], K)
    pm = port_symmetry.symmetric_cut4(g[pre + "posF"], g[pre + "xyz1"], K)
    assert np.array_equal(np.stack(bm), g[pre + "base_masks"]) and np.array_equal(np.stack(pm), g[pre + "pos_masks"])
    np.random.seed(31)
    Tb, cb, Tr, cr, ok = port_symmetry.sym_pose(g[pre + "baseF"], g[pre + "xyz0"], g[pre + "posF"], g[pre + "xyz1"], pos_sym,
                                                k_nn=5, max_corr=0.2, seed=31, registration=reg)
    assert ok and len(calls) == len(g[pre + "T_all"]) == (1 + 2 * K if pos_sym >= 2 else 1 + K)
    np.testing.assert_allclose(np.stack(calls), g[pre + "T_all"], atol=1e-12)
    assert np.allclose(Tb.numpy(), g[pre + "T_best"], atol=1e-6) and cb == pytest.approx(float(g[pre + "ch_best"]), rel=1e-9)
    assert np.allclose(Tr.numpy(), g[pre + "T_ransac"], atol=1e-6) and cr == pytest.approx(float(g[pre + "ch_ransac"]), rel=1e-9)
    assert np.allclose(Tb.numpy(), g[pre + "T_all"][int(g[pre + "best_index"])].astype(np.float32), atol=1e-6)


@pytest.mark.parametrize("name,K", [("chair0", 2), ("chair3", 2), ("table1", 4), ("table6", 4)])
def test_symmetric_cut4_seeded_clouds(name, K):
    g = load("symcut")
    np.random.seed(100 + int(name[-1]))
    try:
        masks = np.stack(port_symmetry.symmetric_cut4(g[name + "_F"], g[name + "_xyz"], K))
    except AttributeError:
        masks = np.zeros((0, len(g[name + "_xyz"])), bool)
    assert np.array_equal(masks, g[name + "_masks"])
