"""GPU parity: the symmetry-order test and the registration worker of evaluation-shapenet.py
 against the reference-generated golden vectors and the oracle."""
import os

import numpy as np
import pytest
import torch

from corsair_b200 import synth
from oracle import port_shapenet as ps

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "shapenet.npz")
TOL = 1e-6  # the neighbour is chosen in fp32, its distance recomputed in float64


def test_symmetry_errors_and_labels_golden(ctx):
    from corsair_b200 import shapenet_eval as se
    g = np.load(GOLDEN)
    for c in range(len(g["labels"])):
        pc = g[f"pc{c}"]
        errs = se.symmetry_errors(pc)
        flat = [e for sym in se.SYM_ORDERS for e in errs[sym]]
        np.testing.assert_allclose(flat, g["errors"][c], rtol=0, atol=TOL)
        assert se.get_symmetry_label(pc, float(g["threshold"])) == g["labels"][c]
        assert se.test_symmetry_label(4, pc, 0.05) == bool(g["test_label4_thr005"][c])
        assert se.test_symmetry_label(1, pc, 0.0) is True  # order 1 tries no rotation (:139)


@pytest.mark.parametrize("n0,n1", [(1, 7), (257, 1000), (15000, 15000)])
def test_chamfer_max_vs_oracle(ctx, rng, n0, n1):
    from corsair_b200 import shapenet_eval as se
    a = rng.standard_normal((n0, 3)).astype(np.float32)
    b = (rng.standard_normal((n1, 3)) * 0.8 + 0.1).astype(np.float32)
    got = se.chamfer_max(a, None, b, None)
    ref = ps.chamfer_max(a.astype(np.float64), b.astype(np.float64))
    assert abs(got - ref) <= TOL


def test_symmetry_label_full_size_clouds(ctx):
    """15 000-point clouds (ShapeNetCore.v2.PC15k size): labels and all 17 errors vs the oracle."""
    from corsair_b200 import shapenet_eval as se
    rng = np.random.default_rng(5)
    for maker in (synth.make_chair, synth.make_table):
        pc = synth.normalize_cloud(maker(rng, 15000), 15000).astype(np.float32)
        errs = se.symmetry_errors(pc)
        for sym in se.SYM_ORDERS:
            np.testing.assert_allclose(errs[sym], ps.rotation_errors(sym, pc), rtol=0, atol=TOL)
        assert se.get_symmetry_label(pc, 0.1) == ps.get_symmetry_label(pc, 0.1)


def _items(n_items):
    items = []
    for i in range(n_items):
        p = synth.make_pair(300 + i, 16, "table" if i % 2 else "chair", n_raw=4000)
        # (feats_pc, feats_pc_transformed, coords_pc, pc_quant, pc_transformed_quant, pose_gt), :316-323
        items.append((p["baseF"], p["posF"], p["xyz0"], p["xyz0"], p["xyz1"], p["T_gt"] if "T_gt" in p else np.eye(4)))
    return items


def test_registration_worker_and_register_many(ctx):
    from corsair_b200 import eval_pose, shapenet_eval as se, symmetry
    items = _items(4)
    kw = dict(k_nn=5, max_corr=0.4, seed=0)
    seq = [se.registration_worker(*x, rng=0, **kw) for x in items]
    par = se.register_many(items, workers=3, **kw)
    keys = ["symmetry_label", "T_est_sym", "chamfer_dist_sym", "T_est_ransac", "chamfer_dist_ransac", "sym_success",
            "rte_sym", "rre_sym", "rte_ransac", "rre_ransac"]
    for x, a, b in zip(items, seq, par):
        assert list(a) == keys == list(b)
        assert a["symmetry_label"] == ps.get_symmetry_label(x[2], 0.1) == b["symmetry_label"]
        # the worker is get_symmetry_label -> sym_pose -> eval_pose x 2 (:244-258)
        ref = symmetry.sym_pose(x[0], x[3], x[1], x[4], a["symmetry_label"], 5, 0.4, seed=0, rng=0)
        assert torch.equal(a["T_est_sym"], ref[0]) and a["chamfer_dist_sym"] == ref[1]
        assert torch.equal(a["T_est_ransac"], ref[2]) and a["chamfer_dist_ransac"] == ref[3]
        rte, rre = eval_pose.eval_pose(ref[0], np.eye(4), x[5], axis_symmetry=a["symmetry_label"])
        assert a["rte_sym"] == rte and a["rre_sym"] == rre
        for k in keys:
            if isinstance(a[k], torch.Tensor):
                assert torch.equal(a[k], b[k])
            else:
                assert a[k] == b[k] or (np.isnan(a[k]) and np.isnan(b[k]))


@pytest.mark.parametrize("n,voxel", [(1, 0.03), (257, 0.03), (10000, 0.03), (50000, 0.01), (4000, 0.5)])
def test_quantize_pc_vs_oracle(ctx, rng, n, voxel):
    """cb_voxel_unique: first point of every occupied voxel, ascending (numpy.unique rule)."""
    from corsair_b200 import preprocess
    pc = (rng.standard_normal((n, 3)) * 0.4).astype(np.float32)
    if n > 100:
        pc[n // 2:n // 2 + 50] = pc[:50]          # exact duplicates
        pc[-10:] = -np.abs(pc[-10:]) - 0.3         # negative voxels
    ref_pc, ref_grid, _ = ps.quantize_pc(pc, np.float32(voxel))
    got_pc, got_grid = preprocess.quantize_pc(pc, voxel)
    # NumPy in -> CPU tensors out, like the reference (torch.from_numpy(...).float() / .int())
    assert got_pc.dtype == torch.float32 and got_grid.dtype == torch.int32 and not got_pc.is_cuda and not got_grid.is_cuda
    assert np.array_equal(got_pc.numpy(), ref_pc) and np.array_equal(got_grid.numpy(), ref_grid)
    # CUDA input stays where it is; the synthetic generator's dedup is the same rule
    again, g2 = preprocess.quantize_pc(torch.from_numpy(pc).cuda(), voxel)
    assert again.is_cuda and g2.is_cuda and torch.equal(again.cpu(), got_pc)
    assert np.array_equal(synth.voxel_dedup(pc, voxel), ref_pc)


@pytest.mark.parametrize("n,voxel", [(257, 0.03), (20000, 0.03), (20000, 0.01)])
def test_quantize_pc_float64_keys(ctx, rng, n, voxel):
    """The reference floors pc / voxel_size in the INPUT dtype (evaluation-shapenet.py:97-99); its posed
    clouds are float64 (:116-120).  A float32 division can put a point on the other side of a voxel face,
    so float64 input must be keyed in float64: equal to np.unique(np.floor(pc / v)) of the float64 array,
    including points planted a few float64 ulps either side of a face."""
    from corsair_b200 import preprocess
    pc = rng.standard_normal((n, 3)) * 0.4
    faces = np.round(pc[:200] / voxel) * voxel
    pc[:100] = np.nextafter(faces[:100], np.inf)       # just above a face
    pc[100:200] = np.nextafter(faces[100:], -np.inf)    # just below
    keys = np.floor(pc / voxel)
    _, first = np.unique(keys, axis=0, return_index=True)
    first = np.sort(first)
    got_pc, got_grid = preprocess.quantize_pc(pc, voxel)
    assert np.array_equal(got_grid.numpy(), keys[first].astype(np.int32))
    assert np.array_equal(got_pc.numpy(), pc[first].astype(np.float32))
    got32, _ = preprocess.quantize_pc(pc.astype(np.float32), voxel)   # float32 input: float32 keys, as NumPy would
    k32 = np.floor(pc.astype(np.float32) / np.float32(voxel))
    assert len(got32) == len(np.unique(k32, axis=0))


def test_quantize_pc_edge_cases(ctx):
    from corsair_b200 import preprocess
    e_pc, e_grid = preprocess.quantize_pc(np.zeros((0, 3), np.float32), 0.03)
    assert e_pc.shape == (0, 3) and e_grid.shape == (0, 3)
    same = np.tile(np.array([[0.011, -0.02, 0.5]], np.float32), (300, 1))
    p, g = preprocess.quantize_pc(same, 0.03)
    assert p.shape == (1, 3) and g.cpu().tolist() == [[0, -1, 16]]
    with pytest.raises(ValueError):
        preprocess.quantize_pc(np.array([[0.0, np.inf, 0.0]], np.float32), 0.03)
    with pytest.raises(ValueError):
        preprocess.quantize_pc(np.array([[1e9, 0.0, 0.0]], np.float32), 0.03)


def test_split_batch_matches_mask_loop(ctx, rng):
    from corsair_b200 import preprocess
    b = torch.from_numpy(rng.integers(0, 5, 3000)).int()
    b[b == 3] = 4                                  # an empty batch index in the middle
    C = torch.cat([b[:, None], torch.zeros((3000, 3), dtype=torch.int32)], 1).cuda()
    F = torch.from_numpy(rng.standard_normal((3000, 16)).astype(np.float32)).cuda()
    got = preprocess.split_batch(C, F)
    ref = [F[C[:, 0] == i, :] for i in range(1 + int(C[:, 0].max()))]   # model/fc.py:23-29
    assert len(got) == len(ref) and all(torch.equal(a, r) for a, r in zip(got, ref))
