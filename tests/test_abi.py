"""CPU: the C-ABI library loads, exports every symbol include/corsair_b200.h declares, the ctypes
table covers the header, and the product path fails loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "corsair_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from corsair_b200 import _capi
    lib = _capi.load_library()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_capi.SIGNATURES) == names, "ctypes signature table and header disagree"
    assert lib.cb_version() == 100


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np
    from corsair_b200 import _capi, find_nn, preprocess, symmetry
    h = ctypes.c_void_p()
    assert _capi.load_library().cb_create(ctypes.byref(h), 0, 0) != 0 and not h
    f = np.zeros((4, 16), np.float32)
    with pytest.raises(RuntimeError):
        find_nn.find_knn_cpu(f, f, 1)
    with pytest.raises(RuntimeError):
        preprocess.chamfer_kdtree_1direction(f[:, :3], f[:, :3])
    with pytest.raises(RuntimeError):
        symmetry.sym_pose(f, f[:, :3], f, f[:, :3], 1)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "corsair_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src and "sklearn.cluster" not in src and "KDTree" not in src, fn


def test_dropin_install_aliases_reference_module_names():
    import sys
    import corsair_b200
    corsair_b200.install("utils_dropin_test")
    from utils_dropin_test import find_nn, symmetry  # noqa: F401
    for mod, names in {
        "find_nn": ["pdist", "find_nn_cpu", "find_knn_cpu", "find_nn_gpu", "find_knn_gpu"],
        "eval_pose": ["find_corr", "find_kcorr", "registration_based_on_corr", "eval_pose"],
        "symmetry": ["split_corr", "symmetric_cut4", "sym_pose", "find_kcorr", "chamfer_kdtree_1direction"],
        "preprocess": ["apply_transform", "chamfer_gpu", "chamfer_gpu_1direction", "chamfer_kdtree_1direction",
                       "load_norm_pc"],
        "retrieval": ["compute_mAP", "retrieval_dist", "retrieval_eval", "get_rank", "scan2cad_retrieval_eval_dist",
                      "scan2cad_retrieval_eval"],
        "pc_dist": ["chamfer", "compute_chamfer", "compute_dist", "shapenet_cat"],
        "_transform_estimation": ["rot_x", "rot_y", "rot_z", "get_trans", "update_pcd", "build_linear_system",
                                  "solve_linear_system", "compute_weights", "est_quad_linear_robust",
                                  "pose_estimation"],
    }.items():
        m = sys.modules["utils_dropin_test." + mod]
        for n in names:
            assert callable(getattr(m, n)), f"{mod}.{n}"


def test_np_seeded_choices_replays_numpy():
    """cb_np_seeded_choices (host only) against NumPy's legacy RandomState: the seed rows symmetric_cut4
    draws (utils/symmetry.py:193) must be the same whether Python or the C call draws them."""
    import numpy as np
    from corsair_b200 import _capi
    f = _capi.load_library().cb_np_seeded_choices
    for seed in (0, 1, 31, 20260101, 2 ** 32 - 1):
        for ns in ([4096, 3000], [100, 100, 101], [2048], [70000, 257], [1]):
            size = min(100, min(ns))
            n = np.asarray(ns, np.int32)
            out = np.full((len(ns), size), -1, np.int32)
            assert f(seed, n.ctypes.data, len(ns), size, out.ctypes.data) == 0
            rs = np.random.RandomState(seed)
            for i, k in enumerate(ns):
                assert np.array_equal(rs.choice(k, size, replace=False), out[i]), (seed, ns, i)
    n = np.asarray([4096, 99], np.int32)
    out = np.zeros((2, 100), np.int32)
    assert f(3, n.ctypes.data, 2, 100, out.ctypes.data) != 0  # NumPy raises: sample larger than population
