#!/usr/bin/env python
"""Golden vectors for the symmetry-order test of evaluation-shapenet.py:122-154, produced by the
REFERENCE'S OWN SOURCE: the three functions are cut out of /root/reference/evaluation-shapenet.py
with `ast` (the app itself cannot be imported: MinkowskiEngine, open3d, vedo ... at module scope) and
executed with scipy's KDTree and the rotation-about-y restatement of transforms3d.euler.euler2mat
from oracle/ref_import.py.  Runs in the build container only (needs /root/reference); the output
tests/golden/shapenet.npz is committed.

    python tests/golden/make_shapenet_golden.py
"""
import ast
import os
import sys
import types

import numpy as np
from scipy.spatial import KDTree

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402

WANTED = ("chamfer_max", "test_symmetry_label", "get_symmetry_label")


def reference_functions():
    path = os.path.join(ref_import.REF_ROOT, "evaluation-shapenet.py")
    tree = ast.parse(open(path).read(), path)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANTED]
    assert sorted(n.name for n in body) == sorted(WANTED)
    t3d = types.SimpleNamespace(euler=types.SimpleNamespace(euler2mat=ref_import._euler2mat))
    ns = {"np": np, "KDTree": KDTree, "t3d": t3d}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return ns


def ring(rng, order, n, jitter):
    """n points of an `order`-fold rotationally symmetric object about y (order 0: no symmetry)."""
    if order == 0:
        pc = rng.standard_normal((n, 3)) * np.array([1.0, 0.4, 0.25]) + np.array([0.3, 0.0, 0.0])
    else:
        base = max(n // order, 1)
        r = 0.4 + 0.6 * rng.random(base)
        a = (rng.random(base) - 0.5) * (1.2 * np.pi / order)
        y = rng.random(base) - 0.5
        parts = []
        for k in range(order):
            ang = a + 2 * np.pi * k / order
            parts.append(np.stack([r * np.cos(ang), y, r * np.sin(ang)], 1))
        pc = np.concatenate(parts)
    pc = pc + jitter * rng.standard_normal(pc.shape)
    pc -= pc.mean(0, keepdims=True)
    pc /= np.linalg.norm(pc, axis=1).max()
    return pc.astype(np.float32)


def main():
    ns = reference_functions()
    rng = np.random.default_rng(2026)
    out = {}
    cases = [(0, 600, 0.0), (1, 500, 0.0), (2, 600, 0.004), (3, 600, 0.004), (4, 640, 0.004), (6, 600, 0.003),
             (8, 640, 0.003), (12, 720, 0.002), (4, 640, 0.05), (12, 720, 0.03)]
    labels, errs = [], []
    for c, (order, n, jitter) in enumerate(cases):
        pc = ring(rng, order, n, jitter)
        out[f"pc{c}"] = pc
        labels.append(ns["get_symmetry_label"](pc, 0.1))
        # every rotation's chamfer_max, in the order test_symmetry_label tries them
        e = []
        for sym in (12, 8, 6, 4, 3, 2, 1):
            for i in range(1, int(sym / 2) + 1):
                R = ref_import._euler2mat(0, i * (2 * np.pi) / sym, 0)
                rot = pc @ R.T
                e.append(ns["chamfer_max"](pc, KDTree(pc), rot, KDTree(rot)))
        errs.append(e)
        print(c, order, jitter, "->", labels[-1], np.round(e[:6], 4))
    out["labels"] = np.asarray(labels, np.int32)
    out["errors"] = np.asarray(errs, np.float64)
    out["threshold"] = np.float64(0.1)
    # test_symmetry_label alone, with a threshold that splits the cases
    out["test_label4_thr005"] = np.asarray([ns["test_symmetry_label"](4, out[f"pc{c}"], 0.05) for c in range(len(cases))])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "shapenet.npz"), **out)
    print("labels", labels)


if __name__ == "__main__":
    main()
