#!/usr/bin/env python
"""Golden vectors for the shape-estimation branch of evaluation-scan2cad.py (:322-330, :343-362) and
the cross table of utils/precompute_chamfer_splats.py:113-149.  The branch is inline code of the
app's constructor, so it is replayed here statement for statement with scipy's cdist / NumPy's
argsort / argmin and THE REFERENCE'S OWN chamfer_kdtree_1direction (utils/preprocess.py, imported
from /root/reference through oracle/ref_import.py).  Build container only; the output
tests/golden/scan2cad.npz is committed.

    python tests/golden/make_scan2cad_golden.py
"""
import os
import sys

import numpy as np
from scipy.spatial.distance import cdist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402


def cloud(rng, n, scale):
    p = rng.standard_normal((n, 3)) * scale
    p -= p.mean(0)
    p /= np.linalg.norm(p, axis=1).max()
    return p.astype(np.float32)


def main():
    pp = ref_import.load("preprocess")
    rng = np.random.default_rng(77)
    # cross table: 5 "GT" clouds x 4 "reconstructions" (noisy copies of some of them), 1500 points
    gt = np.stack([cloud(rng, 1500, np.array([1.0, 0.3 + 0.15 * i, 0.5])) for i in range(5)])
    recon = np.stack([(gt[i % 5] + 0.01 * rng.standard_normal(gt[0].shape)).astype(np.float32)[rng.permutation(1500)]
                      for i in (0, 2, 3, 1)])
    table = np.array([[pp.chamfer_kdtree_1direction(a, b) + pp.chamfer_kdtree_1direction(b, a) for b in recon] for a in gt])
    # re-rank: 12 scene descriptors vs 40 library descriptors, a 40 x 40 Chamfer table, use_best = 6
    lib = rng.standard_normal((40, 64)).astype(np.float32)
    lib /= np.linalg.norm(lib, axis=1, keepdims=True)
    best = rng.integers(0, 40, 12)
    desc = (lib[best] + 0.35 * rng.standard_normal((12, 64))).astype(np.float32)
    desc /= np.linalg.norm(desc, axis=1, keepdims=True)
    ctab = rng.random((40, 40))
    ctab = ctab + ctab.T
    use_best = 6
    feature_dist = cdist(desc, lib)                                   # evaluation-scan2cad.py:322
    topN_idx = np.argsort(feature_dist, axis=-1)                      # :323
    retrieved = np.array([topN_idx[s, np.argmin(ctab[best[s], topN_idx[s, :use_best].flatten()])]
                          for s in range(len(best))])                 # :324-329
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "scan2cad.npz"), gt=gt, recon=recon, table=table,
                        lib=lib, desc=desc, best=best, ctab=ctab, use_best=use_best, topN=topN_idx[:, :use_best],
                        retrieved=retrieved)
    print(table.round(4))
    print(retrieved)


if __name__ == "__main__":
    main()
