"""CPU restatement of the shape-estimation branch of evaluation-scan2cad.py (:322-330, :343-362) and
of the table utils/precompute_chamfer_splats.py:113-149 builds.  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.

The branch is inline code of App.__init__ (not callable on its own); its two numeric ingredients
are scipy's ``cdist`` + ``np.argsort`` / ``np.argmin`` and the reference's
``chamfer_kdtree_1direction`` (utils/preprocess.py:67-70).  tests/golden/scan2cad.npz was produced
with the reference's own chamfer_kdtree_1direction (imported from /root/reference) and these very
NumPy / scipy calls: tests/golden/make_scan2cad_golden.py.
"""
import numpy as np
from scipy.spatial import cKDTree
from scipy.spatial.distance import cdist


def chamfer_1direction(pc0, pc1):
    """utils/preprocess.py:67-70: mean over pc0 of the distance to its nearest neighbour in pc1."""
    d, _ = cKDTree(pc1).query(pc0, k=1)
    return float(np.mean(d))


def chamfer_2direction(pc0, pc1):
    """evaluation-scan2cad.py:358-360 / precompute_chamfer_splats.py:146-147."""
    return chamfer_1direction(pc0, pc1) + chamfer_1direction(pc1, pc0)


def rerank_by_chamfer(descriptors, lib_descriptors, chamfer_table, best_matches_idx, use_best=30):
    """evaluation-scan2cad.py:322-330."""
    feature_dist = cdist(descriptors, lib_descriptors)
    topN_idx = np.argsort(feature_dist, axis=-1)
    retrieved = np.array([
        topN_idx[s, np.argmin(chamfer_table[best_matches_idx[s], topN_idx[s, :use_best].flatten()])]
        for s in range(len(best_matches_idx))])
    return topN_idx[:, :use_best], retrieved


def chamfer_cross_table(gt_clouds, recon_clouds):
    """precompute_chamfer_splats.py:113-149 over loaded clouds."""
    return np.array([[chamfer_2direction(a, b) for b in recon_clouds] for a in gt_clouds])
