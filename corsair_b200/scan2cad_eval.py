"""Drop-in for the shape-estimation branch of the reference's Scan2CAD app and its precompute step
(SURVEY.md §8(f) rank 3): the Chamfer-table re-rank of the top ``use_best`` retrieved CADs
(evaluation-scan2cad.py:322-330), ``evaluate_retrieval`` (:343-362, two-direction Chamfer of two
15 000-point clouds, run under ``thread_map``) and the GT x reconstruction Chamfer table of
utils/precompute_chamfer_splats.py:113-149.  Mesh reading / sampling and the dataset objects stay
on the caller's side: these functions take the arrays the app already holds.
"""
import numpy as np
import torch

from . import _capi, retrieval
from ._capi import ptr
from .preprocess import chamfer_device


def rerank_by_chamfer(descriptors, lib_descriptors, chamfer_table, best_matches_idx, use_best=30):
    """evaluation-scan2cad.py:322-330.  For every scene descriptor: the ``use_best`` library rows
    nearest in Euclidean distance (``cdist`` + ``argsort`` there, cb_retrieval_topk here: ascending
    distance, ties by lower index), then the one among them whose precomputed Chamfer distance to the
    scene's ground-truth CAD (row ``best_matches_idx[scene]`` of the table) is smallest.

    Returns (topN_idx [Q, use_best] int64, retrieved_object_idx [Q] int64) as NumPy arrays."""
    best_matches_idx = np.asarray(best_matches_idx, dtype=np.int64)
    table = np.asarray(chamfer_table)
    idx, _ = retrieval.topk(descriptors, lib_descriptors, int(use_best))
    idx = np.asarray(idx, dtype=np.int64)
    cand = table[best_matches_idx[:, None], idx]            # [Q, use_best] Chamfer to the GT CAD
    pick = np.argmin(cand, axis=1)                           # first minimum, like np.argmin there
    return idx, idx[np.arange(len(idx)), pick]


def evaluate_retrieval(align_cad_xyz, retrieved_xyz):
    """evaluation-scan2cad.py:343-362 for the two clouds it loads: chamfer_kdtree_1direction(a, b) +
    chamfer_kdtree_1direction(b, a) -> float."""
    return float(chamfer_device(align_cad_xyz, retrieved_xyz, None, True).item())


def evaluate_retrieval_many(pairs):
    """The ``thread_map(self.evaluate_retrieval, ...)`` loop (:332-336): all pairs are enqueued on the
    stream, one read at the end.  ``pairs``: iterable of (align_cad_xyz, retrieved_xyz)."""
    outs = [chamfer_device(a, b, None, True) for a, b in pairs]
    if not outs:
        return []
    return torch.cat(outs).cpu().numpy().tolist()


def chamfer_cross_table(gt_clouds, recon_clouds):
    """utils/precompute_chamfer_splats.py:113-149 for stacks of already loaded clouds: [LA, LB]
    float64 NumPy table of two-direction Chamfer distances (entry (i, j): GT cloud i vs reconstruction
    j), the array evaluation-scan2cad.py:311-317 reshapes out of ``chamfer_dist_list.csv``."""
    ctx = _capi.get_context()
    a = ctx.dev(np.asarray(gt_clouds) if not isinstance(gt_clouds, torch.Tensor) else gt_clouds, torch.float32)
    b = ctx.dev(np.asarray(recon_clouds) if not isinstance(recon_clouds, torch.Tensor) else recon_clouds, torch.float32)
    if a.dim() != 3 or b.dim() != 3 or a.shape[2] != 3 or b.shape[2] != 3:
        raise ValueError("chamfer_cross_table expects [L, P, 3] stacks")
    out = ctx.empty((a.shape[0], b.shape[0]), torch.float64)
    ctx.call("cb_chamfer_cross", ptr(a), a.shape[0], a.shape[1], ptr(b), b.shape[0], b.shape[1], ptr(out), ctx.stream)
    return out.cpu().numpy()
