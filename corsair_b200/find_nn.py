"""Drop-in for the reference's utils/find_nn.py (same names, arguments and return types).

All five functions run the fused kNN engine of the C-ABI library (cb_knn): a candidate pass
(tensor-core or SIMT) followed by an exact float64 difference-form re-rank, so indices equal the
float64 KD-tree answer of the reference's CPU functions with ties broken by lower index.
"""
import numpy as np
import torch

from . import _capi
from ._capi import ptr

_METRIC = {"SquareL2": 0, "L2": 1}


def _knn(F0, F1, k, metric, want_dist, mutual=False):
    ctx = _capi.get_context()
    q = ctx.dev(F0, torch.float32)
    r = ctx.dev(F1, torch.float32)
    if q.dim() != 2 or r.dim() != 2 or q.shape[1] != r.shape[1]:
        raise ValueError(f"feature arrays must be [N,D] with equal D, got {tuple(q.shape)} {tuple(r.shape)}")
    nq, d = q.shape
    nr = r.shape[0]
    if nr < k:
        raise ValueError(f"k={k} neighbours requested from {nr} reference points")
    idx = ctx.empty((nq, k), torch.int64)
    dist = ctx.empty((nq, k), torch.float32) if want_dist else None
    mut = ctx.empty((nq,), torch.uint8) if mutual else None
    ctx.call("cb_knn", ptr(q), nq, ptr(r), nr, d, k, metric, ptr(idx), ptr(dist), ptr(mut), ctx.stream)
    return idx, dist, mut


def knn_device(F0, F1, k, metric=0, return_distance=False, mutual=False):
    """Device-resident variant: returns CUDA tensors (idx [N,k] int64[, dist [N,k] f32][, mutual])."""
    idx, dist, mut = _knn(F0, F1, k, metric, return_distance, mutual)
    out = (idx,)
    if return_distance:
        out += (dist,)
    if mutual:
        out += (mut,)
    return out if len(out) > 1 else idx


def pdist(A, B, dist_type="L2"):
    """utils/find_nn.py:24-31: dense [n,m] distance matrix on A's device."""
    if dist_type not in _METRIC:
        raise NotImplementedError("Not implemented")
    ctx = _capi.get_context()
    a = ctx.dev(A, torch.float32)
    b = ctx.dev(B, torch.float32)
    out = ctx.empty((a.shape[0], b.shape[0]), torch.float32)
    if out.numel():
        ctx.call("cb_pdist", ptr(a), a.shape[0], ptr(b), b.shape[0], a.shape[1], _METRIC[dist_type],
                 ptr(out), ctx.stream)
    if isinstance(A, torch.Tensor):
        return out.to(device=A.device, dtype=A.dtype)
    return out


def find_nn_cpu(feat0, feat1, return_distance=False):
    """utils/find_nn.py:34-40: nn_inds [N0] int64 (and Euclidean dists [N0] float64), NumPy."""
    idx, dist, _ = _knn(feat0, feat1, 1, 0, return_distance)
    inds = idx[:, 0].cpu().numpy()
    if return_distance:
        return inds, np.sqrt(dist[:, 0].cpu().numpy().astype(np.float64))
    return inds


def find_knn_cpu(feat0, feat1, k, return_distance=False):
    """utils/find_nn.py:43-49: nn_inds [N0,k] int64 (and Euclidean dists [N0,k] float64), NumPy."""
    idx, dist, _ = _knn(feat0, feat1, k, 0, return_distance)
    inds = idx.cpu().numpy()
    if k == 1:
        inds = inds[:, 0]  # scipy squeezes k=1
    if return_distance:
        d = np.sqrt(dist.cpu().numpy().astype(np.float64))
        return inds, (d[:, 0] if k == 1 else d)
    return inds


def find_nn_gpu(F0, F1, nn_max_n=-1, return_distance=False, dist_type="SquareL2"):
    """utils/find_nn.py:52-82: inds [N0] int64 CPU tensor, dists [N0,1] CPU tensor.

    ``nn_max_n`` only bounded the reference's [n,m,D] broadcast; the fused kernel never
    materialises the distance matrix, so it is accepted and ignored."""
    if dist_type not in _METRIC:
        raise NotImplementedError("Not implemented")
    idx, dist, _ = _knn(F0, F1, 1, _METRIC[dist_type], return_distance)
    inds = idx[:, 0].cpu()
    if return_distance:
        return inds, dist.cpu()
    return inds


def find_knn_gpu(F0, F1, k, nn_max_n=-1, return_distance=False, dist_type="SquareL2"):
    """utils/find_nn.py:86-119: inds [N0*k] int64 CPU tensor, dists [N0*k,1] CPU tensor."""
    if dist_type not in _METRIC:
        raise NotImplementedError("Not implemented")
    idx, dist, _ = _knn(F0, F1, k, _METRIC[dist_type], return_distance)
    inds = idx.reshape(-1).cpu()
    assert len(inds) == len(F0) * k
    if return_distance:
        return inds, dist.reshape(-1, 1).cpu()
    return inds
