"""Drop-in for the hot-path functions of the reference's utils/preprocess.py
(apply_transform :39-48, chamfer_gpu :51-56, chamfer_gpu_1direction :59-64,
chamfer_kdtree_1direction :67-70, load_norm_pc :32-36).  The label/split readers of that
file are dataset IO and out of scope."""
import numpy as np
import torch

from . import _capi
from ._capi import ptr


def load_norm_pc(path, samples):
    """utils/preprocess.py:32-36 (host IO; unchanged semantics)."""
    pc = np.load(path)[:samples, :]
    pc -= np.mean(pc, 0)
    pc /= np.max(np.linalg.norm(pc, 2, 1))
    return pc


def apply_transform(pointcloud, T):
    """utils/preprocess.py:39-48: homogeneous T.[p;1] -> [N,3] float64 NumPy (host; tiny)."""
    if isinstance(T, torch.Tensor):
        T = T.detach().cpu().numpy()
    pc = np.asarray(pointcloud)
    homo = np.concatenate([pc, np.ones([pc.shape[0], 1])], axis=1)
    return np.matmul(T, homo.T).T[:, :3]


def _chamfer(pc0, pc1, T, two_dir, ctx=None):
    ctx = ctx or _capi.get_context()
    a = ctx.dev(pc0, torch.float32)
    b = ctx.dev(pc1, torch.float32)
    if a.dim() != 2 or a.shape[1] != 3 or b.dim() != 2 or b.shape[1] != 3:
        raise ValueError("point clouds must be [N,3]")
    if a.shape[0] == 0 or b.shape[0] == 0:
        raise ValueError("empty point cloud")
    Td = ctx.dev(np.asarray(T, dtype=np.float64).reshape(16), torch.float64) if T is not None else None
    out = ctx.empty((1,), torch.float64)
    ctx.call("cb_chamfer", ptr(a), a.shape[0], ptr(b), b.shape[0], ptr(Td), int(two_dir), ptr(out), None,
             ctx.stream)
    return out


def chamfer_device(pc0, pc1, T=None, two_dir=False):
    """Device-resident variant: 1-element float64 CUDA tensor; ``T`` (4x4) is applied to pc0."""
    if isinstance(T, torch.Tensor):
        T = T.detach().cpu().numpy()
    return _chamfer(pc0, pc1, T, two_dir)


def chamfer_gpu(pc0, pc1):
    """utils/preprocess.py:51-56: two-direction Chamfer, 0-d tensor on pc0's device."""
    out = _chamfer(pc0, pc1, None, True)[0]
    if isinstance(pc0, torch.Tensor):
        return out.to(device=pc0.device, dtype=pc0.dtype)
    return out.to(torch.float32)


def chamfer_gpu_1direction(pc0, pc1):
    """utils/preprocess.py:59-64: mean over pc0 of the NN distance into pc1 -> python float."""
    return float(_chamfer(pc0, pc1, None, False).item())


def chamfer_kdtree_1direction(pc0, pc1):
    """utils/preprocess.py:67-70 -> float (float64 mean of nearest-neighbour distances).

    float64 inputs (e.g. the output of apply_transform) are rounded to float32 for the search;
    use ``chamfer_device(xyz, pc1, T)`` to apply a pose in float64 on the device instead."""
    return float(_chamfer(pc0, pc1, None, False).item())


def quantize_pc(pc, voxel_size):
    """evaluation-shapenet.py:97-107 (and the library-side dedup of utils/Info/CADLib.py:106-121) on
    the device: one point per occupied voxel -> (pc [M,3] float32, pc_grid [M,3] int32) as tensors on
    the input's device (CUDA for NumPy input).  Kept point = the first (lowest index) of its voxel,
    output in ascending index order: ``np.unique(np.floor(pc / v), axis=0, return_index=True)`` with
    the indices sorted.  MinkowskiEngine's own choice among duplicates cannot be checked here (not
    installed): DESIGN.md lists this call as parity-unpinned."""
    ctx = _capi.get_context()
    p = ctx.dev(pc, torch.float32)
    n = p.shape[0]
    idx = ctx.empty((max(n, 1),), torch.int64)
    grid = ctx.empty((max(n, 1), 3), torch.int32)
    count = ctx.empty((1,), torch.int32)
    ctx.call("cb_voxel_unique", ptr(p), n, float(voxel_size), ptr(idx), ptr(grid), ptr(count), ctx.stream)
    m = int(count.item())
    if m < 0:
        raise ValueError("quantize_pc: a coordinate is not finite or beyond 2^20 voxels")
    return p[idx[:m]], grid[:m]


def split_batch(C, F):
    """model/fc.py:23-29 for the tensors of a sparse output (``sp.C`` [N,1+3] int, ``sp.F`` [N,D]):
    the feature rows of every batch index, in their original order, as a list of device tensors —
    the hand-off from the feature producer to sym_pose without a ``.cpu().numpy()`` round trip.
    One stable sort + one bincount (a single host read of the counts) instead of one boolean mask,
    and one synchronisation, per batch item."""
    b = C[:, 0].long()
    counts = torch.bincount(b)
    order = torch.argsort(b, stable=True)
    return list(F[order].split(counts.tolist()))
