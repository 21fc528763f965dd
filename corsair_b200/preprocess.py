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
