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
    return out.to(torch.float32).cpu()  # NumPy clouds: a host value (pc_dist.compute_dist stores it in a NumPy table)


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
    the input's device (CPU tensors for NumPy input, like the reference).  Kept point = the first
    (lowest index) of its voxel, output in ascending index order.  That is MinkowskiEngine's rule:
    sparse_quantize(return_index=True, return_maps_only=True) on the CPU goes through
    CoordinateMapManagerCPU.insert_and_map (deps/MinkowskiEngine/src/coordinate_map_manager.cpp:56-69)
    into CoordinateMapCPU::insert_and_map (src/coordinate_map_cpu.hpp:355-380), a sequential loop over
    the rows that pushes a row index into the mapping only when its key is newly inserted.
    The voxel keys ``floor(pc / voxel_size)`` are computed in the input's dtype like the reference
    (float64 for the posed clouds of evaluation-shapenet.py:116-120)."""
    ctx = _capi.get_context()
    src = pc if isinstance(pc, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(pc))
    on_cuda = src.is_cuda
    if src.dtype == torch.float64:
        # keys in float64 (a float32 division could put a point on the other side of a voxel face); the
        # kernel then only has to floor exact integers
        p64 = src.to(ctx.device)
        if not on_cuda:
            ctx.h2d_bytes += src.numel() * 8
        # a DEVICE tensor as divisor: torch turns division by a host scalar into multiplication by its
        # reciprocal, which is not the correctly rounded quotient NumPy computes
        vox64 = torch.full((1,), float(voxel_size), dtype=torch.float64, device=ctx.device)
        keys = torch.floor(p64 / vox64).to(torch.float32).contiguous()
        p, vox = p64.to(torch.float32), 1.0
    else:
        p = ctx.dev(src, torch.float32)
        keys, vox = p, float(voxel_size)
    n = p.shape[0]
    idx = ctx.empty((max(n, 1),), torch.int64)
    grid = ctx.empty((max(n, 1), 3), torch.int32)
    count = ctx.empty((1,), torch.int32)
    ctx.call("cb_voxel_unique", ptr(keys), n, vox, ptr(idx), ptr(grid), ptr(count), ctx.stream)
    m = int(count.item())
    if m < 0:
        raise ValueError("quantize_pc: a coordinate is not finite or beyond 2^20 voxels")
    out_pc, out_grid = p[idx[:m]], grid[:m]
    if not on_cuda:
        ctx.d2h_bytes += out_pc.numel() * 4 + out_grid.numel() * 4
        return out_pc.cpu(), out_grid.cpu()
    return out_pc, out_grid


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
