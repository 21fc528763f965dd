"""Drop-in for the reference's utils/pc_dist.py (chamfer :45-52, compute_chamfer :55-58,
compute_dist :61-99, shapenet_cat :102-125)."""
import os

import numpy as np
import torch

from . import _capi
from ._capi import ptr
from .preprocess import chamfer_gpu, load_norm_pc


def chamfer(pc0, pc1):
    """utils/pc_dist.py:45-52: two-direction Chamfer distance (0-d tensor)."""
    return chamfer_gpu(pc0, pc1)


def compute_chamfer(base_idx, pc0, pcs, domain, table):
    """utils/pc_dist.py:55-58: fill table[base_idx, i] for i in [domain[0], domain[1])."""
    for i in range(domain[0], domain[1]):
        table[base_idx, i] = chamfer(pc0, pcs[i])


def chamfer_rows_device(pts, row0=0, row1=None, ctx=None, half=False):
    """Rows [row0,row1) of the all-pairs table for a [L,P,3] stack -> [rows, L] float64 CUDA tensor.
    ``half``: only the circulant half of those rows (cb_chamfer_matrix_half: every unordered pair once
    over all rows, zeros elsewhere) — what a row-sharded run computes before symmetrising."""
    ctx = ctx or _capi.get_context()
    p = ctx.dev(pts, torch.float32)
    L, P = p.shape[0], p.shape[1]
    row1 = L if row1 is None else row1
    out = ctx.empty((row1 - row0, L), torch.float64)
    ctx.call("cb_chamfer_matrix_half" if half else "cb_chamfer_matrix", ptr(p), L, P, row0, row1, ptr(out), ctx.stream)
    return out


def compute_dist(pcs):
    """utils/pc_dist.py:61-99 -> [L,L] float64 NumPy, symmetric, diagonal 200 (eye*100 + its transpose).

    Clouds of equal size go through the all-pairs kernel; ragged lists fall back to one
    cb_chamfer call per pair (still on the device)."""
    L = len(pcs)
    sizes = {int(len(p)) for p in pcs}
    if L and len(sizes) == 1:
        stack = torch.stack([p if isinstance(p, torch.Tensor) else torch.from_numpy(np.asarray(p)) for p in pcs])
        return chamfer_rows_device(stack.float()).cpu().numpy()
    table = np.eye(L) * 100
    for i in range(L):
        compute_chamfer(i, pcs[i], pcs, [i + 1, L], table)
    table += table.T
    return table


def shapenet_cat(split, catid):
    """utils/pc_dist.py:102-121 (host IO around compute_dist; same hard-coded paths)."""
    root = "/scannet/ShapeNetCore.v2.PC15k"
    target = "/scannet/tables"
    files = os.listdir(os.path.join(root, catid, split))
    print(catid, len(files))
    files.sort()
    pcs = [torch.Tensor(load_norm_pc(os.path.join(root, catid, split, objid), 2000)) for objid in files]
    table = compute_dist(pcs)
    np.save(os.path.join(target, "{}_{}.npy".format(catid, split)), table)


if __name__ == "__main__":
    shapenet_cat("train", "03001627")
    shapenet_cat("val", "03001627")
    shapenet_cat("test", "03001627")
