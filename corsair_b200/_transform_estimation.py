"""Drop-in for the reference's utils/_transform_estimation.py (rot_x/y/z :21-54, get_trans :57-61,
update_pcd :64-71, build_linear_system :74-95, solve_linear_system :98-103, compute_weights
:106-107, est_quad_linear_robust :110-137, pose_estimation :140-159).

The small helpers keep their torch semantics (they are API surface, a few flops each); the two
hot pieces run on the device: the F0.F1^T row-max correspondence search (cb_knn with the
inner-product metric) and the 20-round IRLS solve (cb_irls_pose, one fused kernel)."""
import numpy as np
import torch

from . import _capi
from ._capi import ptr


def rot_x(x):
    out = torch.zeros((3, 3))
    c, s = torch.cos(x), torch.sin(x)
    out[0, 0] = 1
    out[1, 1], out[1, 2], out[2, 1], out[2, 2] = c, -s, s, c
    return out


def rot_y(x):
    out = torch.zeros((3, 3))
    c, s = torch.cos(x), torch.sin(x)
    out[0, 0], out[0, 2], out[1, 1], out[2, 0], out[2, 2] = c, s, 1, -s, c
    return out


def rot_z(x):
    out = torch.zeros((3, 3))
    c, s = torch.cos(x), torch.sin(x)
    out[0, 0], out[0, 1], out[1, 0], out[1, 1], out[2, 2] = c, -s, s, c, 1
    return out


def get_trans(x):
    trans = torch.eye(4)
    trans[:3, :3] = rot_z(x[2]).mm(rot_y(x[1])).mm(rot_x(x[0]))
    trans[:3, 3] = x[3:, 0]
    return trans


def update_pcd(pts, trans):
    trans = trans.float()
    R = trans[:3, :3]
    T = trans[:3, 3]
    return torch.t(R @ torch.t(pts.float())) + T


def build_linear_system(pts0, pts1, weight):
    npts0 = pts0.shape[0]
    A0 = torch.zeros((npts0, 6))
    A1 = torch.zeros((npts0, 6))
    A2 = torch.zeros((npts0, 6))
    A0[:, 1], A0[:, 2], A0[:, 3] = pts0[:, 2], -pts0[:, 1], 1
    A1[:, 0], A1[:, 2], A1[:, 4] = -pts0[:, 2], pts0[:, 0], 1
    A2[:, 0], A2[:, 1], A2[:, 5] = pts0[:, 1], -pts0[:, 0], 1
    ww1 = weight.repeat(3, 6)
    ww2 = weight.repeat(3, 1)
    A = ww1 * torch.cat((A0, A1, A2), 0)
    b = ww2 * torch.cat((pts1[:, 0] - pts0[:, 0], pts1[:, 1] - pts0[:, 1], pts1[:, 2] - pts0[:, 2]), 0).unsqueeze(1)
    return A, b


def solve_linear_system(A, b):
    A, b = A.float(), b.float()
    return torch.inverse(A.t().mm(A)).mm(A.t()).mm(b)


def compute_weights(pts0, pts1, par):
    return par / (torch.norm(pts0 - pts1, dim=1).unsqueeze(1) + par)


def irls_pose_device(pts0, pts1, weight=None, iters=20, ctx=None):
    """cb_irls_pose -> [4,4] float32 CUDA tensor."""
    ctx = ctx or _capi.get_context()
    p0 = ctx.dev(pts0, torch.float32)
    p1 = ctx.dev(pts1, torch.float32)
    if p0.shape != p1.shape or p0.dim() != 2 or p0.shape[1] != 3:
        raise ValueError("pts0 and pts1 must both be [N,3]")
    w = ctx.dev(weight, torch.float32).reshape(-1) if weight is not None else None
    if w is not None and w.numel() != p0.shape[0]:
        raise ValueError("weight must be [N,1]")
    T = ctx.empty((4, 4), torch.float32)
    ctx.call("cb_irls_pose", ptr(p0), ptr(p1), ptr(w), p0.shape[0], int(iters), ptr(T), ctx.stream)
    return T


def est_quad_linear_robust(pts0, pts1, weight=None):
    """utils/_transform_estimation.py:110-137 -> [4,4] float32 CPU tensor."""
    return irls_pose_device(pts0, pts1, weight, 20).cpu()


def pose_from_features(xyz0, xyz1, F0, F1, return_corr=False):
    """The part of pose_estimation after the two model forwards (:149-159)."""
    ctx = _capi.get_context()
    f0 = ctx.dev(F0, torch.float32)
    f1 = ctx.dev(F1, torch.float32)
    idx = ctx.empty((f0.shape[0], 1), torch.int64)
    wgt = ctx.empty((f0.shape[0], 1), torch.float32)
    ctx.call("cb_knn", ptr(f0), f0.shape[0], ptr(f1), f1.shape[0], f0.shape[1], 1, 2, ptr(idx), ptr(wgt), None,
             ctx.stream)
    x1 = ctx.dev(xyz1, torch.float32)[idx[:, 0]]
    trans = irls_pose_device(xyz0, x1, wgt, 20).cpu()
    weight = wgt.cpu()
    if return_corr:
        corr = ctx.empty((f0.shape[0], f1.shape[0]), torch.float32)
        ctx.call("cb_dot_matrix", ptr(f0), f0.shape[0], ptr(f1), f1.shape[0], f0.shape[1], ptr(corr), ctx.stream)
        return trans, weight, corr
    return trans, weight


def pose_estimation(model, device, xyz0, xyz1, coord0, coord1, feats0, feats1, return_corr=False):
    """utils/_transform_estimation.py:140-159.  The MinkowskiEngine ResUNet stays the reference
    feature producer (BASELINE.json); everything after its two forwards runs in this library."""
    import MinkowskiEngine as ME  # the reference's own dependency, only needed for this call
    sinput0 = ME.SparseTensor(feats0, coords=coord0).to(device)
    F0 = model(sinput0).F
    sinput1 = ME.SparseTensor(feats1, coords=coord1).to(device)
    F1 = model(sinput1).F
    return pose_from_features(xyz0, xyz1, F0, F1, return_corr)
