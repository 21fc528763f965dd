"""Seeded synthetic inputs of the shapes BASELINE.json names: procedural
chairs / tables normalised and voxel-deduplicated like the reference's loaders
(utils/preprocess.py:32-36, utils/Info/CADLib.py:106-121), rigidly moved query copies
(evaluation-shapenet.py:79-94), symmetric per-point features standing in for the ResUNet output,
and unit-norm 256-d global descriptors.  Host NumPy only; used by tests, smoke() and bench.py.
"""
import numpy as np


def _box(rng, n, lo, hi):
    return rng.uniform(lo, hi, (n, 3))


def make_chair(rng, n_pts=15000, bulk=1.0):
    """Mirror-symmetric (about x = 0) chair: seat, back rest, four legs.  ``bulk`` thickens the seat,
    the back and the legs (and adds arm rests above 3): the same 10 000 raw points then fall into more
    0.03-voxels, so clouds after de-duplication span the 2.6k-8.0k points SURVEY.md F5 measured on
    the real ShapeNet chairs instead of sitting at the light end (bulk = 1: 3.6k-4.7k)."""
    w, dpt = rng.uniform(0.35, 0.5), rng.uniform(0.35, 0.5)
    seat_h, back_h, leg = rng.uniform(0.35, 0.5), rng.uniform(0.4, 0.7), rng.uniform(0.02, 0.04)
    st, bt = min(0.04 * bulk, 0.6 * seat_h), min(0.05 * bulk, 1.5 * dpt)
    leg = leg * (1.0 + 0.25 * (bulk - 1.0)) if bulk > 1.0 else leg
    parts = [
        (0.35, [-w, seat_h - st, -dpt], [w, seat_h, dpt]),
        (0.35, [-w, seat_h, -dpt], [w, seat_h + back_h, -dpt + bt]),
    ]
    for sx in (-1, 1):
        for sz in (-1, 1):
            cx, cz = sx * (w - leg), sz * (dpt - leg)
            parts.append((0.075, [cx - leg, 0, cz - leg], [cx + leg, seat_h - st, cz + leg]))
    if bulk > 3.0:
        for sx in (-1, 1):
            parts.append((0.1, [sx * w - 0.06, seat_h, -dpt], [sx * w + 0.06, seat_h + 0.25, dpt]))
    tot = sum(f for f, _, _ in parts)
    pts = [_box(rng, max(1, int(f / tot * n_pts)), lo, hi) for f, lo, hi in parts]
    pc = np.concatenate(pts)[:n_pts]
    return pc[rng.permutation(len(pc))].astype(np.float32)


def make_table(rng, n_pts=15000):
    """4-fold rotationally symmetric (about y) table: square top, four legs."""
    w, h, leg = rng.uniform(0.45, 0.6), rng.uniform(0.4, 0.6), rng.uniform(0.03, 0.05)
    parts = [(0.6, [-w, h - 0.05, -w], [w, h, w])]
    for sx in (-1, 1):
        for sz in (-1, 1):
            cx, cz = sx * (w - 2 * leg), sz * (w - 2 * leg)
            parts.append((0.1, [cx - leg, 0, cz - leg], [cx + leg, h - 0.05, cz + leg]))
    pts = [_box(rng, max(1, int(f * n_pts)), lo, hi) for f, lo, hi in parts]
    pc = np.concatenate(pts)[:n_pts]
    return pc[rng.permutation(len(pc))].astype(np.float32)


def normalize_cloud(pc, samples=10000):
    """load_norm_pc (utils/preprocess.py:32-36) on an in-memory array."""
    pc = np.array(pc[:samples], dtype=np.float32, copy=True)
    pc -= pc.mean(0)
    pc /= np.max(np.linalg.norm(pc, 2, 1))
    return pc


def voxel_dedup(pc, voxel=0.03):
    """One point per voxel, first occurrence kept (quantize_pc, evaluation-shapenet.py:97-107)."""
    keys = np.floor(pc / voxel).astype(np.int64)
    _, first = np.unique(keys, axis=0, return_index=True)
    return pc[np.sort(first)]


def random_pose(rng, max_angle=360.0, max_trans=1.0):
    """generate_random_pose (evaluation-shapenet.py:79-94): Euler angles and translation, uniform."""
    a = np.deg2rad(rng.uniform(-max_angle, max_angle, 3))
    cx, sx, cy, sy, cz, sz = np.cos(a[0]), np.sin(a[0]), np.cos(a[1]), np.sin(a[1]), np.cos(a[2]), np.sin(a[2])
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    T = np.eye(4)
    T[:3, :3] = Rz @ Ry @ Rx
    T[:3, 3] = rng.uniform(-max_trans, max_trans, 3)
    return T


def symmetric_features(pc, W, kind, noise, rng):
    """Unit-norm features shared by symmetric points: kind 'mirror' uses (|x|, y, z), kind 'rot4'
    uses (r, y, cos 4t / 2, sin 4t / 2) so the four rotated copies of a point coincide."""
    if kind == "mirror":
        phi = np.stack([np.abs(pc[:, 0]), pc[:, 1], pc[:, 2]], 1)
    else:
        r = np.hypot(pc[:, 0], pc[:, 2])
        t = np.arctan2(pc[:, 2], pc[:, 0])
        phi = np.stack([r, pc[:, 1], 0.5 * np.cos(4 * t) * r, 0.5 * np.sin(4 * t) * r], 1)
    F = np.sin(4.0 * phi @ W[: phi.shape[1]]) + noise * rng.standard_normal((len(pc), W.shape[1]))
    F /= np.linalg.norm(F, axis=1, keepdims=True)
    return F.astype(np.float32)


def make_pair(seed, D=32, category="chair", n_raw=10000, voxel=0.03, dropout=0.2, jitter=0.005,
              feat_noise=0.05, bulk=1.0, n_cad=None):
    """One query-CAD pair: dict(baseF, xyz0, posF, xyz1, T_gt, pos_sym).  xyz1/posF is the CAD in its
    canonical frame, xyz0/baseF the moved, thinned, jittered query (T_gt maps query -> CAD)."""
    rng = np.random.default_rng(seed)
    n_pts = max(15000, n_raw)
    if category == "chair":
        raw, kind, sym = make_chair(rng, n_pts, bulk), "mirror", 1
    else:
        raw, kind, sym = make_table(rng, n_pts), "rot4", 4
    W = rng.standard_normal((4, D))
    norm = normalize_cloud(raw, n_raw)
    if n_cad is None:
        cad = voxel_dedup(norm, voxel)
    else:
        # BASELINE.json configs[4]: a dense scan against a fixed-size CAD cloud, neither de-duplicated
        cad = norm[rng.choice(len(norm), n_cad, replace=False)]
        keep = np.ones(len(norm), bool)
        q_can = norm + jitter * rng.standard_normal(norm.shape).astype(np.float32)
        T = random_pose(rng)
        xyz0 = (q_can @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
        return dict(baseF=symmetric_features(q_can, W, kind, feat_noise, rng), xyz0=xyz0,
                    posF=symmetric_features(cad, W, kind, feat_noise, rng), xyz1=cad.astype(np.float32),
                    T_gt=np.linalg.inv(T), pos_sym=sym)
    keep = rng.random(len(cad)) > dropout
    q_can = cad[keep] + jitter * rng.standard_normal((int(keep.sum()), 3)).astype(np.float32)
    T = random_pose(rng)  # canonical -> query frame
    xyz0 = (q_can @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    return dict(baseF=symmetric_features(q_can, W, kind, feat_noise, rng), xyz0=xyz0,
                posF=symmetric_features(cad, W, kind, feat_noise, rng), xyz1=cad.astype(np.float32),
                T_gt=np.linalg.inv(T), pos_sym=sym)


def descriptor_db(seed, L=6778, Q=1, d=256, noise=0.3):
    """Unit-norm library descriptors and queries = normalize(row[g] + noise*N(0,1)/sqrt(d)...)."""
    rng = np.random.default_rng(seed)
    lib = rng.standard_normal((L, d)).astype(np.float32)
    lib /= np.linalg.norm(lib, axis=1, keepdims=True)
    gt = rng.integers(0, L, Q)
    q = lib[gt] + noise * rng.standard_normal((Q, d)).astype(np.float32) / np.sqrt(d)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return lib.astype(np.float32), q.astype(np.float32), gt
