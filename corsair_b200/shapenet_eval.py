"""Drop-in for the compute functions of the reference's ShapeNet benchmark app
(evaluation-shapenet.py): ``load_pc`` (:70-76), ``generate_random_pose``
(:79-94), ``generate_test_pc_pair`` (:116-120), ``chamfer_max`` (:122-134), ``test_symmetry_label``
(:137-147), ``get_symmetry_label`` (:150-154) and ``registration_worker`` (:243-275), plus
``register_many`` for the ``joblib.Parallel`` loop of ``generate_test_results`` (:341-343).

The reference finds an object's rotational-symmetry order with a per-point Python loop over a scipy
KD-tree (2 x 15 000 queries for each of up to 17 rotations); here all 17 rotated copies go through
ONE C call (cb_self_hausdorff: brute-force NN on the device, the winning neighbour's distance
recomputed in float64), and the sequential rule of :150-154 is applied to the 17 maxima.

Model loading, MinkowskiEngine quantisation / inference and the GUI of that app are out of scope
(DESIGN.md §7): the producer hands ``registration_worker`` the same six arrays as the reference's.
"""
import numpy as np
import torch

from . import _capi
from ._capi import ptr
from .eval_pose import eval_pose
from .symmetry import sym_pose


SYM_ORDERS = (12, 8, 6, 4, 3, 2, 1)  # evaluation-shapenet.py:151 (1 = no symmetry)


def euler2mat(ai, aj, ak):
    """transforms3d.euler.euler2mat(ai, aj, ak) for its default 'sxyz' axes, same closed form."""
    si, sj, sk = np.sin(ai), np.sin(aj), np.sin(ak)
    ci, cj, ck = np.cos(ai), np.cos(aj), np.cos(ak)
    cc, cs = ci * ck, ci * sk
    sc, ss = si * ck, si * sk
    M = np.eye(3)
    M[0, 0], M[0, 1], M[0, 2] = cj * ck, sj * sc - cs, sj * cc + ss
    M[1, 0], M[1, 1], M[1, 2] = cj * sk, sj * ss + cc, sj * cs - sc
    M[2, 0], M[2, 1], M[2, 2] = -sj, cj * si, cj * ci
    return M


def load_pc(path):
    """evaluation-shapenet.py:70-76: centre on the mean, scale to unit max norm."""
    pc = np.load(path)
    pc -= pc.mean(axis=0, keepdims=True)
    pc /= np.linalg.norm(pc, axis=1).max()
    return pc


def generate_random_pose(config):
    """evaluation-shapenet.py:79-94: six uniform draws from the global NumPy RNG, in that order."""
    roll = np.deg2rad(np.random.uniform(-config.max_roll_deg, config.max_roll_deg))
    pitch = np.deg2rad(np.random.uniform(-config.max_pitch_deg, config.max_pitch_deg))
    yaw = np.deg2rad(np.random.uniform(-config.max_yaw_deg, config.max_yaw_deg))
    pose = np.eye(4)
    pose[0, 3] = np.random.uniform(-config.max_translation_x, config.max_translation_x)
    pose[1, 3] = np.random.uniform(-config.max_translation_y, config.max_translation_y)
    pose[2, 3] = np.random.uniform(-config.max_translation_z, config.max_translation_z)
    pose[:3, :3] = euler2mat(roll, pitch, yaw)
    return pose


def generate_test_pc_pair(config, pc_file):
    """evaluation-shapenet.py:116-120."""
    pc = load_pc(pc_file)
    pose = generate_random_pose(config)
    return pc, pc @ pose[:3, :3].T + pose[:3, [3]].T, pose


def chamfer_max(pc0, pc0_tree, pc1, pc1_tree):
    """evaluation-shapenet.py:122-134: the larger of the two directed Hausdorff distances.  The KD-tree
    arguments of the reference's signature are accepted and ignored."""
    ctx = _capi.get_context()
    a, b = ctx.dev(pc0, torch.float32), ctx.dev(pc1, torch.float32)
    out = ctx.empty((2,), torch.float64)
    ctx.call("cb_nn_max", ptr(a), a.shape[0], ptr(b), b.shape[0], None, ptr(out), ctx.stream)
    ctx.call("cb_nn_max", ptr(b), b.shape[0], ptr(a), a.shape[0], None, ptr(out[1:]), ctx.stream)
    return float(out.max().item())


def _rotation_poses(orders):
    """[(order, i, 4x4)] for every rotation test_symmetry_label tries (:139-142), in its order."""
    poses = []
    for sym_label in orders:
        for i in range(1, int(sym_label / 2) + 1):
            T = np.eye(4)
            T[:3, :3] = euler2mat(0, i * (2 * np.pi) / sym_label, 0)  # in ShapeNet, y is up
            poses.append((sym_label, i, T))
    return poses


def symmetry_errors(pc, orders=SYM_ORDERS):
    """chamfer_max(pc, pc @ R.T) for every rotation of every order in one device call ->
    {order: [error for i = 1 .. order // 2]}."""
    ctx = _capi.get_context()
    poses = _rotation_poses(orders)
    errs = {o: [] for o in orders}
    if not poses:
        return errs
    xyz = ctx.dev(pc, torch.float32)
    T = ctx.dev(np.stack([p[2] for p in poses]), torch.float64)
    out = ctx.empty((len(poses),), torch.float64)
    ctx.call("cb_self_hausdorff", ptr(xyz), xyz.shape[0], ptr(T), len(poses), ptr(out), ctx.stream)
    for (o, _, _), e in zip(poses, out.cpu().numpy()):
        errs[o].append(float(e))
    return errs


def test_symmetry_label(sym_label, pc, cd_threshold):
    """evaluation-shapenet.py:137-147."""
    return all(e <= cd_threshold for e in symmetry_errors(pc, (sym_label,))[sym_label])


test_symmetry_label.__test__ = False  # a reference function name, not a pytest test


def get_symmetry_label(pc, cd_threshold):
    """evaluation-shapenet.py:150-154: the highest order in 12, 8, 6, 4, 3, 2, 1 all of whose
    rotations keep the cloud within cd_threshold of itself (1: none needed)."""
    errs = symmetry_errors(pc)
    for sym_label in SYM_ORDERS:
        if all(e <= cd_threshold for e in errs[sym_label]):
            return sym_label
    return 0


def registration_worker(feats_pc, feats_pc_transformed, coords_pc, pc_quant, pc_transformed_quant, pose_gt, *,
                        k_nn=5, max_corr=0.4, seed=0, cd_threshold=0.1, rng=None):
    """evaluation-shapenet.py:243-275 (``self.config.k_nn / max_corr / random_seed`` as keywords;
    ``rng`` as in symmetry.sym_pose).  Returns the reference's result dict."""
    symmetry_label = get_symmetry_label(coords_pc, cd_threshold)
    T_est_sym, chamfer_dist_sym, T_est_ransac, chamfer_dist_ransac, sym_success = sym_pose(
        feats_pc, pc_quant, feats_pc_transformed, pc_transformed_quant, symmetry_label, k_nn, max_corr, seed=seed,
        rng=rng)
    rte_sym, rre_sym = eval_pose(T_est_sym, np.eye(4), pose_gt, axis_symmetry=symmetry_label)
    rte_ransac, rre_ransac = eval_pose(T_est_ransac, np.eye(4), pose_gt, axis_symmetry=symmetry_label)
    return dict(symmetry_label=symmetry_label, T_est_sym=T_est_sym, chamfer_dist_sym=chamfer_dist_sym,
                T_est_ransac=T_est_ransac, chamfer_dist_ransac=chamfer_dist_ransac, sym_success=sym_success,
                rte_sym=rte_sym, rre_sym=rre_sym, rte_ransac=rte_ransac, rre_ransac=rre_ransac)


def register_many(items, workers=8, **kwargs):
    """The ``Parallel(n_jobs=-1)(delayed(self.registration_worker)(*x) for x in producer)`` loop of
    evaluation-shapenet.py:341-343 without host threads: the symmetry labels of all items first (one
    cb_self_hausdorff call each), then ONE symmetry.sym_pose_many call that keeps ``workers`` pairs in
    flight (cb_sym_pose_many), then eval_pose of both poses per item.  ``items``: iterable of the
    producer's 6-tuples; keywords as registration_worker; ``rng`` defaults to the int ``seed`` so that
    every pair's segmentation draws are its own (an explicit RandomState keeps the sequential loop).
    Results equal registration_worker's, in item order."""
    from .symmetry import sym_pose_many
    kwargs.setdefault("rng", int(kwargs.get("seed", 0)))
    items = list(items)
    if workers <= 1 or len(items) <= 1 or not isinstance(kwargs["rng"], (int, np.integer)):
        return [registration_worker(*x, **kwargs) for x in items]
    k_nn, max_corr = kwargs.get("k_nn", 5), kwargs.get("max_corr", 0.4)
    seed, cd_threshold = kwargs.get("seed", 0), kwargs.get("cd_threshold", 0.1)
    labels = [get_symmetry_label(x[2], cd_threshold) for x in items]
    jobs = [((x[0], x[3], x[1], x[4], lab), dict(k_nn=k_nn, max_corr=max_corr, seed=seed, rng=int(kwargs["rng"])))
            for x, lab in zip(items, labels)]
    out = []
    for x, lab, (T_sym, ch_sym, T_r, ch_r, ok) in zip(items, labels, sym_pose_many(jobs, workers=workers)):
        rte_sym, rre_sym = eval_pose(T_sym, np.eye(4), x[5], axis_symmetry=lab)
        rte_ransac, rre_ransac = eval_pose(T_r, np.eye(4), x[5], axis_symmetry=lab)
        out.append(dict(symmetry_label=lab, T_est_sym=T_sym, chamfer_dist_sym=ch_sym, T_est_ransac=T_r,
                        chamfer_dist_ransac=ch_r, sym_success=ok, rte_sym=rte_sym, rre_sym=rre_sym,
                        rte_ransac=rte_ransac, rre_ransac=rre_ransac))
    return out
