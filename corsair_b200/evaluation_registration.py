"""Drop-in for the registration sweep, result cache and metric tables of the reference's main app
(evaluation.py:286-383, 390-441).  The app keeps this as inline code of
``App.__init__`` / ``_load_data`` / ``_save_data``; here it is three functions over the arrays the
app already holds, so a maintainer replaces the loop body and keeps loading, inference and the GUI.

* ``register_queries``  the ``for i in range(len(self.base_outputs))`` loop (:297-334): sym_pose of
  every query against its chosen CAD, then eval_pose of both poses.
* ``save_results`` / ``load_results``  the nine ``.npy`` files of ``_save_data`` / ``_load_data``
  (:390-441), same names, same shapes, so caches are interchangeable with the reference's.
* ``summarize``  RTE / RRE tables and means (:336-358) and the two log blocks (:359-383).
"""
import os

import numpy as np
import torch

from .eval_pose import eval_pose
from .symmetry import sym_pose, sym_pose_many

KEYS = ("Ts_est_ransac", "Ts_est_best", "t_losses_ransac", "t_losses_sym", "r_losses_ransac", "r_losses_sym",
        "sym_ransac_success", "chamfer_dist_ransac", "chamfer_dist_sym")  # evaluation.py:288-296


def _np(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def register_queries(base_outputs, base_origins, base_Ts, lib_outputs, lib_origins, lib_Ts, sym_label, pos_idx,
                     k_nn=5, max_corr=0.2, *, workers=1, seeds=None, on_device=False):
    """evaluation.py:297-334.  ``pos_idx[i]`` is the CAD registered against query ``i``
    (``self.stat["top1_predict"]`` or ``["gt"]``, :302).  Returns a dict of the nine result lists.

    ``workers == 1`` and ``seeds is None`` is the reference's loop: one pair after the other, seed
    points from the global NumPy RNG.  ``seeds`` (one int per query) makes every pair's segmentation
    independent of the others (``rng=seed`` of symmetry.sym_pose), which ``workers > 1`` requires.
    ``on_device``: hand the inputs to sym_pose as they are (CUDA tensors stay on the device) instead
    of the reference's ``.detach().cpu().numpy()`` round trip (:306-309)."""
    n = len(base_outputs)
    if workers > 1 and seeds is None:
        raise ValueError("workers > 1 needs per-query seeds: the global NumPy stream is sequential")
    conv = (lambda x: x) if on_device else _np
    jobs = []
    for i in range(n):
        p = int(pos_idx[i])
        kw = {} if seeds is None else {"rng": int(seeds[i])}
        jobs.append(((conv(base_outputs[i]), conv(base_origins[i]), conv(lib_outputs[p]), conv(lib_origins[p]),
                      sym_label[p], k_nn, max_corr), kw))
    if workers > 1:
        poses = sym_pose_many(jobs, workers=workers)
    else:
        poses = [sym_pose(*a, **k) for a, k in jobs]
    out = {k: [] for k in KEYS}
    for i, (T_est_best, chamfer_dist_best, T_est_ransac, chamfer_dist_ransac, sym_success) in enumerate(poses):
        p = int(pos_idx[i])
        T0, T1, cad_sym = _np(base_Ts[i]), _np(lib_Ts[p]), sym_label[p]
        t_loss_ransac, r_loss_ransac = eval_pose(T_est_ransac, T0, T1, axis_symmetry=cad_sym)
        t_loss_sym, r_loss_sym = eval_pose(T_est_best, T0, T1, axis_symmetry=cad_sym)
        out["Ts_est_ransac"].append(T_est_ransac.detach().cpu().numpy())
        out["Ts_est_best"].append(T_est_best.detach().cpu().numpy())
        out["t_losses_ransac"].append(t_loss_ransac)
        out["t_losses_sym"].append(t_loss_sym)
        out["r_losses_ransac"].append(r_loss_ransac)
        out["r_losses_sym"].append(r_loss_sym)
        out["sym_ransac_success"].append(sym_success)
        out["chamfer_dist_ransac"].append(chamfer_dist_ransac)
        out["chamfer_dist_sym"].append(chamfer_dist_best)
    return out


def _suffix(register_top1):
    return "_top1.npy" if register_top1 else "_gt.npy"  # evaluation.py:401-404, 425-428


def save_results(cache_dir, category, register_top1, results):
    """evaluation.py:420-441: nine ``<name>_<category>_{top1,gt}.npy`` files, poses flattened to 16."""
    sfx = _suffix(register_top1)
    for k in KEYS:
        data = results[k]
        if k.startswith("Ts_est"):
            data = np.array([np.asarray(x).flatten() for x in data])
        np.save(os.path.join(cache_dir, f"{k}_{category}{sfx}"), data)


def load_results(cache_dir, category, register_top1):
    """evaluation.py:390-418: the dict of ``save_results`` (poses back as 4x4), or None when a file
    is missing (the app then recomputes)."""
    sfx = _suffix(register_top1)
    out = {}
    for k in KEYS:
        name = os.path.join(cache_dir, f"{k}_{category}{sfx}")
        if not os.path.exists(name):
            return None
        out[k] = np.load(name)
        if k.startswith("Ts_est"):
            out[k] = [x.reshape(4, 4) for x in out[k]]
    return out


def _compute_rte(t_losses):
    t = np.array(t_losses)
    return tuple(np.sum(t <= thr) / len(t) for thr in (0.02, 0.05, 0.10, 0.15))  # evaluation.py:336-341


def _compute_rre(r_losses):
    r = np.rad2deg(np.array(r_losses))
    return tuple(np.sum(r <= thr) / len(r) for thr in (5, 15, 45))  # evaluation.py:343-348


def summarize(results):
    """evaluation.py:350-383 -> (stats dict, the two log blocks + success-rate line as one string)."""
    s = {}
    for tag, tk, rk, ck in (("ransac", "t_losses_ransac", "r_losses_ransac", "chamfer_dist_ransac"),
                            ("sym", "t_losses_sym", "r_losses_sym", "chamfer_dist_sym")):
        s[f"t_loss_{tag}"] = np.mean(results[tk])
        s[f"rte_002_{tag}"], s[f"rte_005_{tag}"], s[f"rte_010_{tag}"], s[f"rte_015_{tag}"] = _compute_rte(results[tk])
        s[f"r_loss_{tag}"] = np.mean(results[rk])
        s[f"rre_005_{tag}"], s[f"rre_015_{tag}"], s[f"rre_045_{tag}"] = _compute_rre(results[rk])
        s[f"chamfer_dist_{tag}"] = np.mean(results[ck])
    s["sym_success_rate"] = np.mean(results["sym_ransac_success"])
    bar, dash = "=" * 66, "-" * 66
    text = ""
    for title, tag in (("vanilla ransac", "ransac"), ("sym ransac", "sym")):
        text += (f"\n{bar}\n{title}:\ntranslation error: {s[f't_loss_{tag}']},\n"
                 f"rte 0.02: {s[f'rte_002_{tag}']}, rte 0.05: {s[f'rte_005_{tag}']}, "
                 f"rte 0.10: {s[f'rte_010_{tag}']}, rte 0.15: {s[f'rte_015_{tag}']}\n{dash}\n"
                 f"rotation error: {s[f'r_loss_{tag}']},\n"
                 f"rre 5: {s[f'rre_005_{tag}']}, rre 15: {s[f'rre_015_{tag}']}, rre 45: {s[f'rre_045_{tag}']},\n"
                 f"chamfer distance: {s[f'chamfer_dist_{tag}']}")
    text += f"\n{bar}\n\nsym success rate: {s['sym_success_rate']}"
    return s, text
