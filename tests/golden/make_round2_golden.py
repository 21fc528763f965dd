"""Round-2 golden vectors, produced by the REFERENCE's own Python like make_golden.py (run here):

    python tests/golden/make_round2_golden.py

tests/golden/find_corr.npz   utils/eval_pose.py:28-45 find_corr with and without the seeded subsample
tests/golden/sympose_k4.npz  utils/symmetry.py:262-358 sym_pose for table pairs with pos_sym = 4 (K = 4, both
                             sweeps, nine registrations) and chair pairs (K = 2): best / vanilla pose and
                             Chamfer, the segmentation masks of both clouds, and — by recording every call the
                             reference makes to registration_based_on_corr — each registration's pose and
                             Chamfer distance and the index of the winning one.  The Open3D call is patched
                             with the restated RANSAC (C/OpenMP build of oracle/ransac_c.c, bit-identical to
                             oracle/ransac_np.py on counts: tests/test_oracle_golden.py) — `restated`, parity
                             unpinned for that call only.
tests/golden/symcut.npz      utils/symmetry.py:182-259 symmetric_cut4 masks of 10 seeded clouds per category
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from corsair_b200 import synth  # noqa: E402
from oracle import ransac_c, ransac_np, ref_import  # noqa: E402


def main():
    ep = ref_import.load("eval_pose")
    pp = ref_import.load("preprocess")
    sy = ref_import.load("symmetry")

    # ---- find_corr
    rng = np.random.default_rng(77)
    out = {}
    F0 = rng.standard_normal((600, 16)).astype(np.float32)
    F1 = rng.standard_normal((750, 16)).astype(np.float32)
    F0 /= np.linalg.norm(F0, axis=1, keepdims=True)
    F1 /= np.linalg.norm(F1, axis=1, keepdims=True)
    x0 = rng.uniform(-1, 1, (600, 3)).astype(np.float32)
    x1 = rng.uniform(-1, 1, (750, 3)).astype(np.float32)
    a, b = ep.find_corr(x0, x1, torch.from_numpy(F0), torch.from_numpy(F1), nn_max_n=250, subsample_size=-1)
    np.random.seed(5)
    c, d = ep.find_corr(x0, x1, torch.from_numpy(F0), torch.from_numpy(F1), nn_max_n=250, subsample_size=200)
    np.random.seed(6)
    e, f = ep.find_corr(x0, x1, torch.from_numpy(F0), torch.from_numpy(F1), nn_max_n=-1, subsample_size=700)
    out.update(F0=F0, F1=F1, xyz0=x0, xyz1=x1, full_0=a, full_1=b, sub200_0=c, sub200_1=d, sub700_0=e, sub700_1=f)
    np.savez_compressed(os.path.join(HERE, "find_corr.npz"), **out)

    # ---- sym_pose, K = 4 and K = 2, every registration recorded
    calls = []

    def reg(src, tgt, max_corr=0.03, seed=0):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        T = ransac_c.ransac(src.astype(np.float64), tgt.astype(np.float64), max_corr, ms, adaptive=True)["T"]
        calls.append(T.copy())
        return T

    ep.registration_based_on_corr = reg
    sy.registration_based_on_corr = reg
    sp = {}
    cases = [("table", 23, 4), ("table", 24, 4), ("table", 27, 12), ("chair", 25, 1), ("chair", 26, 1)]
    for ci, (cat, seed, pos_sym) in enumerate(cases):
        p = synth.make_pair(seed, 16, cat, n_raw=3000)
        K = 4 if pos_sym >= 2 else 2
        np.random.seed(31)
        bm = sy.symmetric_cut4(p["baseF"], p["xyz0"], K)
        pm = sy.symmetric_cut4(p["posF"], p["xyz1"], K)
        calls.clear()
        np.random.seed(31)
        Tb, cb, Tr, cr, ok = sy.sym_pose(p["baseF"], p["xyz0"], p["posF"], p["xyz1"], pos_sym, k_nn=5, max_corr=0.2, seed=31)
        Ts = np.stack(calls)
        ch = np.array([pp.chamfer_kdtree_1direction(pp.apply_transform(p["xyz0"], torch.from_numpy(T.astype(np.float32))), p["xyz1"])
                       for T in Ts])
        best = 0
        for i in range(1, len(ch)):
            if ch[best] > ch[i]:
                best = i
        assert ok and len(Ts) == (1 + 2 * K if pos_sym >= 2 else 1 + K) and ch[best] == cb
        pre = f"c{ci}_"
        sp.update({pre + "baseF": p["baseF"], pre + "xyz0": p["xyz0"], pre + "posF": p["posF"], pre + "xyz1": p["xyz1"],
                   pre + "pos_sym": pos_sym, pre + "base_masks": np.stack(bm), pre + "pos_masks": np.stack(pm),
                   pre + "T_best": Tb.numpy(), pre + "ch_best": cb, pre + "T_ransac": Tr.numpy(), pre + "ch_ransac": cr,
                   pre + "T_all": Ts, pre + "ch_all": ch, pre + "best_index": best})
        print(cat, seed, pos_sym, "registrations", len(Ts), "best", best, "chamfer", cb, "vanilla", cr)
    sp["n_cases"] = len(cases)
    np.savez_compressed(os.path.join(HERE, "sympose_k4.npz"), **sp)

    # ---- symmetric_cut4 on 10 seeded clouds per category
    sc = {}
    for cat, K in (("chair", 2), ("table", 4)):
        for i in range(10):
            p = synth.make_pair(300 + i, 16 if i % 2 else 32, cat, n_raw=3000)
            np.random.seed(100 + i)
            try:
                masks = np.stack(sy.symmetric_cut4(p["posF"], p["xyz1"], K))
            except AttributeError:
                masks = np.zeros((0, len(p["xyz1"])), bool)  # no model accepted: the reference raises
            sc[f"{cat}{i}_F"], sc[f"{cat}{i}_xyz"], sc[f"{cat}{i}_masks"] = p["posF"], p["xyz1"], masks
    np.savez_compressed(os.path.join(HERE, "symcut.npz"), **sc)
    print("written")


if __name__ == "__main__":
    main()
