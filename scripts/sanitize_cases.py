#!/usr/bin/env python
"""Small-shape cases of the three multi-role kernels (knn_tc_kernel, ransac_tc_score_kernel, the
cb_sym_pose / cb_sym_pose_many stream choreography) for compute-sanitizer (scripts/sanitize.sh).
Every case is checked against the oracle, so a sanitizer run is also a parity run."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from corsair_b200 import eval_pose, find_nn, get_context, symmetry, synth  # noqa: E402
from oracle import port, ransac_np  # noqa: E402


def main():
    ctx = get_context(0)
    rng = np.random.default_rng(3)
    # ---- kNN: tensor-core candidate kernel + re-rank + (forced) fallback rows
    for (n0, n1, d, k) in ((257, 301, 16, 5), (300, 777, 32, 12), (129, 513, 32, 1)):
        F0 = rng.standard_normal((n0, d)).astype(np.float32)
        F1 = rng.standard_normal((n1, d)).astype(np.float32)
        F1[7] = F1[3]  # an exact tie: goes through the exact fallback
        idx, dist = find_nn.knn_device(F0, F1, k, return_distance=True)
        ref_idx, ref_d2 = port.find_knn_exact(F0, F1, k)
        assert np.array_equal(idx.cpu().numpy(), ref_idx), "kNN indices"
        print(f"knn {n0}x{n1} d={d} k={k}: ok, fallback rows {ctx.knn_last_fallback_rows()}")
    # the shape and call of round 2's driver-side failure: find_nn_gpu(dist_type="L2") = k 1, metric 1
    F0 = rng.standard_normal((1203, 32)).astype(np.float32)
    F1 = rng.standard_normal((997, 32)).astype(np.float32)
    F0 /= np.linalg.norm(F0, axis=1, keepdims=True)
    F1 /= np.linalg.norm(F1, axis=1, keepdims=True)
    idx, dist = find_nn.knn_device(F0, F1, 1, metric=1, return_distance=True)
    ref_idx, ref_d2 = port.find_knn_exact(F0, F1, 1)
    assert np.array_equal(idx.cpu().numpy(), ref_idx), "kNN indices (L2, k=1)"
    want = np.sqrt(ref_d2.astype(np.float32) + np.float32(1e-7))
    assert np.allclose(dist.cpu().numpy(), want, rtol=2e-6, atol=0), "kNN L2 distances (k=1)"
    print(f"knn 1203x997 d=32 k=1 metric=L2: ok, fallback rows {ctx.knn_last_fallback_rows()}")
    # ---- RANSAC: tensor-core scoring with the in-band queue, fold, adaptive stop
    for (m, h) in ((999, 300), (3001, 700)):
        src = rng.uniform(-1, 1, (m, 3)).astype(np.float32)
        T = synth.random_pose(np.random.default_rng(4))
        tgt = (src @ T[:3, :3].T + T[:3, 3] + 0.01 * rng.standard_normal((m, 3))).astype(np.float32)
        tgt[::3] = rng.uniform(-2, 2, (len(tgt[::3]), 3)).astype(np.float32)
        ms = ransac_np.draw_minimal_sets(m, h, 10, seed=5)
        ref = ransac_np.ransac(src, tgt, 0.2, ms, adaptive=False, return_all=True)
        got = eval_pose.ransac_device(src, tgt, 0.2, minimal_sets=ms, adaptive=False, return_all=True)
        assert np.array_equal(got["counts"].cpu().numpy(), ref["counts"]), "inlier counts"
        assert int(got["stats"][0].item()) == int(ref["best_h"])
        ad = eval_pose.ransac_device(src, tgt, 0.2, seed=7, n_hypotheses=6000, adaptive=True)
        print(f"ransac m={m} h={h}: counts bit-exact, adaptive n_evaluated {int(ad['stats'][3].item())}")
    # ---- sym_pose: one call, and three pairs in flight from one thread
    pairs = [synth.make_pair(60 + i, 16, "table" if i == 1 else "chair", n_raw=1500) for i in range(3)]
    jobs = [((p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"]), dict(k_nn=5, max_corr=0.2, seed=3, rng=9 + i, n_hypotheses=3000))
            for i, p in enumerate(pairs)]
    one = [symmetry.sym_pose(*a, **k) for a, k in jobs]
    many = symmetry.sym_pose_many(jobs, workers=3)
    for a, b in zip(one, many):
        assert np.array_equal(a[0].numpy(), b[0].numpy()) and a[1] == b[1] and a[4] == b[4]
    print("sym_pose / sym_pose_many: ok", [round(r[1], 5) for r in many])
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
