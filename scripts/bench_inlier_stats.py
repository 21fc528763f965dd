#!/usr/bin/env python
"""Inlier ratio / winning hypothesis index / evaluated hypotheses of every registration of a few
bench steps (vanilla + pairings), via cb_register_pairings' stats output.  Diagnostic only."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    from corsair_b200 import get_context, retrieval, symmetry
    from corsair_b200._capi import ptr
    ctx = get_context(0)
    lib, cads, queries = bench.build_workload(0)
    rows = []
    for qi in range(3):
        q = queries[qi]
        idx, _ = retrieval.topk(q["desc"], lib, bench.TOPK)
        for c in idx[0]:
            cad = cads[int(c) % bench.CAD_POOL]
            fd = [ctx.dev(v, torch.float32) for v in (q["F"], q["xyz"], cad["F"], cad["xyz"])]
            K = 4 if cad["sym"] >= 2 else 2
            rs = np.random.RandomState(bench.SEED)
            A_perm, offA = symmetry._segments_device(fd[0], fd[1], K, 100, ctx, rs)
            B_perm, offB = symmetry._segments_device(fd[2], fd[3], K, 100, ctx, rs)
            off = torch.stack([offA, offB]).cpu().numpy()
            if off[0, K + 1] < 0 or off[1, K + 1] < 0:
                continue
            maps = symmetry._pair_maps(K, cad["sym"] >= 2)
            P = len(maps)
            T_all, ch = ctx.empty((1 + P, 4, 4), torch.float64), ctx.empty((1 + P,), torch.float64)
            st = ctx.empty((1 + P, 4), torch.float64)
            q_off, r_off = np.ascontiguousarray(off[0, :K + 1]), np.ascontiguousarray(off[1, :K + 1])
            n0, n1 = len(q["xyz"]), len(cad["xyz"])
            ctx.call("cb_register_pairings", ptr(fd[0]), ptr(fd[1]), n0, ptr(fd[2]), ptr(fd[3]), n1, fd[0].shape[1],
                     bench.K_NN, ptr(A_perm), q_off.ctypes.data, ptr(B_perm), r_off.ctypes.data, K, maps.ctypes.data, P,
                     bench.MAX_CORR, bench.SEED, 100000, 1, symmetry.CONFIDENCE, ptr(T_all), ptr(ch), ptr(st), ctx.stream)
            for r in st.cpu().numpy():
                rows.append((r[1] / (n0 * bench.K_NN), r[0], r[3]))
    rows = np.asarray(rows)
    print(f"{len(rows)} registrations: inlier ratio of the winner min {rows[:,0].min():.3f} median {np.median(rows[:,0]):.3f} "
          f"mean {rows[:,0].mean():.3f} max {rows[:,0].max():.3f}")
    print("ratio deciles:", np.round(np.quantile(rows[:, 0], np.linspace(0, 1, 11)), 3))
    print("winner index deciles:", np.quantile(rows[:, 1], np.linspace(0, 1, 11)).astype(int))
    print("hypotheses evaluated deciles:", np.quantile(rows[:, 2], np.linspace(0, 1, 11)).astype(int))


if __name__ == "__main__":
    main()
