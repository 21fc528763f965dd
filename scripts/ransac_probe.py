#!/usr/bin/env python
"""Times cb_ransac alone on one bench-shaped pair (kNN correspondences of a synthetic chair pair):
fixed H hypotheses, no adaptive stop, both scoring variants.  Used for ncu captures."""
import sys
import os
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from corsair_b200 import eval_pose, find_nn, get_context, synth  # noqa: E402


def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [2, 1]
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    ctx = get_context(0)
    p = synth.make_pair(100, 32, "chair")
    idx = find_nn.knn_device(p["baseF"], p["posF"], 5)
    n0 = len(p["baseF"])
    si = torch.repeat_interleave(torch.arange(n0, device="cuda"), 5)
    ti = idx.reshape(-1)
    x0 = torch.from_numpy(p["xyz0"]).cuda()
    x1 = torch.from_numpy(p["xyz1"]).cuda()
    print(f"M = {len(si)} correspondences, H = {H}")
    for v in variants:
        ctx.set_ransac_variant(v)
        for _ in range(3):
            r = eval_pose.ransac_device(x0, x1, 0.2, seed=31, src_idx=si, tgt_idx=ti, n_hypotheses=H, adaptive=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            r = eval_pose.ransac_device(x0, x1, 0.2, seed=31, src_idx=si, tgt_idx=ti, n_hypotheses=H, adaptive=False)
        e1.record()
        torch.cuda.synchronize()
        st = r["stats"].cpu().numpy()
        print(f"variant {ctx.ransac_variant()}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per cb_ransac, "
              f"best_h {int(st[0])} inliers {int(st[1])} sum {st[2]:.9f}")
    ctx.set_ransac_variant(0)


if __name__ == "__main__":
    main()
