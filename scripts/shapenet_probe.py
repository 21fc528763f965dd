#!/usr/bin/env python
"""get_symmetry_label on a 15 000-point cloud: device call vs the oracle's vectorised cKDTree port vs
the reference's per-point query loop (timed on a 1 500-query sample and scaled).  Diagnostic only."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from corsair_b200 import get_context, shapenet_eval as se, synth  # noqa: E402
from oracle import port_shapenet as ps  # noqa: E402


def main():
    get_context(0)
    rng = np.random.default_rng(5)
    pc = synth.normalize_cloud(synth.make_table(rng, 15000), 15000).astype(np.float32)
    for _ in range(3):
        lab = se.get_symmetry_label(pc, 0.1)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(20):
        lab = se.get_symmetry_label(pc, 0.1)
    dt = (time.perf_counter() - t) / 20
    t = time.perf_counter()
    ref = ps.get_symmetry_label(pc, 0.1)
    dt_port = time.perf_counter() - t
    from scipy.spatial import KDTree
    tree = KDTree(pc)
    t = time.perf_counter()
    for i in range(1500):
        tree.query(pc[i], k=1)
    per_query = (time.perf_counter() - t) / 1500
    n_rot = sum(1 for s in se.SYM_ORDERS for _ in range(1, s // 2 + 1))
    print(f"label {lab} (oracle {ref}); device {dt*1e3:.2f} ms for all {n_rot} rotations; oracle port (early exit, "
          f"vectorised cKDTree) {dt_port*1e3:.0f} ms; reference loop: {per_query*1e6:.0f} us per query -> "
          f"{per_query*2*15000*1e3:.0f} ms per rotation tried")


if __name__ == "__main__":
    main()
