"""Prints cb_knn's fallback-row count on the shapes the tests use (to set exact assertions)."""
import sys

import numpy as np

sys.path.insert(0, ".")
from corsair_b200 import find_nn, get_context  # noqa: E402


def unit_features(rng, n, d, clusters=0):
    f = rng.standard_normal((n, d)).astype(np.float32)
    if clusters:
        c = rng.standard_normal((clusters, d)).astype(np.float32)
        f = c[rng.integers(0, clusters, n)] + 0.3 * f
    f /= np.linalg.norm(f, axis=1, keepdims=True)
    return f.astype(np.float32)


ctx = get_context(0)
for clusters in (0, 25, 40):
    for d in (16, 32):
        for n0, n1, k in [(257, 301, 5), (4513, 3999, 5), (1000, 2000, 1), (777, 300, 12), (10000, 10000, 5),
                          (50000, 10000, 1), (50000, 10000, 5), (1203, 997, 1), (1203, 997, 5)]:
            fbs = []
            for seed in range(3):
                rng = np.random.default_rng(31 + seed)
                F0, F1 = unit_features(rng, n0, d, clusters), unit_features(rng, n1, d, clusters)
                find_nn.knn_device(F0, F1, k, return_distance=True)
                fbs.append(ctx.knn_last_fallback_rows())
            print(f"clusters {clusters:2d} d {d} {n0}x{n1} k {k}: fallback rows {fbs}", flush=True)
