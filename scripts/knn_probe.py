#!/usr/bin/env python
"""Times cb_knn alone (CUDA events) for a few shapes; used for ncu captures of knn_tc_kernel."""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from corsair_b200 import find_nn, get_context  # noqa: E402


def main():
    shapes = [(128, 256), (128, 1024), (3300, 4300), (10000, 10000), (50000, 10000)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]]
    ctx = get_context(0)
    rng = np.random.default_rng(0)
    D = int(os.environ.get("KNN_D", "32"))  # feature width
    for n0, n1 in shapes:
        F0 = rng.standard_normal((n0, D)).astype(np.float32)
        F1 = rng.standard_normal((n1, D)).astype(np.float32)
        F0 /= np.linalg.norm(F0, axis=1, keepdims=True)
        F1 /= np.linalg.norm(F1, axis=1, keepdims=True)
        a, b = torch.from_numpy(F0).cuda(), torch.from_numpy(F1).cuda()
        for _ in range(3):
            find_nn.knn_device(a, b, 5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            find_nn.knn_device(a, b, 5)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        print(f"knn {n0} x {n1} (D={D}, k=5): {us:.1f} us per call, {2.0 * n0 * n1 * D / us / 1e6:.1f} TFLOP/s algorithmic, "
              f"fallback rows {ctx.knn_last_fallback_rows()}")


if __name__ == "__main__":
    main()
