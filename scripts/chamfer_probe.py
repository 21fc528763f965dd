#!/usr/bin/env python
"""Times the K4 kernels alone: the all-pairs table (cb_chamfer_matrix, P = 2048) and the two-direction
Chamfer of two 15 000-point clouds (cb_chamfer), in point-pair distance evaluations per second
against the FP32 roof (148 SMs x 128 lanes x SM clock; one evaluation = 8 flop in the survey's
counting, i.e. 3 sub + 3 multiply-add + min)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from corsair_b200 import get_context, pc_dist, preprocess  # noqa: E402


def timed(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


def main():
    ctx = get_context(0)
    rng = np.random.default_rng(0)
    L, P = (int(sys.argv[1]) if len(sys.argv) > 1 else 384), 2048
    pts = torch.from_numpy(rng.standard_normal((L, P, 3)).astype(np.float32)).cuda()
    dt = timed(lambda: pc_dist.chamfer_rows_device(pts, 0, L, ctx), 3)
    ev = L * (L - 1) * P * P  # unordered pairs x 2 directions
    print(f"chamfer table L={L} P={P}: {dt*1e3:.2f} ms, {ev/dt/1e12:.2f}e12 evaluations/s "
          f"({ev/dt/1e12/ (148*128*1.965e9/1e12) * 100:.1f} % of one evaluation per lane-clock)")
    a = torch.from_numpy(rng.standard_normal((15000, 3)).astype(np.float32)).cuda()
    b = torch.from_numpy((rng.standard_normal((15000, 3)) * 0.9).astype(np.float32)).cuda()
    dt = timed(lambda: preprocess.chamfer_device(a, b, None, True), 20)
    ev = 2 * 15000 * 15000
    print(f"chamfer 15000 x 15000 two directions: {dt*1e6:.1f} us, {ev/dt/1e12:.2f}e12 evaluations/s")
    a = torch.from_numpy(rng.standard_normal((3300, 3)).astype(np.float32)).cuda()
    b = torch.from_numpy((rng.standard_normal((4300, 3)) * 0.9).astype(np.float32)).cuda()
    dt = timed(lambda: preprocess.chamfer_device(a, b, None, False), 50)
    print(f"chamfer 3300 -> 4300 one direction (bench shape): {dt*1e6:.1f} us")


if __name__ == "__main__":
    main()
