"""A/B of the all-pairs Chamfer table kernels: the one-tile kernel (both directions from one distance
tile) against the two-sweep kernel of rounds 1-2 (CORSAIR_B200_CHAMFER_TWO_SWEEPS=1): tables must be
bit-equal; prints kernel seconds of both at a C4-like shape."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from corsair_b200 import get_context, pc_dist  # noqa: E402

ctx = get_context(0)
rng = np.random.default_rng(5)


def run(pts, two, *args, **kw):
    os.environ["CORSAIR_B200_CHAMFER_TWO_SWEEPS"] = "1" if two else "0"
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = pc_dist.chamfer_rows_device(pts, *args, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy(), time.perf_counter() - t0


for L, P in ((7, 100), (5, 2049), (33, 2048), (12, 4266), (9, 31), (6, 1)):
    pts = torch.from_numpy(rng.standard_normal((L, P, 3)).astype(np.float32)).cuda()
    for args, kw in (((), {}), ((), {"half": True}), ((1, 4), {})):
        a, _ = run(pts, False, *args, **kw)
        b, _ = run(pts, True, *args, **kw)
        print(f"L {L} P {P} rows {args} {kw}: equal {np.array_equal(a, b)} max {np.abs(a - b).max():.3e}", flush=True)
        assert np.array_equal(a, b)
L, P = 768, 2048
pts = torch.from_numpy(rng.standard_normal((L, P, 3)).astype(np.float32) * 0.3).cuda()
for two in (False, True, False, True):
    out, sec = run(pts, two, half=True)
    pairs = L * (L - 1) // 2
    print(f"L {L} P {P} two_sweeps {two}: {sec:.3f} s, {pairs * P * P / sec / 1e12:.3f} T evaluations/s "
          f"({pairs * P * P * 8 / sec / 1e12:.1f} TFLOP/s at 8 flop)", flush=True)
