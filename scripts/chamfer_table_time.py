"""Times the all-pairs Chamfer table kernel of the library named by CORSAIR_B200_LIB (build variants of
chamfer_matrix.cu's CM_* knobs) at 768 x 2048 (half table) and checks it bit for bit against the
two-sweep kernel of the same library on a small ragged case."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from corsair_b200 import pc_dist  # noqa: E402

rng = np.random.default_rng(5)
L = int(sys.argv[1]) if len(sys.argv) > 1 else 768
small = torch.from_numpy(rng.standard_normal((6, 2500, 3)).astype(np.float32)).cuda()
os.environ["CORSAIR_B200_CHAMFER_TWO_SWEEPS"] = "1"
ref = pc_dist.chamfer_rows_device(small).cpu().numpy()
os.environ["CORSAIR_B200_CHAMFER_TWO_SWEEPS"] = "0"
ok = np.array_equal(pc_dist.chamfer_rows_device(small).cpu().numpy(), ref)
pts = torch.from_numpy(rng.standard_normal((L, 2048, 3)).astype(np.float32) * 0.3).cuda()
best = 1e9
for _ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pc_dist.chamfer_rows_device(pts, half=True)
    torch.cuda.synchronize()
    best = min(best, time.perf_counter() - t0)
ev = L * (L - 1) // 2 * 2048.0 * 2048.0
print(f"{os.environ.get('CORSAIR_B200_LIB', 'default')}: bit-equal {ok}, {best:.4f} s, {ev / best / 1e12:.3f} T eval/s, "
      f"{ev / best / (148 * 1.965e9):.2f} eval/clk/SM, {8 * ev / best / 1e12:.1f} TFLOP/s")
