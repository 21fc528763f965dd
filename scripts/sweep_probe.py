#!/usr/bin/env python
"""A mixed sweep through the public API: chairs (K = 2: 3 registrations per pair) and tables
(K = 4 with the second sweep: 9 registrations per pair), registered with sym_pose_many; checks that the
threaded results equal the sequential ones on a sample and reports pairs/s and device memory.
Diagnostic only."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from corsair_b200 import get_context, symmetry, synth  # noqa: E402


def main():
    n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    get_context(0)
    pairs = [synth.make_pair(500 + i, 32, "table" if i % 3 == 0 else "chair") for i in range(12)]
    dev = torch.device("cuda", 0)
    pd = [{k: (torch.from_numpy(v).to(dev) if isinstance(v, np.ndarray) and v.ndim == 2 and k != "T_gt" else v)
           for k, v in p.items()} for p in pairs]
    jobs = [((pd[i % 12]["baseF"], pd[i % 12]["xyz0"], pd[i % 12]["posF"], pd[i % 12]["xyz1"], pairs[i % 12]["pos_sym"]),
             dict(k_nn=5, max_corr=0.2, seed=i, rng=i)) for i in range(n_pairs)]
    print("pos_sym of the 12 distinct pairs:", [p["pos_sym"] for p in pairs])
    symmetry.sym_pose_many(jobs[:10], workers=5)
    torch.cuda.synchronize()
    m0 = torch.cuda.memory_allocated()
    t = time.perf_counter()
    out = symmetry.sym_pose_many(jobs, workers=5)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"{n_pairs} pairs in {dt*1e3:.1f} ms = {n_pairs/dt:.1f} pairs/s; torch memory delta {torch.cuda.memory_allocated()-m0} B; "
          f"success {sum(o[4] for o in out)}/{n_pairs}")
    seq = symmetry.sym_pose_many(jobs[:12], workers=1)
    same = all(torch.equal(a[0], b[0]) and a[1] == b[1] and torch.equal(a[2], b[2]) and a[3] == b[3] and a[4] == b[4]
               for a, b in zip(seq, out[:12]))
    print("threaded == sequential on the first 12:", same)
    free, total = torch.cuda.mem_get_info()
    print(f"device memory in use {(total-free)/2**30:.2f} GiB")


if __name__ == "__main__":
    main()
