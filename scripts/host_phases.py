#!/usr/bin/env python
"""Host-side timeline of one threaded bench step (perf_counter stamps relative to the step start):
where the host spends the time before the device gets its first registration kernels.  Diagnostic only."""
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

stamps = []
lock = threading.Lock()
T0 = [0.0]


def stamp(tag):
    with lock:
        stamps.append((time.perf_counter() - T0[0], threading.current_thread().name, tag))


def wrap(mod, name, tag=None):
    f = getattr(mod, name)

    def g(*a, **k):
        stamp(f"> {tag or name}")
        r = f(*a, **k)
        stamp(f"< {tag or name}")
        return r
    setattr(mod, name, g)


def main():
    from corsair_b200 import get_context, retrieval, symmetry
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    get_context(0)
    lib, cads, queries = bench.build_workload(0)
    lib_d = torch.from_numpy(lib).to(dev)
    cads_d = [dict(F=torch.from_numpy(c["F"]).to(dev), xyz=torch.from_numpy(c["xyz"]).to(dev), sym=c["sym"]) for c in cads]
    qs = [dict(desc=torch.from_numpy(q["desc"]).to(dev), F=torch.from_numpy(q["F"]).to(dev),
               xyz=torch.from_numpy(q["xyz"]).to(dev)) for q in queries]
    for i in range(4):
        bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[i])
    torch.cuda.synchronize()
    wrap(retrieval, "topk")
    wrap(symmetry, "sym_pose_many")
    wrap(symmetry, "sym_pose")
    wrap(symmetry, "_segments_device")
    wrap(symmetry, "cut_model_device")
    for i in range(2):
        stamps.clear()
        T0[0] = time.perf_counter()
        bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[4 + i])
        stamp("step done")
        torch.cuda.synchronize()
    for t, th, tag in sorted(stamps):
        print(f"{t*1e3:8.3f} ms  {th:18s} {tag}")


if __name__ == "__main__":
    main()
