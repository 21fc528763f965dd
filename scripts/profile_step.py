#!/usr/bin/env python
"""Host/device timeline of bench.py's step with torch.profiler (kineto/CUPTI; nsys is not in the
image).  Prints per-kernel device totals, the busiest CPU ops and the device-busy fraction of the
step, and writes the tables to gpurun_out/.  Diagnostic only: numbers taken under the profiler
are never bench values."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    from corsair_b200 import get_context, retrieval, symmetry
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    get_context(0)
    lib, cads, queries = bench.build_workload(0)
    lib_d = torch.from_numpy(lib).to(dev)
    cads_d = [dict(F=torch.from_numpy(c["F"]).to(dev), xyz=torch.from_numpy(c["xyz"]).to(dev), sym=c["sym"]) for c in cads]
    qs = [dict(desc=torch.from_numpy(q["desc"]).to(dev), F=torch.from_numpy(q["F"]).to(dev),
               xyz=torch.from_numpy(q["xyz"]).to(dev)) for q in queries]
    for i in range(3):
        bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[i])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[3 + i])
    torch.cuda.synchronize()
    wall_plain = (time.perf_counter() - t0) / steps
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        t0 = time.perf_counter()
        for i in range(steps):
            bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[3 + i])
        torch.cuda.synchronize()
        wall_prof = (time.perf_counter() - t0) / steps
    ev = prof.key_averages()
    dev_total = sum(e.self_device_time_total for e in ev) / steps / 1e3
    with open(os.path.join(out_dir, "timeline_summary.txt"), "w") as f:
        f.write(f"steps {steps}  wall/step plain {wall_plain*1e3:.2f} ms  under profiler {wall_prof*1e3:.2f} ms  "
                f"device busy/step {dev_total:.2f} ms\n\n")
        f.write(ev.table(sort_by="self_device_time_total", row_limit=40, max_name_column_width=70))
        f.write("\n\n")
        f.write(ev.table(sort_by="self_cpu_time_total", row_limit=40, max_name_column_width=70))
    print(open(os.path.join(out_dir, "timeline_summary.txt")).read()[:6000])


if __name__ == "__main__":
    main()
