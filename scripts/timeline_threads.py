#!/usr/bin/env python
"""GPU occupancy of the threaded bench step: kineto trace of a few steps, then the union of the
kernel intervals (busy time), the summed kernel time (overlap) and the per-kernel totals.
Diagnostic only."""
import json
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
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        t0 = time.perf_counter()
        for i in range(steps):
            bench.gpu_step(retrieval, symmetry, lib_d, cads_d, qs[3 + i])
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / steps
    path = os.path.join(ROOT, "gpurun_out", "trace_threads.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    iv = sorted((e["ts"], e["ts"] + e["dur"]) for e in ev)
    busy, cur_s, cur_e = 0.0, None, None
    for s, e in iv:
        if cur_e is None or s > cur_e:
            if cur_e is not None:
                busy += cur_e - cur_s
            cur_s, cur_e = s, e
        else:
            cur_e = max(cur_e, e)
    busy += cur_e - cur_s
    total = sum(e["dur"] for e in ev)
    span = iv[-1][1] - iv[0][0]
    print(f"wall/step {wall*1e3:.2f} ms; per step: span {span/steps/1e3:.2f} ms, busy (union) {busy/steps/1e3:.2f} ms, "
          f"summed kernel time {total/steps/1e3:.2f} ms")
    agg = {}
    for e in ev:
        n = e["name"][:60]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += e["dur"]
    for n, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f"  {d/steps/1e3:8.3f} ms/step  {c/steps:7.1f} launches/step  avg {d/c:7.1f} us  {n}")
    # how much of the busy time has a scoring kernel resident, and what runs when none is
    def union(ivs):
        tot, cs, ce = 0.0, None, None
        for a, b in sorted(ivs):
            if ce is None or a > ce:
                if ce is not None:
                    tot += ce - cs
                cs, ce = a, b
            else:
                ce = max(ce, b)
        return tot + (ce - cs if ce is not None else 0.0)
    sc = [(e["ts"], e["ts"] + e["dur"]) for e in ev if "ransac_tc_score" in e["name"]]
    print(f"scoring kernel resident (union) {union(sc)/steps/1e3:.2f} ms/step of {busy/steps/1e3:.2f} busy; "
          f"alone it sums to {sum(b - a for a, b in sc)/steps/1e3:.2f} ms/step")
    # time with k scoring kernels in flight
    pts = sorted([(a, 1) for a, b in sc] + [(b, -1) for a, b in sc])
    hist, k, last_t = {}, 0, None
    for t, dlt in pts:
        if last_t is not None:
            hist[k] = hist.get(k, 0.0) + t - last_t
        k += dlt
        last_t = t
    print("  scoring kernels in flight: " + ", ".join(f"{k}: {v/steps/1e3:.2f} ms" for k, v in sorted(hist.items())))
    # idle gaps: where in the run, how long, and which kernels sit on either side
    evs = sorted(ev, key=lambda e: e["ts"])
    gaps, end, last = [], None, None
    for e in evs:
        if end is not None and e["ts"] > end:
            gaps.append((e["ts"] - end, end - evs[0]["ts"], last["name"][:40], e["name"][:40]))
        if end is None or e["ts"] + e["dur"] > end:
            end, last = e["ts"] + e["dur"], e
    print(f"idle gaps: {len(gaps)} totalling {sum(g[0] for g in gaps)/steps/1e3:.2f} ms/step; "
          f"{sum(1 for g in gaps if g[0] > 100)} above 100 us = {sum(g[0] for g in gaps if g[0] > 100)/steps/1e3:.2f} ms/step")
    for g in sorted(gaps, key=lambda g: -g[0])[:24]:
        print(f"  {g[0]:8.1f} us at t={g[1]/1e3:8.2f} ms   after {g[2]:40s} before {g[3]}")
    os.remove(path)


if __name__ == "__main__":
    main()
