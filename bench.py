#!/usr/bin/env python
"""Benchmark of the correspondence-and-registration hot path (BASELINE.json metric: query-CAD
pairs registered per second).

One step = one query of BASELINE.json configs[1]: a 256-d global descriptor is ranked against a
synthetic ShapeNet-chair-sized library (6778 CADs), and the query scan is registered against each
of the top-10 candidates with the full symmetry-aided pipeline (sym_pose: feature 5-NN ->
correspondence RANSAC (budget 100000, Open3D's adaptive stop) -> Chamfer, two symmetry
segmentations, K more {per-segment kNN, RANSAC, Chamfer} pairings).  So a step registers 10 pairs.

  value  device-resident inputs (query + CAD library already in HBM), timed with CUDA events
  e2e    the same step through the reference-facing Python API with HOST NumPy buffers
         (host->device copies and the device->host result reads inside the timed region)
Both register a step's 10 independent pairs through symmetry.sym_pose_many on WORKERS host threads
(own cb_ctx + CUDA stream each), like the reference's thread_map loops; the roofline / kernel-class
numbers come from a separate instrumented pass that runs the pairs one after the other.

`python bench.py --impl reference` times the CPU path instead (the oracle's restatement of the
reference's algorithm on all host cores), one pair per step as a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

L_DB, D_DESC, TOPK, D_FEAT, CAD_POOL, N_QUERIES = 6778, 256, 10, 32, 32, 8
K_NN, MAX_CORR, SEED = 5, 0.2, 31
WORKLOAD = ("C2: 1 query descriptor vs 6778-CAD chair DB, top-10, then sym_pose registration of each "
            "candidate (10 pairs/step)")


def build_workload(rank):
    """Seeded synthetic inputs (corsair_b200/synth.py): the CAD library holds CAD_POOL distinct
    procedural chairs (DB row i uses chair i % CAD_POOL); each query is a moved, thinned, jittered
    copy of its ground-truth CAD with a noisy copy of that row's descriptor."""
    from corsair_b200 import synth
    lib, _, _ = synth.descriptor_db(1000, L=L_DB, Q=1, d=D_DESC)
    pairs = [synth.make_pair(100 + c, D_FEAT, "chair") for c in range(CAD_POOL)]
    cads = [dict(F=p["posF"], xyz=p["xyz1"], sym=p["pos_sym"]) for p in pairs]
    rng = np.random.default_rng(7000 + rank)
    queries = []
    for i in range(N_QUERIES):
        g = int(rng.integers(0, L_DB))
        desc = lib[g] + 0.3 * rng.standard_normal(D_DESC).astype(np.float32) / np.sqrt(D_DESC)
        desc = (desc / np.linalg.norm(desc)).astype(np.float32)[None]
        src = pairs[g % CAD_POOL]
        keep = rng.random(len(src["xyz1"])) > 0.2
        T = synth.random_pose(rng)
        xyz0 = ((src["xyz1"][keep] + 0.005 * rng.standard_normal((int(keep.sum()), 3))) @ T[:3, :3].T
                + T[:3, 3]).astype(np.float32)
        F0 = src["posF"][keep] + 0.05 * rng.standard_normal((int(keep.sum()), D_FEAT)).astype(np.float32)
        F0 = (F0 / np.linalg.norm(F0, axis=1, keepdims=True)).astype(np.float32)
        queries.append(dict(desc=desc, F=F0, xyz=xyz0, gt=g))
    return lib, cads, queries


class ClockSampler:
    """nvidia-smi sampler for the `clocks` object (B200_PROFILING.md recipe)."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_pair(q, cad):
    """One registered pair on the CPU: the oracle's restatement of sym_pose (reference Python
    algorithm: scipy KD-tree kNN with workers=-1, sklearn KMeans segmentation, KD-tree Chamfer)
    with the Open3D RANSAC call restated in C/OpenMP on the same seeded minimal sets."""
    from oracle import port_symmetry, ransac_c, ransac_np

    def reg(src, tgt, max_corr, seed):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        return ransac_c.ransac(src, tgt, max_corr, ms, adaptive=True)["T"]

    np.random.seed(SEED)
    return port_symmetry.sym_pose(q["F"], q["xyz"], cad["F"], cad["xyz"], cad["sym"], k_nn=K_NN, max_corr=MAX_CORR,
                                  seed=SEED, registration=reg)


def cpu_step(lib, cads, q, n_pairs):
    from scipy.spatial.distance import cdist
    order = np.argsort(cdist(q["desc"], lib), axis=1, kind="stable")[0, :TOPK]
    for c in order[:n_pairs]:
        cpu_pair(q, cads[int(c) % CAD_POOL])
    return n_pairs


def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import ransac_c
    ransac_c.build()
    lib, cads, queries = build_workload(0)
    for i in range(args.warmup):
        cpu_step(lib, cads, queries[i % N_QUERIES], 1)
    t0 = time.perf_counter()
    pairs = 0
    for i in range(args.steps):
        pairs += cpu_step(lib, cads, queries[(args.warmup + i) % N_QUERIES], 1)
    dt = time.perf_counter() - t0
    v = pairs / dt
    cores = os.cpu_count()
    print(json.dumps({
        "impl": "reference", "metric": "query-CAD pairs registered/sec", "value": v, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": "retrieval + the first of the 10 candidate pairs per step"},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port",
                         "sample": f"{pairs} pairs (1 per step) of the same workload, all host threads"},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


WORKERS = int(os.environ.get("CB_BENCH_WORKERS", "5"))  # host threads (each with its own cb_ctx + CUDA stream) registering a step's pairs


def gpu_step(retrieval, symmetry, lib_in, cads_in, q, workers=WORKERS):
    """One step through the public drop-in API; inputs are whatever the caller holds (CUDA tensors
    for `value`, host NumPy for `e2e`).  The 10 candidate pairs are independent; sym_pose_many
    registers them on `workers` host threads (workers=1: one after the other, as instrumented)."""
    idx, _ = retrieval.topk(q["desc"], lib_in, TOPK)          # D2H of the 10 candidate ids
    jobs = []
    for c in idx[0]:
        cad = cads_in[int(c) % CAD_POOL]
        jobs.append(((q["F"], q["xyz"], cad["F"], cad["xyz"], cad["sym"]),
                     dict(k_nn=K_NN, max_corr=MAX_CORR, seed=SEED, rng=SEED)))
    return symmetry.sym_pose_many(jobs, workers=workers)


def nbytes(*arrays):
    return int(sum(a.size * a.itemsize for a in arrays))


def run_ours(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from corsair_b200 import get_context, retrieval, symmetry
    ctx = get_context(local_rank)
    lib, cads, queries = build_workload(rank)
    lib_d = torch.from_numpy(lib).to(dev)
    cads_d = [dict(F=torch.from_numpy(c["F"]).to(dev), xyz=torch.from_numpy(c["xyz"]).to(dev), sym=c["sym"]) for c in cads]
    queries_d = [dict(desc=torch.from_numpy(q["desc"]).to(dev), F=torch.from_numpy(q["F"]).to(dev),
                      xyz=torch.from_numpy(q["xyz"]).to(dev)) for q in queries]
    queries_h = [dict(desc=torch.from_numpy(q["desc"]).pin_memory().numpy(), F=torch.from_numpy(q["F"]).pin_memory().numpy(),
                      xyz=torch.from_numpy(q["xyz"]).pin_memory().numpy()) for q in queries]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, lib_in, cads_in, qs, steps, warmup):
        for i in range(warmup):
            fn(retrieval, symmetry, lib_in, cads_in, qs[i % N_QUERIES])
        barrier()
        evs = []
        for i in range(steps):
            flush.fill_(i & 0xff)  # L2 flush between timed iterations (outside the event pair)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(retrieval, symmetry, lib_in, cads_in, qs[(warmup + i) % N_QUERIES])
            e1.record()
            evs.append((e0, e1))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    import functools
    from corsair_b200 import _capi
    # (1) instrumented pass, one pair after the other on one stream: per-class CUDA-event timing of
    #     the library's kernels (meaningless when several streams overlap) -> roofline, shares
    seq_step = functools.partial(gpu_step, workers=1)
    seq_step(retrieval, symmetry, lib_d, cads_d, queries_d[0])
    ctx.timing_reset()
    ctx.timing(True)
    n_inst = max(1, min(2, args.steps))
    for i in range(n_inst):
        seq_step(retrieval, symmetry, lib_d, cads_d, queries_d[(1 + i) % N_QUERIES])
    ctx.timing(False)
    classes = ctx.timing_read()
    ms_seq = timed(seq_step, lib_d, cads_d, queries_d, n_inst, 0)
    # (2) the timed runs: WORKERS host threads, each with its own context and stream
    for w in range(2):
        gpu_step(retrieval, symmetry, lib_d, cads_d, queries_d[w])  # create the worker contexts
    count = lambda: sum(c.launches() for c in _capi.all_contexts())
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev = timed(gpu_step, lib_d, cads_d, queries_d, args.steps, args.warmup)
    clocks = sampler.stop()
    barrier()
    l0 = count()
    gpu_step(retrieval, symmetry, lib_d, cads_d, queries_d[0])
    barrier()
    launches = (count() - l0) * args.steps  # library kernels of one step x timed steps
    ms_e2e = timed(gpu_step, lib, cads, queries_h, args.steps, args.warmup)
    # the same pairs without the per-query step boundary (BASELINE.json configs[2]'s shape at N_QUERIES
    # queries per rank): one batched retrieval, then all N_QUERIES x TOPK pairs in ONE sym_pose_many call
    def sweep():
        desc = torch.cat([q["desc"] for q in queries_d])
        idx, _ = retrieval.topk(desc, lib_d, TOPK)
        jobs = []
        for qi, q in enumerate(queries_d):
            for c in idx[qi]:
                cad = cads_d[int(c) % CAD_POOL]
                jobs.append(((q["F"], q["xyz"], cad["F"], cad["xyz"], cad["sym"]),
                             dict(k_nn=K_NN, max_corr=MAX_CORR, seed=SEED, rng=SEED)))
        return symmetry.sym_pose_many(jobs, workers=WORKERS)
    sweep()
    barrier()
    se0, se1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.fill_(1)
    se0.record()
    sweep()
    se1.record()
    barrier()
    t_sw = torch.tensor([se0.elapsed_time(se1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_sw, op=dist.ReduceOp.MAX)
    sweep_pairs_s = world * N_QUERIES * TOPK / (float(t_sw.item()) * 1e-3)
    pairs = world * args.steps * TOPK
    value = pairs / (ms_dev * 1e-3)
    e2e = pairs / (ms_e2e * 1e-3)
    if rank != 0:
        return

    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk_path):
        peaks = json.load(open(pk_path))
    hbm = peaks.get("hbm_gbs", 6650.0)
    bf16 = peaks.get("bf16_tflops_sustained", 1400.0)
    src = "measured" if peaks else "fallback"
    traffic = {}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        traffic = json.load(open(tr_path))
    tracked = {k: v for k, v in classes.items() if v[2] > 0}
    top = max(("knn_candidates", "ransac_score"), key=lambda k: tracked.get(k, (0, 0, 0))[0])
    ms_k, work_k, n_k = tracked[top]
    if top == "ransac_score":
        roof = {"bound": "hbm", "kernel": "ransac_tc_score_kernel (" + ctx.ransac_variant() + ")",
                "achieved": work_k / (ms_k * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                "note": f"effective streamed bytes (hypotheses scored x M x 32 B, BASELINE.json's definition); peak = {src} "
                        "HBM copy bandwidth.  The correspondences live in L2/shared memory, so frac > 1; the kernel is "
                        "a split-bf16 GEMM (K = 96) whose accumulators are drained by tcgen05.ld beside the next tile's "
                        "MMAs: evals_per_clk_per_sm against the tensor pipe's 42.7 (12 MMAs x 64 clk per 32768 "
                        "evaluations) is tensor_frac; with its arithmetic compiled out the kernel still needs 70 % of "
                        "its time (profiles/r01_tmem_ld_microbench.txt)",
                "evals_per_s": work_k / 32.0 / (ms_k * 1e-3)}
        roof["evals_per_clk_per_sm"] = roof["evals_per_s"] / (148 * (clocks.get("sm_mhz") or 1965.0) * 1e6)
        roof["tensor_frac"] = roof["evals_per_clk_per_sm"] / (128.0 * 128.0 * 2 / (12 * 64.0))
    else:
        roof = {"bound": "tensor", "kernel": "knn candidates (" + ctx.knn_variant() + ")",
                "achieved": work_k / (ms_k * 1e-3) / 1e12, "peak": bf16 / 3.0, "unit": "TFLOP/s",
                "note": f"2*N0*N1*D flop; peak = {src} sustained bf16 dense / 3 (split-bf16 issues 3 products)"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["traffic"] = traffic.get(top)
    roof["launches"] = n_k
    roof["avg_launch_us"] = 1e3 * ms_k / max(1, n_k)
    total_tracked = sum(v[0] for v in tracked.values())
    shares = {k: {"ms": round(v[0], 3), "launches": v[2], "share_of_tracked": round(v[0] / max(total_tracked, 1e-9), 4)}
              for k, v in tracked.items()}

    # the metric's second half: the feature-kNN distance GEMM against the split-precision tensor peak, on
    # the largest single-pair shape BASELINE.json names (configs[4]: 50k-point scan vs a 10k-point CAD)
    from corsair_b200 import find_nn
    g = torch.Generator(device=dev).manual_seed(5)
    fa = torch.nn.functional.normalize(torch.randn(50000, D_FEAT, device=dev, generator=g), dim=1)
    fb = torch.nn.functional.normalize(torch.randn(10000, D_FEAT, device=dev, generator=g), dim=1)
    for _ in range(3):
        find_nn.knn_device(fa, fb, K_NN)
    ke0, ke1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ke0.record()
    for _ in range(10):
        find_nn.knn_device(fa, fb, K_NN)
    ke1.record()
    torch.cuda.synchronize()
    knn_tf = 2.0 * 50000 * 10000 * D_FEAT / (ke0.elapsed_time(ke1) / 10 * 1e-3) / 1e12
    knn_tensor = {"shape": f"50000 x 10000, D = {D_FEAT}, k = {K_NN}", "achieved": knn_tf, "unit": "TFLOP/s",
                  "peak": bf16 / 3.0, "frac": knn_tf / (bf16 / 3.0), "us_per_call": ke0.elapsed_time(ke1) / 10 * 1e3,
                  "note": "whole cb_knn call (norms, packing, two-pass tcgen05 candidate kernel, exact re-rank), 2*N0*N1*D "
                          f"flop; peak = {src} sustained bf16 dense / 3 products.  Every score is read out of tensor "
                          "memory twice, beside the MMAs' accumulator traffic (measured 88-125 B/clk/SM), which caps this "
                          "near 22-31 % of that peak (DESIGN.md K1)"}
    del fa, fb

    cpu_base = None
    if world == 1:
        from oracle import ransac_c
        ransac_c.build()
        cpu_step(lib, cads, queries[0], 1)  # warm the KD-tree / BLAS / OpenMP pools
        t0 = time.perf_counter()
        n = 0
        while n < 3 or (time.perf_counter() - t0 < 12 and n < 10):
            n += cpu_step(lib, cads, queries[n % N_QUERIES], 1)
        dt = time.perf_counter() - t0
        cpu_base = {"value": n / dt, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port",
                    "sample": f"{n} pairs of the same workload (retrieval + first candidate per query), all host threads"}

    h2d = nbytes(queries[0]["desc"]) + TOPK * (nbytes(queries[0]["F"], queries[0]["xyz"]) +
                                               int(np.mean([nbytes(c["F"], c["xyz"]) for c in cads]))) + nbytes(lib)
    # per pair: the two seed-row draws and k-means++ draws go up (cb_sym_pose's staging), the segment
    # offsets and the 1 + P poses / Chamfer values come back
    h2d += TOPK * (2 * 100 * 4 + 10 * 4 + 10 * 1 * 2 * 8)
    d2h = TOPK * 8 * 2 + TOPK * (16 * 4 + 3 * 17 * 8)
    print(json.dumps({
        "metric": "query-CAD pairs registered/sec", "value": value, "unit": "pairs/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 search / f64 re-rank, fit and in-band scoring",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "db": L_DB, "descriptor_dim": D_DESC, "feature_dim": D_FEAT, "k_nn": K_NN,
                   "ransac": "budget 100000, adaptive stop (confidence 0.999), ransac_n 10, max_corr 0.2",
                   "points_per_cloud": "2.6k-4.3k after 0.03 voxel dedup of 10k", "cad_pool": CAD_POOL,
                   "l2": "256 MiB flush between timed steps", "host_threads": WORKERS,
                   "sequential_ms_per_step": ms_seq / n_inst, "sharding": f"queries sharded over {world} rank(s), no data-path collective"},
        "e2e": {"value": e2e, "unit": "pairs/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "roofline": roof, "knn_tensor": knn_tensor,
        "sweep": {"value": sweep_pairs_s, "unit": "pairs/s",
                  "workload": f"the same pairs without the per-query step boundary: {N_QUERIES} queries per rank, one batched "
                              f"retrieval, {N_QUERIES * TOPK} pairs in one sym_pose_many call (configs[2]'s shape)"},
        "kernel_classes": shares,
        "cpu_baseline": cpu_base,
        "clocks": clocks}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
