#!/usr/bin/env python
"""Benchmark of the correspondence-and-registration hot path (BASELINE.json metric: query-CAD
pairs registered per second at 1/2/4/8 B200; kNN % of tensor peak).

Headline workload = BASELINE.json configs[2] ("C3", the full chair evaluation sweep of
evaluation.py:297-313): every rank registers its share of synthetic Scan2CAD-style chair queries
against their top-10 retrieved CADs.  One step = QUERIES_PER_STEP queries of this rank: their 256-d
descriptors are ranked against the 6778-CAD library in one batch, then every (query, candidate) pair
goes through the full symmetry-aided pipeline (sym_pose: feature 5-NN -> correspondence RANSAC
(budget 100000, Open3D's adaptive stop) -> Chamfer, two symmetry segmentations, K more {per-segment
kNN, RANSAC, Chamfer} pairings).  So a step registers QUERIES_PER_STEP x 10 = 100 pairs per rank and
the driver's 20 steps are a 2000-pair sweep per rank (weak scaling: the per-rank sweep is fixed).

  value  device-resident inputs (queries + CAD library already in HBM), timed with CUDA events
  e2e    the same steps through the reference-facing Python API with HOST NumPy buffers
         (host->device copies and the device->host result reads inside the timed region; the bytes
         are counted where the copies are made)
Sub-objects on the same line: `d16` (the same sweep at the feature width the shipped scripts use,
evaluation.py:183), `c5` (table category, 50k-point scans, 4096 hypotheses, nine registrations per
pair), `c4` (all-pairs Chamfer table 6778 x 2048 points, rows sharded over the ranks), `sharded`
(the three exchanges of SURVEY.md 8(e) run over the ranks and checked bit for bit against the
single-GPU result), `knn_tensor` (the metric's second half), `roofline`, `cpu_baseline`.

`python bench.py --impl reference` times the CPU path instead (the oracle's restatement of the
reference's algorithm on all host cores): per step ONE pair drawn uniformly (same seed) from the
step's 100 pairs, as a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# the CPU arm must see every host core: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
# silently single-thread the OpenMP restatement of the RANSAC call, sklearn and the BLAS underneath
if "reference" in sys.argv or int(os.environ.get("WORLD_SIZE", "1")) == 1:
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count())

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

L_DB, D_DESC, TOPK, CAD_POOL, QUERIES_PER_STEP = 6778, 256, 10, 48, 10
K_NN, MAX_CORR, SEED = 5, 0.2, 31
PAIRS_PER_STEP = QUERIES_PER_STEP * TOPK
WORKLOAD = ("C3: chair evaluation sweep — per step 10 synthetic Scan2CAD chair queries x top-10 of a 6778-CAD "
            "library, sym_pose registration of each pair (100 pairs/step/rank)")
L_TABLE, C4_L, C4_P = 8436, 6778, 2048
TABLE_SYM_HIST = {1: 233, 2: 422, 3: 7, 4: 128, 12: 40}  # configs/04379243_scan2cad_rot_sym_label.txt


# ------------------------------------------------------------------------------------ workloads
class Sweep:
    """Seeded synthetic inputs of the C3 sweep (corsair_b200/synth.py).  The CAD library holds
    CAD_POOL distinct procedural chairs (DB row i uses chair i % CAD_POOL) whose bulk is drawn so that
    the de-duplicated clouds span 2.6k-8.0k points; a query is a moved, thinned,
    jittered copy of its ground-truth CAD with a noisy copy of that row's descriptor."""

    def __init__(self, rank, d_feat, n_steps):
        from corsair_b200 import synth
        self.d = d_feat
        self.lib, _, _ = synth.descriptor_db(1000, L=L_DB, Q=1, d=D_DESC)
        brng = np.random.default_rng(4242)
        bulks = np.exp(brng.uniform(np.log(0.5), np.log(10.0), CAD_POOL))
        pairs = [synth.make_pair(100 + c, d_feat, "chair", bulk=float(bulks[c])) for c in range(CAD_POOL)]
        self.cads = [dict(F=p["posF"], xyz=p["xyz1"], sym=p["pos_sym"]) for p in pairs]
        rng = np.random.default_rng(7000 + rank)
        self.steps = []
        for _ in range(n_steps):
            qs = []
            for _ in range(QUERIES_PER_STEP):
                g = int(rng.integers(0, L_DB))
                desc = self.lib[g] + 0.3 * rng.standard_normal(D_DESC).astype(np.float32) / np.sqrt(D_DESC)
                desc = (desc / np.linalg.norm(desc)).astype(np.float32)
                src = pairs[g % CAD_POOL]
                keep = rng.random(len(src["xyz1"])) > 0.2
                T = synth.random_pose(rng)
                xyz0 = ((src["xyz1"][keep] + 0.005 * rng.standard_normal((int(keep.sum()), 3))) @ T[:3, :3].T
                        + T[:3, 3]).astype(np.float32)
                F0 = src["posF"][keep] + 0.05 * rng.standard_normal((int(keep.sum()), d_feat)).astype(np.float32)
                F0 = (F0 / np.linalg.norm(F0, axis=1, keepdims=True)).astype(np.float32)
                qs.append(dict(desc=desc, F=F0, xyz=xyz0, gt=g))
            self.steps.append(dict(desc=np.stack([q["desc"] for q in qs]), queries=qs))

    def point_counts(self):
        n0 = [len(q["xyz"]) for s in self.steps for q in s["queries"]]
        n1 = [len(c["xyz"]) for c in self.cads]
        return dict(query_min=int(min(n0)), query_max=int(max(n0)), cad_min=int(min(n1)), cad_max=int(max(n1)))

    def to(self, fn):
        """Copy of the sweep with every array mapped through fn (device upload / pinning)."""
        import copy
        o = copy.copy(self)
        o.lib = fn(self.lib)
        o.cads = [dict(F=fn(c["F"]), xyz=fn(c["xyz"]), sym=c["sym"]) for c in self.cads]
        o.steps = [dict(desc=fn(s["desc"]), queries=[dict(F=fn(q["F"]), xyz=fn(q["xyz"]), gt=q["gt"]) for q in s["queries"]])
                   for s in self.steps]
        return o


def candidate_pairs(order, step):
    """(query index, CAD pool index) of a step's pairs, in registration order."""
    return [(qi, int(c) % CAD_POOL) for qi in range(len(step["queries"])) for c in order[qi]]


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
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                pw.append(float(r[3]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                pass
        # "under load": the samples drawing more than half of the highest power seen
        load = [s for s, p in zip(sm, pw) if pw and p >= 0.5 * max(pw)] or sm
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(pw) if pw else None}


def base_config(args, sweep):
    """The workload description both arms print (same keys, same values)."""
    return {"workload": WORKLOAD, "db": L_DB, "descriptor_dim": D_DESC, "feature_dim": args.feature_dim, "k_nn": K_NN,
            "pairs_per_step_per_rank": PAIRS_PER_STEP,
            "ransac": "budget 100000, adaptive stop (confidence 0.999), ransac_n 10, max_corr 0.2",
            "points_per_cloud_after_0.03_dedup": sweep.point_counts(), "cad_pool": CAD_POOL}


# ------------------------------------------------------------------------------------ CPU arm
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


def cpu_sample(sweep, step_index, rng, n_pairs, deadline=None, seconds=None):
    """Retrieval of the whole step (cdist + stable argsort, utils/retrieval.py:170-177) and up to n_pairs
    of its 100 pairs drawn uniformly without replacement (at least one; it stops drawing once
    time.perf_counter() passes `deadline`).  Appends each pair's wall seconds to `seconds`."""
    from scipy.spatial.distance import cdist
    step = sweep.steps[step_index % len(sweep.steps)]
    order = np.argsort(cdist(step["desc"], sweep.lib), axis=1, kind="stable")[:, :TOPK]
    pairs = candidate_pairs(order, step)
    chosen = rng.choice(len(pairs), n_pairs, replace=False)
    ran = []
    for i in chosen:
        if ran and deadline is not None and time.perf_counter() > deadline:
            break
        qi, ci = pairs[int(i)]
        t = time.perf_counter()
        cpu_pair(step["queries"][qi], sweep.cads[ci])
        if seconds is not None:
            seconds.append(time.perf_counter() - t)
        ran.append(int(i) % TOPK)
    return ran


REF_MIN_PAIRS, REF_MIN_SECONDS = 48, 60.0


def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import ransac_c
    ransac_c.build()
    sweep = Sweep(0, args.feature_dim, args.warmup + args.steps)
    rng = np.random.default_rng(SEED)  # same draw sequence as the cpu_baseline leg of the GPU arm
    for i in range(args.warmup):
        cpu_sample(sweep, i, rng, 1)
    # the sample: pairs spread evenly over the K steps until 48 pairs or 60 s, whichever comes first
    per_step = -(-REF_MIN_PAIRS // max(1, args.steps))
    t0 = time.perf_counter()
    ranks_drawn, secs = [], []
    for i in range(args.steps):
        ranks_drawn += cpu_sample(sweep, args.warmup + i, rng, per_step, deadline=t0 + REF_MIN_SECONDS, seconds=secs)
    dt = time.perf_counter() - t0
    pairs = len(ranks_drawn)
    v = pairs / dt
    cores = os.cpu_count()
    sample = (f"{pairs} pairs over {args.steps} steps: after each step's retrieval up to {per_step} of its "
              f"{PAIRS_PER_STEP} (query, top-10 candidate) pairs drawn uniformly (seed {SEED}), until {REF_MIN_PAIRS} "
              f"pairs or {REF_MIN_SECONDS:.0f} s; candidate ranks drawn: {ranks_drawn}; OpenMP threads {ransac_c.num_threads()}")
    print(json.dumps({
        "impl": "reference", "metric": "query-CAD pairs registered/sec", "value": v, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": base_config(args, sweep),
        "run": {"sample": sample, "pairs_timed": pairs,
                "seconds_per_pair": {"min": min(secs), "median": float(np.median(secs)), "max": max(secs)}},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------ GPU arm
def gpu_step(retrieval, symmetry, sweep, step, workers):
    """One step through the public drop-in API; inputs are whatever the sweep holds (CUDA tensors for
    `value`, host NumPy for `e2e`).  The step's pairs are independent; sym_pose_many registers them
    concurrently (workers=1: one after the other, as instrumented)."""
    idx, _ = retrieval.topk(step["desc"], sweep.lib, TOPK)     # D2H of the candidate ids
    jobs = []
    for qi, ci in candidate_pairs(idx, step):
        q, cad = step["queries"][qi], sweep.cads[ci]
        jobs.append(((q["F"], q["xyz"], cad["F"], cad["xyz"], cad["sym"]),
                     dict(k_nn=K_NN, max_corr=MAX_CORR, seed=SEED, rng=SEED)))
    return symmetry.sym_pose_many(jobs, workers=workers)


def load_peaks():
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peaks = json.load(open(pk_path)) if os.path.exists(pk_path) else {}
    src = "measured" if peaks else "fallback"
    return dict(hbm=peaks.get("hbm_gbs", 6650.0), bf16_burst=peaks.get("bf16_tflops", 1590.0),
                bf16_sustained=peaks.get("bf16_tflops_sustained", 1400.0), src=src)


def run_ours(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from corsair_b200 import _capi, get_context, retrieval, symmetry
    ctx = get_context(local_rank)
    workers = int(os.environ.get("CB_BENCH_WORKERS", "0")) or symmetry.DEFAULT_WORKERS
    peaks = load_peaks()

    up = lambda a: torch.from_numpy(a).to(dev)                       # noqa: E731
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy()         # noqa: E731
    n_total = args.warmup + args.steps
    sweep = Sweep(rank, args.feature_dim, n_total)
    sweep_d, sweep_h = sweep.to(up), sweep.to(pin)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(sw, steps, warmup, nworkers):
        for i in range(warmup):
            gpu_step(retrieval, symmetry, sw, sw.steps[i % len(sw.steps)], nworkers)
        barrier()
        evs = []
        for i in range(steps):
            flush.fill_(i & 0xff)  # L2 flush between timed iterations (outside the event pair)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gpu_step(retrieval, symmetry, sw, sw.steps[(warmup + i) % len(sw.steps)], nworkers)
            e1.record()
            evs.append((e0, e1))
        barrier()
        return max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))

    def copy_bytes():
        a = b = 0
        for c in _capi.all_contexts():
            x, y = c.copy_bytes()
            a, b = a + x, b + y
        return a, b

    count_launches = lambda: sum(c.launches() for c in _capi.all_contexts())  # noqa: E731

    # (1) instrumented pass, one pair after the other on one stream: per-class CUDA-event timing of
    #     the library's kernels (meaningless when several streams overlap) -> roofline, shares
    inst_step = dict(desc=sweep_d.steps[0]["desc"][:2], queries=sweep_d.steps[0]["queries"][:2])  # 20 pairs
    gpu_step(retrieval, symmetry, sweep_d, inst_step, 1)
    ctx.timing_reset()
    ctx.timing(True)
    gpu_step(retrieval, symmetry, sweep_d, inst_step, 1)
    ctx.timing(False)
    classes = ctx.timing_read()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    gpu_step(retrieval, symmetry, sweep_d, inst_step, 1)
    torch.cuda.synchronize()
    ms_seq_pair = 1e3 * (time.perf_counter() - t0) / (2 * TOPK)

    # (2) the timed runs
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev = timed(sweep_d, args.steps, args.warmup, workers)
    clocks = sampler.stop()
    barrier()
    l0 = count_launches()
    gpu_step(retrieval, symmetry, sweep_d, sweep_d.steps[0], workers)
    barrier()
    launches_per_step = count_launches() - l0
    h0, d0 = copy_bytes()
    ms_e2e = timed(sweep_h, args.steps, args.warmup, workers)
    h1, d1 = copy_bytes()
    h2d_step, d2h_step = (h1 - h0) / n_total, (d1 - d0) / n_total
    pairs = world * args.steps * PAIRS_PER_STEP
    value, e2e = pairs / (ms_dev * 1e-3), pairs / (ms_e2e * 1e-3)

    # (3) the same sweep at the feature width the shipped scripts use (evaluation.py:183)
    d16 = None
    if args.feature_dim != 16 and not args.quick:
        n16 = max(2, min(6, args.steps))
        s16 = Sweep(rank, 16, n16 + 1).to(up)
        ms16 = timed(s16, n16, 1, workers)
        d16 = {"value": world * n16 * PAIRS_PER_STEP / (ms16 * 1e-3), "unit": "pairs/s", "feature_dim": 16, "steps": n16,
               "ms_per_step": ms16 / n16, "workload": "the C3 sweep with 16-d per-point features (evaluation.py:183), device-resident"}
        del s16

    # (4) strong scaling: ONE fixed sweep of 8 steps (800 pairs, the same on every run) split over the ranks
    #     (rank r registers steps r, r + N, ...); the headline above is weak scaling (fixed sweep per rank)
    strong = None
    if not args.quick:
        fixed = Sweep(424242, args.feature_dim, 8).to(up)
        mine = [fixed.steps[i] for i in range(rank, 8, world)]
        for st_ in mine[:1]:
            gpu_step(retrieval, symmetry, fixed, st_, workers)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st_ in mine:
            gpu_step(retrieval, symmetry, fixed, st_, workers)
        e1.record()
        barrier()
        ms_fix = max_over_ranks(e0.elapsed_time(e1))
        strong = {"value": 8 * PAIRS_PER_STEP / (ms_fix * 1e-3), "unit": "pairs/s", "ms": ms_fix, "pairs": 8 * PAIRS_PER_STEP,
                  "scaling": "strong", "workload": "one fixed C3 sweep of 800 pairs (8 steps, the same at every N) split over the "
                                                   "ranks step by step, device-resident, max over ranks"}
        del fixed

    extra = {} if args.quick else side_workloads(args, rank, world, dev, dist, ctx, peaks, workers, barrier, max_over_ranks)
    if rank != 0:
        return

    # ---- roofline of the kernel class with the largest device time in the sequential pass
    tracked = {k: v for k, v in classes.items() if v[2] > 0 and k != "ransac_waves"}
    top = max(("knn_candidates", "ransac_score"), key=lambda k: tracked.get(k, (0, 0, 0))[0])
    ms_k, work_k, n_k = tracked[top]
    traffic = {}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        traffic = json.load(open(tr_path))
    if top == "ransac_score":
        evals_s = work_k / (ms_k * 1e-3)
        waves = max(1.0, classes["ransac_waves"][1])
        kk = ctx.ransac_k()
        roof = {"bound": "tensor", "kernel": "ransac_tc_score_kernel (" + ctx.ransac_variant() + ")",
                "achieved": evals_s * 2.0 * kk / 1e12, "peak": peaks["bf16_burst"], "unit": "TFLOP/s",
                "note": f"issued bf16 flop: (hypothesis, correspondence) evaluations/s x 2 x K, K = {kk} columns of the split "
                        f"inner product; evaluations counted on the device for the CTAs that ran; peak = {peaks['src']} "
                        "bf16 dense burst (the kernel is timed launch by launch in the sequential pass)",
                "evals_per_s": evals_s, "effective_gbs": evals_s * 32.0 / 1e9,
                "effective_gbs_note": "BASELINE.json's H x M x 32 B per second; the operands are L2/shared-memory resident, "
                                      "so this is not an HBM fraction (dram traffic below)",
                "launches": int(waves), "avg_launch_us": 1e3 * ms_k / waves,
                "launches_skipped": int(n_k - waves)}
        roof["evals_per_clk_per_sm"] = evals_s / (148 * (clocks.get("sm_mhz") or 1965.0) * 1e6)
    else:
        roof = {"bound": "tensor", "kernel": "knn candidates (" + ctx.knn_variant() + ")",
                "achieved": work_k / (ms_k * 1e-3) / 1e12, "peak": peaks["bf16_sustained"] / 3.0, "unit": "TFLOP/s",
                "note": f"2*N0*N1*D flop; peak = {peaks['src']} sustained bf16 dense / 3 (split-bf16 issues 3 products)",
                "launches": n_k, "avg_launch_us": 1e3 * ms_k / max(1, n_k)}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["traffic"] = traffic.get(top)
    total_tracked = sum(v[0] for v in tracked.values())
    shares = {k: {"ms": round(v[0], 3), "launches": v[2], "share_of_tracked": round(v[0] / max(total_tracked, 1e-9), 4)}
              for k, v in tracked.items()}

    cpu_base = None
    if world == 1 and not args.quick:
        from oracle import ransac_c
        ransac_c.build()
        rng = np.random.default_rng(SEED)
        cpu_sample(sweep, 0, rng, 1)  # warm the KD-tree / BLAS / OpenMP pools
        t0 = time.perf_counter()
        n, drawn, secs = 0, [], []
        while n < 3 or (time.perf_counter() - t0 < 15 and n < 8):
            drawn += cpu_sample(sweep, 1 + n, rng, 1, seconds=secs)
            n += 1
        dt = time.perf_counter() - t0
        cpu_base = {"value": n / dt, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port",
                    "seconds_per_pair": {"min": min(secs), "median": float(np.median(secs)), "max": max(secs)},
                    "sample": f"{n} pairs of the same sweep, one drawn uniformly (seed {SEED}) from each of {n} steps' "
                              f"{PAIRS_PER_STEP} pairs after that step's retrieval; candidate ranks drawn {drawn}; "
                              f"OpenMP threads {ransac_c.num_threads()}"}

    print(json.dumps({
        "metric": "query-CAD pairs registered/sec", "value": value, "unit": "pairs/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 search / f64 re-rank, fit and in-band scoring",
        "data": "synthetic",
        "config": base_config(args, sweep),  # the same object in both arms; what is specific to this arm is in "run"
        "run": {"pairs_timed": pairs, "l2": "256 MiB flush between timed steps", "driver": symmetry.driver_name(),
                "concurrency": workers, "sequential_ms_per_pair": ms_seq_pair,
                "sharding": f"queries sharded over {world} rank(s), no data-path collective"},
        "e2e": {"value": e2e, "unit": "pairs/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d_step,
                "d2h_bytes_per_step": d2h_step, "bytes": "counted where the copies are made (Context.dev / Context.host / "
                                                           "the library's own staging), averaged over warm-up + timed steps"},
        "gpu_launches": launches_per_step * args.steps, "launches_per_pair": launches_per_step / PAIRS_PER_STEP,
        "roofline": roof, "kernel_classes": shares, "d16": d16, "strong": strong, **extra,
        "cpu_baseline": cpu_base, "clocks": clocks}))


def side_workloads(args, rank, world, dev, dist, ctx, peaks, workers, barrier, max_over_ranks):
    """C5, C4, the three sharded exchanges and the kNN tensor figure (see the module docstring)."""
    import torch
    from corsair_b200 import dist as cdist_mod, eval_pose, find_nn, pc_dist, retrieval, symmetry, synth
    out = {}
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    # ---- kNN distance GEMM against the split-precision tensor peak, on the largest single-pair shape
    #      BASELINE.json names (configs[4]: 50k-point scan vs a 10k-point CAD); result checked below
    g = torch.Generator(device=dev).manual_seed(5)
    knn_t = {}
    for d in (32, 16):
        fa = torch.nn.functional.normalize(torch.randn(50000, d, device=dev, generator=g), dim=1)
        fb = torch.nn.functional.normalize(torch.randn(10000, d, device=dev, generator=g), dim=1)
        for _ in range(3):
            idx = find_nn.knn_device(fa, fb, K_NN)
        e0, e1 = ev(), ev()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            idx = find_nn.knn_device(fa, fb, K_NN)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 10 * 1e3
        fb_rows = ctx.knn_last_fallback_rows()
        # the timed call's answer against exact fp64 brute force on a sample of rows (torch, on the device)
        rows = torch.arange(0, 50000, 97, device=dev)
        d2 = torch.cdist(fa[rows].double(), fb.double()) ** 2
        ok = bool((torch.gather(d2, 1, idx[rows]) <= torch.topk(d2, K_NN, largest=False).values * (1 + 1e-9) + 1e-12).all())
        tf = 2.0 * 50000 * 10000 * d / (us * 1e-6) / 1e12
        knn_t[f"d{d}"] = {"shape": f"50000 x 10000, D = {d}, k = {K_NN}", "achieved": tf, "unit": "TFLOP/s",
                          "peak": peaks["bf16_burst"] / 3.0, "frac": tf / (peaks["bf16_burst"] / 3.0), "us_per_call": us,
                          "fallback_rows": fb_rows, "checked_rows_ok": ok}
        del fa, fb, d2
    out["knn_tensor"] = dict(knn_t["d32"], d16=knn_t["d16"],
                             note="whole cb_knn call (packing, tcgen05 candidate kernel, exact re-rank), 2*N0*N1*D flop; "
                                  f"peak = {peaks['src']} BURST bf16 dense / 3 products (ten calls timed alone, ~2 ms)")

    # ---- C5: table category, pos_sym from the real histogram, 50k-point scans, 4096 hypotheses
    labels = np.repeat(list(TABLE_SYM_HIST), list(TABLE_SYM_HIST.values()))
    lrng = np.random.default_rng(900 + rank)
    n_c5 = 6
    c5 = []
    for i in range(n_c5):
        p = synth.make_pair(500 + 10 * rank + i, args.feature_dim, "table", n_raw=50000, n_cad=10000)
        sym = int(lrng.choice(labels))
        c5.append(((torch.from_numpy(p["baseF"]).to(dev), torch.from_numpy(p["xyz0"]).to(dev),
                    torch.from_numpy(p["posF"]).to(dev), torch.from_numpy(p["xyz1"]).to(dev), sym),
                   dict(k_nn=K_NN, max_corr=MAX_CORR, seed=SEED, rng=SEED, n_hypotheses=4096)))
    symmetry.sym_pose_many(c5, workers=workers)
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    res = symmetry.sym_pose_many(c5, workers=workers)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    symmetry.sym_pose_many(c5[:2], workers=1)
    torch.cuda.synchronize()
    ms_seq = 1e3 * (time.perf_counter() - t0) / 2
    # per-class device time of the same two pairs (one stream, CUDA events inside the library) -> roofline of
    # the scoring kernel at this shape (M = 250 000, H = 4 096 per RANSAC) and the classes' shares
    ctx.timing_reset()
    ctx.timing(True)
    symmetry.sym_pose_many(c5[:2], workers=1)
    ctx.timing(False)
    cls = ctx.timing_read()
    sc_ms, sc_evals, _ = cls["ransac_score"]
    tot_ms = sum(v[0] for k, v in cls.items() if k != "ransac_waves")
    c5_roof = {"bound": "tensor", "kernel": "ransac_tc_score_kernel", "unit": "TFLOP/s", "peak": peaks["bf16_burst"],
               "achieved": sc_evals / max(sc_ms, 1e-9) * 1e3 * 2.0 * ctx.ransac_k() / 1e12,
               "evals_per_s": sc_evals / max(sc_ms, 1e-9) * 1e3,
               "kernel_class_share_of_tracked": {k: round(v[0] / max(tot_ms, 1e-9), 4) for k, v in cls.items()
                                                 if k != "ransac_waves" and v[2] > 0},
               "note": "issued bf16 flop of the scoring launches of two pairs registered one after the other; peak = "
                       f"{peaks['src']} bf16 dense burst"}
    c5_roof["frac"] = c5_roof["achieved"] / c5_roof["peak"]
    regs = [1 + (8 if j[0][4] >= 2 else 2) for j in c5]
    out["c5"] = {"value": world * n_c5 / (ms * 1e-3), "unit": "pairs/s", "pairs_per_rank": n_c5, "ms": ms,
                 "sequential_ms_per_pair": ms_seq, "registrations_per_pair": regs,
                 "symmetry_ok": [bool(r[4]) for r in res], "roofline": c5_roof,
                 "workload": "C5: table, 50 000-point scan vs 10 000-point CAD (M = 250 000), 4096 hypotheses per RANSAC, pos_sym "
                             "drawn from the real label histogram (K = 4 and nine registrations when >= 2), device-resident"}

    # ---- e3: one C5-sized RANSAC (M = 250 000, H = 4096) with the hypothesis range split over the ranks
    p3 = synth.make_pair(500, args.feature_dim, "table", n_raw=50000, n_cad=10000)  # the SAME pair on every rank
    a = [torch.from_numpy(p3[k]).to(dev) for k in ("baseF", "xyz0", "posF", "xyz1")]
    idx = find_nn.knn_device(a[0], a[2], K_NN)
    si = torch.repeat_interleave(torch.arange(a[0].shape[0], device=dev), K_NN)
    ti = idx.reshape(-1)
    single = eval_pose.ransac_device(a[1], a[3], MAX_CORR, seed=SEED, src_idx=si, tgt_idx=ti, n_hypotheses=4096, adaptive=False)
    for _ in range(2):
        Tm, stm = cdist_mod.ransac_sharded(a[1], a[3], MAX_CORR, seed=SEED, n_hypotheses=4096, src_idx=si, tgt_idx=ti)
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    Tm, stm = cdist_mod.ransac_sharded(a[1], a[3], MAX_CORR, seed=SEED, n_hypotheses=4096, src_idx=si, tgt_idx=ti)
    e1.record()
    barrier()
    same = bool(torch.equal(Tm, single["T"]) and torch.equal(stm[:3], single["stats"][:3]))
    sharded = {"ransac_hypothesis_sharded": {"ms": max_over_ranks(e0.elapsed_time(e1)), "m": int(len(si)), "hypotheses": 4096,
                                             "equal_to_single_gpu": same, "ranks": world}}

    # ---- e2: retrieval of 1000 queries with the 6778 library rows split over the ranks
    lib, q, _ = synth.descriptor_db(1000, L=L_DB, Q=1000, d=D_DESC)
    lib_d, q_d = torch.from_numpy(lib).to(dev), torch.from_numpy(q).to(dev)
    ref_i, ref_d = retrieval.topk(q_d, lib_d, TOPK)
    cdist_mod.retrieval_topk_sharded(q_d, lib_d, TOPK)
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    gi, gd = cdist_mod.retrieval_topk_sharded(q_d, lib_d, TOPK)
    e1.record()
    barrier()
    sharded["retrieval_db_sharded"] = {"ms": max_over_ranks(e0.elapsed_time(e1)), "queries": 1000, "library": L_DB, "k": TOPK,
                                       "equal_to_single_gpu": bool(np.array_equal(gi, ref_i) and np.array_equal(gd, ref_d)),
                                       "ranks": world}

    # ---- C4 / e4: all-pairs Chamfer table, 6778 clouds x 2048 points, rows sharded over the ranks
    crng = np.random.default_rng(77)
    base = [synth.normalize_cloud(synth.make_chair(np.random.default_rng(300 + c), 4096), C4_P) for c in range(64)]
    stack = np.stack([base[i % 64] + 0.02 * crng.standard_normal((C4_P, 3)).astype(np.float32) for i in range(C4_L)])
    stack_d = torch.from_numpy(stack.astype(np.float32)).to(dev)
    barrier()
    t0 = time.perf_counter()
    c4_phases = {}
    table = cdist_mod.chamfer_matrix_sharded(stack_d, timings=c4_phases)
    barrier()
    sec = max_over_ranks(time.perf_counter() - t0)
    rows = [0, 1, C4_L // 2, C4_L - 1]
    ref_rows = np.stack([pc_dist.chamfer_rows_device(stack_d, r, r + 1).cpu().numpy()[0] for r in rows])
    got = table[rows]
    for k_, r in enumerate(rows):
        ref_rows[k_, r] = got[k_, r]  # the diagonal convention (200) is the table's
    n_pairs = C4_L * (C4_L - 1) // 2
    evals = n_pairs * float(C4_P) * C4_P  # one distance tile feeds both directions
    fp32_peak = 67.1  # measured: scripts/microbench/fp64_rate.cu, profiles/r02_fp64_rate.txt (nominal 74.4)
    out["c4"] = {"value": n_pairs / sec, "unit": "cloud pairs/s", "seconds": sec, "clouds": C4_L, "points": C4_P, "ranks": world,
                 "rows_equal_to_single_gpu": bool(np.array_equal(got, ref_rows)),
                 "symmetric": bool(np.array_equal(table, table.T)), "diagonal": float(table[0, 0]),
                 "roofline": {"bound": "fp32", "achieved": evals * 8 / sec / 1e12 / world, "unit": "TFLOP/s per GPU",
                              "peak": fp32_peak, "frac": evals * 8 / sec / 1e12 / world / fp32_peak,
                              "note": "8 flop per point-pair evaluation (SURVEY 8d), wall time of the whole sharded call incl. the "
                                      "all-gather, H + H^T on the device and the device->host read of the table (split in "
                                      "`phases_s`, this rank's wall seconds); peak = the fp32 FMA rate this repo measured on a "
                                      "B200 (profiles/r02_fp64_rate.txt: 67.1 TFLOP/s; nominal 74.4)"},
                 "phases_s": {k_: round(v_, 4) for k_, v_ in c4_phases.items()},
                 "workload": "C4: all-pairs Chamfer table (utils/pc_dist.py:61-99), row blocks over the ranks, one all-gather"}
    sharded["chamfer_rows_sharded"] = {"seconds": sec, "equal_to_single_gpu": out["c4"]["rows_equal_to_single_gpu"], "ranks": world}
    out["sharded"] = sharded
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--feature-dim", type=int, default=32, choices=[16, 32])
    ap.add_argument("--quick", action="store_true", help="headline + roofline only (profiling runs)")
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
