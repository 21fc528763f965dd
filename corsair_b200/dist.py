"""Sharding helpers for one 8xB200 box (one process per GPU, torch.distributed).

The path shards in three ways; only the last two exchange data, with one
all-gather of small per-shard records:
  * query-CAD pairs / queries: independent units, block-cyclic over ranks, no collective;
  * retrieval: the CAD library rows are split over ranks, per-shard top-k merged by
    (distance, global index);
  * RANSAC hypothesis sets: ranges of the seeded hypothesis stream per rank, best records merged
    with Open3D's rule (count desc, rmse asc, lowest hypothesis index first).
Works with the `nccl` backend on CUDA tensors and with `gloo` on CPU tensors (CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank=None, world_size=None):
    """Contiguous [lo, hi) share of n items for `rank` (sizes differ by at most one)."""
    if rank is None:
        rank, world_size = world()
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def cyclic_items(n, rank=None, world_size=None):
    """Block-cyclic item ids for independent units (pairs, queries)."""
    if rank is None:
        rank, world_size = world()
    return list(range(rank, n, world_size))


def _all_gather(t):
    rank, ws = world()
    if ws == 1:
        return [t]
    if t.is_cuda and dist.get_backend() == "gloo":
        # ranks sharing one GPU in tests: gloo gathers through the host
        host = t.detach().cpu().contiguous()
        out = [torch.empty_like(host) for _ in range(ws)]
        dist.all_gather(out, host)
        return [o.to(t.device) for o in out]
    out = [torch.empty_like(t) for _ in range(ws)]
    dist.all_gather(out, t.contiguous())
    return out


def merge_topk(idx_local, dist_local, shard_lo, k):
    """Merge per-shard top-k lists.  idx_local [Q,k'] (indices local to the shard), dist_local
    [Q,k'] float64 ascending; shard_lo = first global row of this rank's shard.  Every rank gets
    the global (idx [Q,k] int64, dist [Q,k]) ordered by (distance, global index)."""
    gi = idx_local.to(torch.int64) + int(shard_lo)
    all_i = torch.cat(_all_gather(gi), dim=1)
    all_d = torch.cat(_all_gather(dist_local.to(torch.float64)), dim=1)
    # lexicographic (distance, index): stable sort by index first, then stable sort by distance
    o1 = torch.argsort(all_i, dim=1, stable=True)
    all_i, all_d = torch.gather(all_i, 1, o1), torch.gather(all_d, 1, o1)
    o2 = torch.argsort(all_d, dim=1, stable=True)[:, :k]
    return torch.gather(all_i, 1, o2), torch.gather(all_d, 1, o2)


def merge_best_hypothesis(T, stats, h_lo):
    """Merge per-shard RANSAC results.  T [4,4] f64, stats [4] f64 = (best_h local to the shard,
    inliers, sum_d2, n_evaluated); h_lo = first hypothesis of this rank's range.  Returns the
    winning (T, stats with the GLOBAL hypothesis index and the summed n_evaluated).  The records stay
    on the tensors' device: one all-gather of 21 doubles per rank, then Open3D's order (count desc,
    rmse asc) with the lowest global hypothesis index on exact ties as three masked reductions."""
    rec = torch.cat([T.reshape(-1).to(torch.float64), stats.to(torch.float64),
                     torch.tensor([float(h_lo)], dtype=torch.float64, device=T.device)])
    recs = torch.stack(_all_gather(rec))                       # [world, 21]
    h, cnt, s2, nev, lo = recs[:, 16], recs[:, 17], recs[:, 18], recs[:, 19], recs[:, 20]
    total_eval = nev.sum()
    valid = h >= 0
    inf = torch.full_like(cnt, float("inf"))
    cnt_v = torch.where(valid, cnt, -inf)
    rmse = torch.where(cnt > 0, torch.sqrt(s2 / cnt.clamp(min=1.0)), torch.zeros_like(cnt))
    top = valid & (cnt_v == cnt_v.max())
    top = top & (rmse == torch.where(top, rmse, inf).min())
    gh = h + lo
    win = torch.argmin(torch.where(top, gh, inf))
    none = ~valid.any()
    best = recs[win]
    Tm = torch.where(none, torch.eye(4, dtype=torch.float64, device=T.device), best[:16].reshape(4, 4))
    st = torch.stack([torch.where(none, torch.full_like(total_eval, -1.0), best[16] + best[20]),
                      torch.where(none, torch.zeros_like(total_eval), best[17]),
                      torch.where(none, torch.zeros_like(total_eval), best[18]), total_eval])
    return Tm, st


def gather_rows(rows):
    """All-gather equally shaped per-rank result rows ([n_local, C]) -> [world*n_local, C]."""
    return torch.cat(_all_gather(rows), dim=0)


# ---- the three exchanges of SURVEY.md section 8(e), end to end on the device kernels -------------
def retrieval_topk_sharded(queries, library, k, metric="euclidean"):
    """DB-sharded retrieval: `library` is the FULL [L, d] array (every rank slices its own
    shard_range rows; pass a pre-sliced shard with `library=(shard, lo)`).  Per-shard top-k on the
    device (cb_retrieval_topk), one all-gather of [Q, k] records, merge by (distance, global index).
    Returns (idx [Q,k] int64, dist [Q,k] float64) NumPy, identical on every rank."""
    from . import retrieval
    if isinstance(library, tuple):
        shard, lo = library
    else:
        lo, hi = shard_range(len(library))
        shard = library[lo:hi]
    kk = min(k, len(shard))
    idx, d = retrieval.topk(queries, shard, kk, metric=metric)
    dev = torch.device("cuda", torch.cuda.current_device())
    pad = k - kk  # a shard smaller than k contributes +inf fillers
    ti = torch.from_numpy(np.pad(idx, ((0, 0), (0, pad)), constant_values=0)).to(dev)
    td = torch.from_numpy(np.pad(d, ((0, 0), (0, pad)), constant_values=np.inf)).to(dev)
    gi, gd = merge_topk(ti, td, lo, k)
    return gi.cpu().numpy(), gd.cpu().numpy()


def ransac_sharded(src_xyz, tgt_xyz, max_corr_dist, seed=0, n_hypotheses=4096, *, src_idx=None, tgt_idx=None):
    """Hypothesis-sharded RANSAC with a fixed budget (BASELINE.json configs[4]): rank r scores
    hypotheses [lo_r, hi_r) of the SAME seeded minimal-set stream (cb_draw_minimal_sets), the best
    records are merged with Open3D's rule after one all-gather.  Returns (T [4,4] f64, stats [4] f64
    = global best_h, inliers, sum_d2, n_evaluated) CUDA tensors, identical on every rank and equal to
    the single-GPU result with adaptive=False."""
    from . import _capi, eval_pose
    ctx = _capi.get_context()
    m = len(src_idx) if src_idx is not None else (len(tgt_idx) if tgt_idx is not None else len(src_xyz))
    lo, hi = shard_range(int(n_hypotheses))
    sets = ctx.empty((int(n_hypotheses), eval_pose.RANSAC_N), torch.int32)
    ctx.call("cb_draw_minimal_sets", int(seed) & 0xFFFFFFFFFFFFFFFF, int(m), int(n_hypotheses), eval_pose.RANSAC_N,
             _capi.ptr(sets), ctx.stream)
    if hi > lo:
        res = eval_pose.ransac_device(src_xyz, tgt_xyz, max_corr_dist, seed, src_idx=src_idx, tgt_idx=tgt_idx,
                                      minimal_sets=sets[lo:hi].contiguous(), adaptive=False, ctx=ctx)
        T, stats = res["T"], res["stats"]
    else:
        T = torch.eye(4, dtype=torch.float64, device=ctx.device)
        stats = torch.tensor([-1.0, 0.0, 0.0, 0.0], dtype=torch.float64, device=ctx.device)
    return merge_best_hypothesis(T, stats, lo)


def chamfer_matrix_sharded(pcs, timings=None):
    """All-pairs Chamfer table (utils/pc_dist.py:61-99) with the rows split over ranks.  Every rank
    computes the circulant half of its rows on the device (cb_chamfer_matrix_half: each unordered pair
    is computed exactly once across the ranks, the same number per row, so contiguous row blocks are
    balanced), one all-gather of the blocks (padded to equal height), then H + H^T with the
    reference's diagonal of 200 ON THE DEVICE and one device->host read of the finished table.
    Returns the [L,L] float64 table.  `timings` (a dict, optional) receives the wall seconds of the
    three phases: "kernel", "all_gather", "symmetrise_d2h"."""
    import time
    from . import pc_dist
    L = len(pcs)
    stack = torch.stack([p if isinstance(p, torch.Tensor) else torch.from_numpy(np.asarray(p)) for p in pcs]).float()
    rank, ws = world()
    lo, hi = shard_range(L)
    rows = (L + ws - 1) // ws
    dev = torch.device("cuda", torch.cuda.current_device())
    t0 = time.perf_counter()
    block = torch.zeros((rows, L), dtype=torch.float64, device=dev)
    if hi > lo:
        block[:hi - lo] = pc_dist.chamfer_rows_device(stack, lo, hi, half=True)
    torch.cuda.synchronize(dev)
    t1 = time.perf_counter()
    allb = gather_rows(block)
    torch.cuda.synchronize(dev)
    t2 = time.perf_counter()
    del block
    if ws > 1:
        H = torch.cat([allb[r * rows:r * rows + (b - a)] for r in range(ws) for a, b in [shard_range(L, r, ws)]], dim=0)
        del allb
    else:
        H = allb[:L]
    H.add_(H.t().clone())  # every unordered pair sits in exactly one of H[i, j], H[j, i]; x + 0 is exact
    H.fill_diagonal_(200.0)
    out = H.cpu().numpy()
    t3 = time.perf_counter()
    if timings is not None:
        timings.update(kernel=t1 - t0, all_gather=t2 - t1, symmetrise_d2h=t3 - t2)
    return out
