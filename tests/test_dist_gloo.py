"""CPU, world_size 2 over gloo: the host-side sharding / merge logic of corsair_b200/dist.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, ws, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    from corsair_b200 import dist as cdist_
    from scipy.spatial.distance import cdist
    rng = np.random.default_rng(5)
    lib = rng.standard_normal((101, 16)).astype(np.float32)
    lib[40] = lib[7]  # a duplicate row across shards: tie must go to the lower global index
    qry = rng.standard_normal((9, 16)).astype(np.float32)
    lo, hi = cdist_.shard_range(len(lib))
    d = cdist(qry, lib[lo:hi])
    order = np.argsort(d, axis=1, kind="stable")[:, :10]
    gi, gd = cdist_.merge_topk(torch.from_numpy(order), torch.from_numpy(np.take_along_axis(d, order, 1)), lo, 10)
    full = cdist(qry, lib)
    ref = np.argsort(full, axis=1, kind="stable")[:, :10]
    ok_topk = np.array_equal(gi.numpy(), ref) and np.allclose(gd.numpy(), np.take_along_axis(full, ref, 1))
    # hypothesis-sharded RANSAC merge: rank 1 holds the better record, equal counts break on rmse
    T = torch.eye(4, dtype=torch.float64) * (rank + 1)
    stats = torch.tensor([3.0, 50.0, 2.0 - rank, 64.0], dtype=torch.float64)
    Tm, st = cdist_.merge_best_hypothesis(T, stats, 64 * rank)
    ok_best = Tm[0, 0].item() == 2.0 and st.tolist() == [67.0, 50.0, 1.0, 128.0]
    rows = cdist_.gather_rows(torch.full((2, 3), float(rank)))
    ok_rows = rows.shape == (4, 3) and rows[:, 0].tolist() == [0.0, 0.0, 1.0, 1.0]
    items = cdist_.cyclic_items(7)
    q.put((rank, ok_topk, ok_best, ok_rows, items, (lo, hi)))
    dist.destroy_process_group()


def test_sharded_merges_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
    assert all(r[1] and r[2] and r[3] for r in res), res
    assert res[0][4] == [0, 2, 4, 6] and res[1][4] == [1, 3, 5]
    assert res[0][5] == (0, 51) and res[1][5] == (51, 101)


def test_shard_range_covers_everything():
    from corsair_b200 import dist as cdist_
    for n in (0, 1, 7, 8, 6778):
        for ws in (1, 2, 3, 8):
            spans = [cdist_.shard_range(n, r, ws) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
