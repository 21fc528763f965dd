"""GPU, world_size 2: the three exchanges of SURVEY.md section 8(e) end to end on the device
kernels (DB-sharded retrieval, hypothesis-sharded RANSAC, row-sharded all-pairs Chamfer) equal the
single-process results.  Two tests, so that a green result says which transport ran:
test_sharded_paths_world2_nccl needs two GPUs (one rank per GPU over NCCL) and is SKIPPED with that reason on
a one-GPU box; test_sharded_paths_world2_one_gpu runs both ranks on cuda:0 and gathers through gloo (it
covers the sharded use of the device kernels, not the transport)."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _problem():
    rng = np.random.default_rng(11)
    lib = rng.standard_normal((652, 256)).astype(np.float32)
    lib[400] = lib[17]  # duplicate across shards: the tie goes to the lower global index
    qry = rng.standard_normal((7, 256)).astype(np.float32)
    src = rng.uniform(-1, 1, (6000, 3)).astype(np.float32)
    R = np.array([[0.0, -1, 0], [1, 0, 0], [0, 0, 1]])
    tgt = (src @ R.T + 0.3 + 0.01 * rng.standard_normal((6000, 3))).astype(np.float32)
    tgt[::3] = rng.uniform(-2, 2, (2000, 3))
    pcs = [rng.uniform(-1, 1, (300, 3)).astype(np.float32) for _ in range(9)]
    return lib, qry, src, tgt, pcs


def _worker(rank, ws, port, q, nccl):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank if nccl else 0)
    if nccl:
        dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=ws)
    from corsair_b200 import dist as cd
    lib, qry, src, tgt, pcs = _problem()
    gi, gd = cd.retrieval_topk_sharded(qry, lib, 10)
    T, st = cd.ransac_sharded(src, tgt, 0.2, seed=5, n_hypotheses=4096)
    tab = cd.chamfer_matrix_sharded(pcs)
    q.put((rank, gi, gd, T.cpu().numpy(), st.cpu().numpy(), tab, dist.get_backend()))
    dist.destroy_process_group()


def test_sharded_paths_world2_nccl(ctx):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs: one rank per GPU over NCCL (bench.py --gpus N runs and checks the same three "
                    "exchanges on N GPUs: its `sharded` object)")
    _run_world2(True)


def test_sharded_paths_world2_one_gpu(ctx):
    _run_world2(False)


def _run_world2(nccl):
    import torch.multiprocessing as mp
    from corsair_b200 import eval_pose, pc_dist, retrieval
    lib, qry, src, tgt, pcs = _problem()
    ref_i, ref_d = retrieval.topk(qry, lib, 10)
    ref = eval_pose.ransac_device(src, tgt, 0.2, seed=5, n_hypotheses=4096, adaptive=False)
    ref_T, ref_st = ref["T"].cpu().numpy(), ref["stats"].cpu().numpy()
    ref_tab = pc_dist.compute_dist([torch.from_numpy(p) for p in pcs])
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = 29600 + os.getpid() % 2000
    procs = [mpc.Process(target=_worker, args=(r, 2, port, q, nccl)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=300) for _ in procs), key=lambda r: r[0])
    for p in procs:
        p.join(60)
    for rank, gi, gd, T, st, tab, backend in res:
        assert backend == ("nccl" if nccl else "gloo")
        assert np.array_equal(gi, ref_i) and np.allclose(gd, ref_d, rtol=0, atol=0)
        assert np.array_equal(T, ref_T), "sharded RANSAC pose differs from the single-GPU run"
        assert st[0] == ref_st[0] and st[1] == ref_st[1] and st[3] == ref_st[3] and abs(st[2] - ref_st[2]) <= 1e-9 * ref_st[2]
        assert np.array_equal(tab, ref_tab)
