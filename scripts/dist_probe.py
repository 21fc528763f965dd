#!/usr/bin/env python
"""torchrun probe: hypothesis-sharded RANSAC at C5 size against the single-GPU result (prints both)."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if int(os.environ.get("WORLD_SIZE", 1)) > 1:
    dist.init_process_group("nccl", device_id=dev)
from corsair_b200 import dist as cd, eval_pose, find_nn, synth
p = synth.make_pair(500, 32, "table", n_raw=50000, n_cad=10000)
a = [torch.from_numpy(p[k]).to(dev) for k in ("baseF", "xyz0", "posF", "xyz1")]
idx = find_nn.knn_device(a[0], a[2], 5)
si = torch.repeat_interleave(torch.arange(a[0].shape[0], device=dev), 5)
ti = idx.reshape(-1)
single = eval_pose.ransac_device(a[1], a[3], 0.2, seed=31, src_idx=si, tgt_idx=ti, n_hypotheses=4096, adaptive=False, return_all=True)
Tm, stm = cd.ransac_sharded(a[1], a[3], 0.2, seed=31, n_hypotheses=4096, src_idx=si, tgt_idx=ti)
c = single["counts"].cpu().numpy()
top = np.argsort(-c)[:5]
print(rank, "single stats", single["stats"].cpu().numpy(), "sharded", stm.cpu().numpy(), "Tdiff", float((Tm - single["T"]).abs().max()),
      "top counts", [(int(h), int(c[h]), float(single["sums"][h])) for h in top], flush=True)
if dist.is_initialized():
    dist.destroy_process_group()
