"""Cuts the head of the reference-held result cache (/root/reference/data/cache_ret, written by the
reference's own evaluation.py:420-441 on the real Scan2CAD data) into small fixtures, and records the
dtype / rank / trailing shape of EVERY file of that cache:

    python tests/golden/make_cache_format_golden.py

tests/golden/cache_ret_head/<name>_chair_gt.npy   first 12 rows of the nine chair files
tests/golden/cache_format.json                    {file name: [dtype, ndim, trailing shape, rows]} + the
                                                  summary values evaluation.py:350-358 prints for them
The files are pipeline OUTPUTS (poses, losses), not source."""
import glob
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/data/cache_ret"


def main():
    fmt = {}
    for f in sorted(glob.glob(os.path.join(SRC, "*.npy"))):
        a = np.load(f)
        fmt[os.path.basename(f)] = [str(a.dtype), a.ndim, list(a.shape[1:]), int(a.shape[0])]
        if f.endswith("_chair_gt.npy"):
            np.save(os.path.join(HERE, "cache_ret_head", os.path.basename(f)), a[:12])
    full = {k: np.load(os.path.join(SRC, f"{k}_chair_gt.npy")) for k in
            ("t_losses_ransac", "t_losses_sym", "r_losses_ransac", "r_losses_sym", "chamfer_dist_ransac",
             "chamfer_dist_sym", "sym_ransac_success")}
    summary = {k: float(np.mean(v)) for k, v in full.items()}
    json.dump({"files": fmt, "chair_gt_means": summary}, open(os.path.join(HERE, "cache_format.json"), "w"), indent=1)
    print(json.dumps(summary, indent=1))


if __name__ == "__main__":
    main()
