"""Generates tests/golden/*.npz by running the REFERENCE's own Python (imported from
/root/reference with stub modules for its absent GUI / Open3D dependencies, oracle/ref_import.py)
on small seeded inputs.  Run here (the reference does not exist on the GPU box):

    python tests/golden/make_golden.py

The one call the reference delegates to Open3D (registration_based_on_corr) has no reference
implementation in the tree: golden vectors for it come from the NumPy restatement and are
labelled `restated` (parity unpinned for that call, see oracle/ransac_np.py)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from corsair_b200 import synth  # noqa: E402
from oracle import ransac_np, ref_import  # noqa: E402


def unit(rng, n, d):
    f = rng.standard_normal((n, d)).astype(np.float32)
    return (f / np.linalg.norm(f, axis=1, keepdims=True)).astype(np.float32)


def main():
    fn = ref_import.load("find_nn")
    ep = ref_import.load("eval_pose")
    pp = ref_import.load("preprocess")
    rt = ref_import.load("retrieval")
    pd = ref_import.load("pc_dist")
    te = ref_import.load("_transform_estimation")
    sy = ref_import.load("symmetry")
    rng = np.random.default_rng(2024)

    # ---- find_nn.py
    out = {}
    for d in (16, 32):
        F0, F1 = unit(rng, 700, d), unit(rng, 900, d)
        i5, d5 = fn.find_knn_cpu(F0, F1, 5, return_distance=True)
        i1, d1 = fn.find_nn_cpu(F0, F1, return_distance=True)
        gi, gd = fn.find_knn_gpu(torch.from_numpy(F0), torch.from_numpy(F1), 5, nn_max_n=300, return_distance=True)
        ni, nd = fn.find_nn_gpu(torch.from_numpy(F0), torch.from_numpy(F1), nn_max_n=300, return_distance=True,
                                dist_type="L2")
        pdm = fn.pdist(torch.from_numpy(F0[:50]), torch.from_numpy(F1[:60]), "L2")
        out.update({f"F0_{d}": F0, f"F1_{d}": F1, f"knn_cpu_idx_{d}": i5, f"knn_cpu_dist_{d}": d5,
                    f"nn_cpu_idx_{d}": i1, f"nn_cpu_dist_{d}": d1, f"knn_gpu_idx_{d}": gi.numpy(),
                    f"knn_gpu_dist_{d}": gd.numpy(), f"nn_gpu_idx_{d}": ni.numpy(), f"nn_gpu_dist_{d}": nd.numpy(),
                    f"pdist_L2_{d}": pdm.numpy()})
    k0, k1 = ep.find_kcorr(out["F0_16"], out["F1_16"], k=5, subsample_size=-1)
    out["kcorr_idx0"], out["kcorr_idx1"] = k0, k1
    np.savez_compressed(os.path.join(HERE, "find_nn.npz"), **out)

    # ---- preprocess.py / pc_dist.py
    a = synth.normalize_cloud(synth.make_chair(np.random.default_rng(1)), 2000)
    b = synth.normalize_cloud(synth.make_chair(np.random.default_rng(2)), 2000)
    c = synth.normalize_cloud(synth.make_table(np.random.default_rng(3)), 2000)
    T = torch.from_numpy(synth.random_pose(np.random.default_rng(4)).astype(np.float32))
    ta = pp.apply_transform(a, T)
    pcs = [torch.from_numpy(x) for x in (a, b, c)]
    table = pd.compute_dist(pcs)
    np.savez_compressed(os.path.join(HERE, "chamfer.npz"), a=a, b=b, c=c, T=T.numpy(), apply_transform=ta,
                        kdtree_1dir=pp.chamfer_kdtree_1direction(ta, b),
                        gpu_1dir=pp.chamfer_gpu_1direction(pcs[0], pcs[1]),
                        gpu_2dir=pp.chamfer_gpu(pcs[0], pcs[1]).item(), table=table)

    # ---- retrieval.py
    lib, q, gt = synth.descriptor_db(5, L=300, Q=40)
    tab = np.abs(np.random.default_rng(6).standard_normal((300, 300)))
    tab = tab + tab.T
    np.fill_diagonal(tab, 0)
    best = np.random.default_rng(7).integers(0, 300, 40)
    st = rt.scan2cad_retrieval_eval(q, lib, best, tab, 30)
    rank = rt.get_rank(lib, top_n=10)
    ev = rt.retrieval_eval(lib, 0.1, tab)
    labels = np.random.default_rng(8).integers(0, 5, 300)
    mAP = rt.compute_mAP(lib, labels)
    np.savez_compressed(os.path.join(HERE, "retrieval.npz"), lib=lib, q=q, table=tab, best_match=best,
                        precision=st["precision"], top1_error=st["top1_error"],
                        top1_predict=np.array(st["top1_predict"]), gt=np.array(st["gt"]), get_rank=rank,
                        eval_mAP=ev["mAP"], eval_percision=ev["percision"], eval_top1=ev["top1_error"],
                        labels=labels, mAP_instance=mAP[0], mAP_class=np.array(mAP[1]))

    # ---- _transform_estimation.py
    p0 = torch.from_numpy(np.random.default_rng(9).uniform(-1, 1, (1500, 3)).astype(np.float32))
    x = torch.tensor([[0.04], [-0.02], [0.06], [0.02], [-0.03], [0.01]])
    Tt = te.get_trans(x)
    p1 = te.update_pcd(p0, Tt) + 0.002 * torch.from_numpy(np.random.default_rng(10).standard_normal((1500, 3)).astype(np.float32))
    w = torch.from_numpy(np.random.default_rng(11).uniform(0.5, 1, (1500, 1)).astype(np.float32))
    np.savez_compressed(os.path.join(HERE, "irls.npz"), p0=p0.numpy(), p1=p1.numpy(), w=w.numpy(),
                        get_trans=Tt.numpy(), T_unweighted=te.est_quad_linear_robust(p0, p1).numpy(),
                        T_weighted=te.est_quad_linear_robust(p0, p1, w).numpy())

    # ---- symmetry.py: masks under the global NumPy seed, per category
    sym_out = {}
    for cat, K in (("chair", 2), ("table", 4)):
        p = synth.make_pair(21, 16, cat, n_raw=4000)
        np.random.seed(31)
        masks = sy.symmetric_cut4(p["posF"], p["xyz1"], K)
        sym_out.update({f"{cat}_F": p["posF"], f"{cat}_xyz": p["xyz1"], f"{cat}_masks": np.stack(masks)})
    # eval_pose
    T0, T1 = synth.random_pose(np.random.default_rng(12)), synth.random_pose(np.random.default_rng(13))
    Te = torch.from_numpy(synth.random_pose(np.random.default_rng(14)).astype(np.float32))
    sym_out.update(T0=T0, T1=T1, T_est=Te.numpy(), eval_pose_sym1=np.array(ep.eval_pose(Te, T0, T1, 1)),
                   eval_pose_sym4=np.array(ep.eval_pose(Te, T0, T1, 4)))
    # sym_pose end to end with the restated RANSAC patched in (restated: parity unpinned)
    p = synth.make_pair(22, 16, "chair", n_raw=3000)

    def reg(src, tgt, max_corr=0.03, seed=0):
        ms = ransac_np.draw_minimal_sets(len(src), 100000, 10, seed)
        return ransac_np.ransac(src.astype(np.float64), tgt.astype(np.float64), max_corr, ms, adaptive=True)["T"]

    ep.registration_based_on_corr = reg
    sy.registration_based_on_corr = reg
    np.random.seed(31)
    Tb, cb, Tr, cr, ok = sy.sym_pose(p["baseF"], p["xyz0"], p["posF"], p["xyz1"], p["pos_sym"], k_nn=5,
                                     max_corr=0.2, seed=31)
    sym_out.update(sp_baseF=p["baseF"], sp_xyz0=p["xyz0"], sp_posF=p["posF"], sp_xyz1=p["xyz1"], sp_T_best=Tb.numpy(),
                   sp_ch_best=cb, sp_T_ransac=Tr.numpy(), sp_ch_ransac=cr, sp_ok=ok)
    np.savez_compressed(os.path.join(HERE, "symmetry.npz"), **sym_out)

    # ---- RANSAC restatement KAT (restated, not reference output)
    src = np.random.default_rng(15).uniform(-1, 1, (2000, 3)).astype(np.float32)
    Tg = synth.random_pose(np.random.default_rng(16))
    tgt = (src @ Tg[:3, :3].T + Tg[:3, 3] + 0.01 * np.random.default_rng(17).standard_normal((2000, 3))).astype(np.float32)
    tgt[::3] = np.random.default_rng(18).uniform(-2, 2, (len(tgt[::3]), 3)).astype(np.float32)
    ms = ransac_np.draw_minimal_sets(2000, 256, 10, seed=3)
    r = ransac_np.ransac(src, tgt, 0.2, ms, adaptive=False, return_all=True)
    np.savez_compressed(os.path.join(HERE, "ransac_restated.npz"), src=src, tgt=tgt, minimal_sets=ms, counts=r["counts"],
                        sums=r["sums"], T=r["T"], best_h=r["best_h"], inliers=r["inliers"])
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
