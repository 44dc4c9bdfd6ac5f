"""Mints tests/golden/bunny_goldens.npz from the CPU oracle (oracle/o3d_oracle.c).

    python tests/golden/make_goldens.py

The reference ships no golden vectors for this path and its arithmetic lives in the
un-vendored open3d==0.19.0 wheel, which cannot be imported here, so these goldens pin the ORACLE,
not Open3D ("parity unpinned").  The input is the reference's own fixture data/bunny.npy
(committed as bunny_xyz_f32.npy).  Values that independent code can check (d2 multisets against
scipy cKDTree, eigenvectors against numpy eigh, inlier counts against numpy) are checked in
tests/test_oracle.py; this file freezes them so that a later change of the oracle is noticed.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

T_FIXED = np.array([[0.36, 0.48, -0.8, 0.1], [-0.8, 0.6, 0.0, -0.2], [0.48, 0.64, 0.6, 0.3], [0.0, 0.0, 0.0, 1.0]])
KNN_ROWS = np.arange(0, 35947, 97)          # 371 query rows kept
SEG = dict(thr=0.01, rn=3, H=450, seed=5)


def main():
    b = np.load(os.path.join(HERE, "bunny_xyz_f32.npy")).astype(np.float64)
    out = {}
    for vs in (0.002, 0.005, 0.01):
        r = orc.voxel_trace(b, vs)
        tag = str(vs).replace(".", "p")
        out[f"vox_{tag}_V"] = np.int64(r["V"])
        out[f"vox_{tag}_rep_sorted"] = np.sort(r["rep"]).astype(np.int32)
        out[f"vox_{tag}_count_sum_sq"] = np.int64((r["count"].astype(np.int64) ** 2).sum())
        out[f"vox_{tag}_centroid_sum"] = r["centroid"].sum(0)
    idx, d2, cnt = orc.knn(b, 30)
    out["knn_rows"] = KNN_ROWS.astype(np.int32)
    out["knn_idx"] = idx[KNN_ROWS]
    out["knn_d2"] = d2[KNN_ROWS]
    # rows whose 30th and 31st neighbour tie exactly (nanoflann's pick there is traversal dependent)
    idx31, d31, _ = orc.knn(b, 31)
    out["knn_boundary_tie_count"] = np.int64((d31[:, 29] == d31[:, 30]).sum())
    nrm = orc.estimate_normals(b, "knn", 30)
    out["normals_rows"] = nrm[KNN_ROWS]
    nh = orc.estimate_normals(b, "hybrid", 30, 0.01)
    out["normals_hybrid_rows"] = nh[KNN_ROWS]
    samples = orc.sample_philox(SEG["seed"], len(b), SEG["rn"], SEG["H"])
    plane, inl, dbg = orc.segment_plane(b, SEG["thr"], SEG["rn"], SEG["H"], samples, debug=True)
    out["seg_samples_head"] = samples[:16]
    out["seg_plane"] = plane
    out["seg_best_plane"] = dbg["best_plane"]
    out["seg_best_iter"] = np.int64(dbg["best_iter"])
    out["seg_counts"] = dbg["counts"]
    out["seg_inliers"] = inl.astype(np.int32)
    out["mt19937_samples_head"] = orc.sample_mt19937(42, len(b), 3, 8)
    tp, _ = orc.transform(b[:64], T_FIXED)
    out["transform_T"] = T_FIXED
    out["transform_head"] = tp
    np.savez_compressed(os.path.join(HERE, "bunny_goldens.npz"), **out)
    print("wrote", os.path.join(HERE, "bunny_goldens.npz"), {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
