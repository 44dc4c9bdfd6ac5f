"""Every kernel of the library on small inputs, written for `compute-sanitizer --tool memcheck|racecheck python
tools/sanitize_run.py small`.  On this pool compute-sanitizer is CLOSED by the operators (gpurun answers
"compute-sanitizer is closed on this pool and stays closed", profiles/README.md r02c), so the script runs plain: the inputs
reach all three KNN tiers, the shrink / halving branches of tier 1, the wide-key voxel sort, the RANSAC early exit, the
slab-sharding pieces, selection, transform and the z-splat, and every result is compared with the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
from oracle import orc

small = len(sys.argv) > 1 and sys.argv[1] == "small"
c = nat.Context(0)
c.set_option(nat.OPT_KEEP_NEIGHBORS, 1)
bunny = np.load(os.path.join(ROOT, "tests", "golden", "bunny_xyz_f32.npy")).astype(np.float64)[:: (8 if small else 3)]
rng = np.random.default_rng(1)
adv = np.concatenate([rng.uniform(-20, 20, size=(3000, 3)), np.repeat(rng.uniform(-5, 5, size=(500, 3)), 4, 0),
                      rng.uniform(-50, 50, size=(8, 3))[rng.integers(0, 8, 3000)] + rng.normal(scale=1e-3, size=(3000, 3)),
                      rng.uniform(-1000, 1000, size=(20, 3))])
clouds = {"bunny": bunny, "lidar": lidar_like(6000 if small else 30000, 5), "adversarial": adv[rng.permutation(len(adv))], "tiny": rng.normal(size=(7, 3))}


def sets(idx, cnt, k):
    big = np.iinfo(np.int32).max
    e = np.where(np.arange(k)[None, :] < cnt[:, None], idx, big); e.sort(1); e[e == big] = -1
    return e


for name, pts in clouds.items():
    c.upload(nat.ATTR_POINTS, pts)
    T = np.eye(4); T[:3, 3] = [0.5, -1.0, 2.0]
    c.transform(T); assert np.array_equal(c.download(nat.ATTR_POINTS), orc.transform(pts, T)[0])
    P = T.copy(); P[3] = [0.001, 0.002, -0.001, 1.25]
    c.upload(nat.ATTR_POINTS, pts); c.transform(P); assert np.array_equal(c.download(nat.ATTR_POINTS), orc.transform(pts, P)[0])
    c.upload(nat.ATTR_POINTS, pts)
    for vs in (0.01 if name == "bunny" else 0.5, 1e-7):
        try:
            V = c.voxel(vs)
        except nat.O3dppError:
            continue
        ref = orc.voxel_trace(pts, vs)
        got = c.voxel_fetch(V, len(pts), ["rep", "count", "cubic_id", "centroid"])
        assert V == ref["V"] and all(np.array_equal(got[k], ref[k]) for k in got)
    V = c.voxel(0.37, pts.min(0) - 0.11, pts.max(0) + 1.0)
    assert V == orc.voxel_trace(pts, 0.37, min_bound=pts.min(0) - 0.11, max_bound=pts.max(0) + 1.0)["V"]
    for mode, radius, opts in ((0, 0.0, {}), (2, 0.02 if name == "bunny" else 0.6, {}), (0, 0.0, {nat.OPT_KNN_GSPLIT: 16, nat.OPT_KNN_MAXSHRINK: 1})):
        for o, v in opts.items(): c.set_option(o, v)
        c.drop(nat.ATTR_NORMALS)
        c.estimate_normals(mode, 30, radius)
        idx, cnt, tier = c.fetch_neighbors(30)
        eidx, _, ecnt = orc.knn(pts, 30, mode=mode, radius=radius)
        assert np.array_equal(cnt, ecnt) and np.array_equal(idx, sets(eidx, ecnt, 30)), (name, mode)
        c.set_option(nat.OPT_KNN_GSPLIT, 0); c.set_option(nat.OPT_KNN_MAXSHRINK, -1)
        print(name, "mode", mode, "tiers", np.bincount(tier, minlength=4)[1:].tolist(), flush=True)
    c.kth_bound()
    for k in (8, 40):
        idx, d2, cnt = c.knn(0, k)
        eidx, ed2, ecnt = orc.knn(pts, k)
        assert np.array_equal(idx, eidx) and np.array_equal(d2, ed2)
    c.drop(nat.ATTR_NORMALS); c.estimate_normals(1, 0, 0.02 if name == "bunny" else 0.6)
    c.knn_mean_distance(10)
    if len(pts) >= 3:
        thr = 0.01 if name == "bunny" else 0.05
        plane, inl, dbg = c.segment_plane(thr, 3, 100, seed=3, debug=True)
        ep, einl = orc.segment_plane(pts, thr, 3, 100, orc.sample_philox(3, len(pts), 3, 100))
        assert np.array_equal(inl, einl)
        s = c.sample_philox(3, len(pts), 3, 16)
        planes = c.planes_from_points(pts[s.reshape(-1)], 3)
        c.score_planes(planes, 0.05)
        m, s3 = c.plane_inliers(planes[0], 0.05)
        if m: c.inlier_moments(s3 / m); c.fetch_inlier_list(m)
        c.select_inliers(invert=True)
        c.upload(nat.ATTR_POINTS, pts); c.select(np.arange(0, len(pts), 3, dtype=np.int32))
        c.upload(nat.ATTR_POINTS, pts); V = c.voxel(0.5); c.voxel_select()
    c.upload(nat.ATTR_POINTS, pts)
    c.zsplat(64, float(pts[:, 0].min()), float(pts[:, 0].max()), float(pts[:, 1].min()), float(pts[:, 1].max()))
    c.drop(nat.ATTR_NORMALS); c.estimate_normals(0, 30, 0.0)
    out6 = np.empty((len(pts), 6)); c.download_rows6(out6)
    assert np.array_equal(out6[:, :3], pts)
    o32 = np.empty((len(pts), 3), np.float32); c.download_as(nat.ATTR_POINTS, o32)
    print(name, "ok", flush=True)
c.close()
print("sanitize_run: all kernels exercised, results equal to the oracle")
