"""Where the time of one o3dpp_voxel call goes: host wall clock against the library's own event time."""
import sys, time
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
pts = lidar_like(n, 3)
c = nat.Context(0)
c.upload(nat.ATTR_POINTS, pts)
for rep in range(8):
    if rep >= 3:
        c.estimate_normals(nat.SEARCH_KNN, 30, 0.0)
        if rep >= 5:
            c.segment_plane(0.05, 3, 1000, seed=1, fetch_inliers=False)
        c.drop(nat.ATTR_NORMALS); c.synchronize()
    t0 = time.perf_counter(); V = c.voxel(0.1); t1 = time.perf_counter()
    print(rep, "V", V, "wall ms %.3f" % ((t1 - t0) * 1e3), "events ms %.3f" % c.timings()[0], flush=True)
