"""Minimal driver: voxel_down_sample(0.1) (points only) and the grid build of estimate_normals on an N-point LiDAR-like scan."""
import sys
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
pts = lidar_like(N, 1235)
c = nat.Context(0)
c.upload(nat.ATTR_POINTS, pts)
for _ in range(4):
    c.drop(nat.ATTR_NORMALS)
    V = c.voxel(0.1)
    tv = c.timings()[0]
    c.estimate_normals(0, 30, 0.0)
    print('voxel ms %.3f V %d | grid ms %.3f search %.3f' % (tv, V, c.timings()[0], c.timings()[1]))
import numpy as np
T = np.eye(4); T[:3, :3] = np.linalg.qr(np.random.default_rng(3).normal(size=(3, 3)))[0]; T[:3, 3] = (0.5, -1.0, 0.25)
c.drop(nat.ATTR_NORMALS)
ts = []
for k in range(12):
    c.transform(T if k % 2 == 0 else np.linalg.inv(T))
    ts.append(c.timings()[0])
print('transform ms', ' '.join('%.4f' % t for t in ts[4:]), '-> %.1f GB/s' % (N * 48 / (min(ts[4:]) * 1e-3) / 1e9))
for k in range(3):
    plane, inl = c.segment_plane(0.05, 3, 1000, seed=1)[:2]
    print('segment ms', ' '.join('%.3f' % t for t in c.timings()[:4]), 'inliers', len(inl))
