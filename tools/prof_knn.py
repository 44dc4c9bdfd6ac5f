"""Minimal driver for ncu: estimate_normals(KNN k=30) on an N-point LiDAR-like scan, twice."""
import sys
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
pts = lidar_like(N, 1235)
c = nat.Context(0)
c.upload(nat.ATTR_POINTS, pts)
for _ in range(2):
    c.drop(nat.ATTR_NORMALS)
    c.estimate_normals(mode, 30, 0.5 if mode == 2 else 0.0)
    print('ms', [round(v, 3) for v in c.timings()[:6]])
