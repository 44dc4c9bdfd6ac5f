import sys, time
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
pts = lidar_like(131072, 2000)
c = nat.Context(0)
def T(f, n=20):
    f(); c.synchronize()
    t=time.perf_counter()
    for _ in range(n): f()
    c.synchronize()
    return (time.perf_counter()-t)/n*1e3
print("upload f32", T(lambda: c.upload(nat.ATTR_POINTS, pts)))
def nrm(): c.drop(nat.ATTR_NORMALS); c.estimate_normals(0, 30, 0.0)
print("normals", T(nrm), [round(v,3) for v in c.timings()[:7]])
l0=c.launch_count(); nrm(); print("launches normals", c.launch_count()-l0)
print("segment", T(lambda: c.segment_plane(0.05, 3, 1000, seed=7, fetch_inliers=False)), [round(v,3) for v in c.timings()])
l0=c.launch_count(); c.segment_plane(0.05, 3, 1000, seed=7, fetch_inliers=False); print("launches segment", c.launch_count()-l0)
print("download normals", T(lambda: c.download(nat.ATTR_NORMALS)))
Tm=np.eye(4)
print("transform", T(lambda: c.transform(Tm)))
