import sys, time
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
t = time.time(); pts = lidar_like(N, 1235); print('gen', time.time() - t)
c = nat.Context(0)
for rep in range(3):
    t = time.time(); c.upload(nat.ATTR_POINTS, pts); c.synchronize(); print('upload s', time.time() - t)
    t = time.time(); V = c.voxel(0.1); print('voxel s', time.time() - t, 'V', V, 'dev ms', c.timings()[0])
    c.drop(nat.ATTR_NORMALS)
    t = time.time(); c.estimate_normals(0, 30, 0.0); tm = c.timings(); print('normals s', time.time() - t, 'build ms', tm[0], 'search ms', tm[1])
    t = time.time(); c.estimate_normals(2, 30, 0.5); tm = c.timings(); print('hybrid s', time.time() - t, 'build ms', tm[0], 'search ms', tm[1])
    t = time.time(); pl, inl, dbg = c.segment_plane(0.05, 3, 1000, seed=1, debug=True); tm = c.timings(); print('segment s', time.time() - t, 'ms', tm[:3], 'scored', tm[7], 'inl', len(inl), pl)
    t = time.time(); c.transform(np.eye(4)); print('transform s', time.time() - t, 'ms', c.timings()[0])
    t = time.time(); nn = c.download(nat.ATTR_NORMALS); print('download s', time.time() - t)
import os
if os.environ.get('O3DPP_STATS') == '1':
    c.drop(nat.ATTR_NORMALS)
    c.estimate_normals(0, 30, 0.0); print('stats knn per query', c.debug_counters()[:8] / N)
    c.estimate_normals(2, 30, 0.5); print('stats hybrid per query', c.debug_counters()[:8] / N)
    b = np.load('tests/golden/bunny_xyz_f32.npy').astype(np.float64)
    c.upload(nat.ATTR_POINTS, b); c.estimate_normals(0, 30, 0.0); print('bunny ms', c.timings()[:2], 'stats', c.debug_counters()[:8] / len(b))
    u = np.random.default_rng(0).uniform(size=(1000000, 3))
    c.upload(nat.ATTR_POINTS, u); c.estimate_normals(0, 30, 0.0); print('uniform 1M ms', c.timings()[:2], 'stats', c.debug_counters()[:8] / len(u))
