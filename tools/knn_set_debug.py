"""Which queries of the tier chain disagree with the oracle's neighbour sets, and how (debugging aid).
    python tools/knn_set_debug.py [bunny|lidar N] [mode radius]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
from oracle import orc
what = sys.argv[1] if len(sys.argv) > 1 else "bunny"
if what == "bunny":
    pts = np.load(os.path.join(ROOT, "tests", "golden", "bunny_xyz_f32.npy")).astype(np.float64)
else:
    pts = lidar_like(int(sys.argv[2]), 77)
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 0
radius = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
K = 30
c = nat.Context(0)
c.set_option(nat.OPT_KEEP_NEIGHBORS, 1)
c.upload(nat.ATTR_POINTS, pts)
c.estimate_normals(mode, K, radius)
idx, cnt, tier = c.fetch_neighbors(K)
eidx, ed2, ecnt = orc.knn(pts, K, mode=mode, radius=radius)
big = np.iinfo(np.int32).max
e = np.where(np.arange(K)[None, :] < ecnt[:, None], eidx, big); e.sort(1); e[e == big] = -1
bad = np.nonzero((idx != e).any(1) | (cnt != ecnt))[0]
print("n", len(pts), "bad", len(bad), "by tier", np.bincount(tier[bad], minlength=4).tolist(), "tiers", np.bincount(tier, minlength=4).tolist())
for q in bad[:12]:
    got, exp = set(idx[q][idx[q] >= 0].tolist()), set(e[q][e[q] >= 0].tolist())
    miss, extra = sorted(exp - got), sorted(got - exp)
    d = lambda j: float(((pts[j] - pts[q]) ** 2).sum())
    print(f"q {q} tier {tier[q]} cnt {cnt[q]}/{ecnt[q]} missing {[(j, d(j)) for j in miss][:4]} extra {[(j, d(j)) for j in extra][:4]} kth {ed2[q, ecnt[q]-1]:.6g} dup {len(idx[q][idx[q]>=0]) - len(got)}")
