"""CPU study (no GPU): how many candidates does a 32-query cell group need under different staging shapes?"""
import sys
import numpy as np
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.'); sys.path.insert(0, 'tools')
from open3dpypro.synthetic import lidar_like
from knn_leaf_study import hilbert3
from scipy.spatial import cKDTree

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
K, G = 30, 32
pts = lidar_like(N, 1235).astype(np.float64)
n = len(pts)
lg = int(np.ceil(np.log2(n))); bits = min(21, max(8, (lg + 1) // 2 + 4))
mn = pts.min(0); ext = (pts.max(0) - mn).max()
cells = np.minimum(((pts - mn) * ((1 << bits) / (ext * 1.000000001))).astype(np.int64), (1 << bits) - 1)
sp = pts[np.argsort(hilbert3(cells, bits), kind='stable')]
tree = cKDTree(sp)
rng = np.random.default_rng(0)
nleaf = n // G
sel = rng.choice(nleaf, 1500, replace=False)
rows = []
for s in sel:
    q = sp[s * G:(s + 1) * G]
    d, idx = tree.query(q, k=K)
    rk = d[:, -1]
    ideal = len(np.unique(idx))
    lo, hi = q.min(0), q.max(0)
    diag = np.linalg.norm(hi - lo)
    out = [ideal]
    for scale in (0.4, 0.5):
        rho = scale * diag
        cand = np.array(tree.query_ball_point((lo + hi) / 2, np.linalg.norm((hi - lo) / 2 + rho) + 1e-9))
        P = sp[cand]
        inbox = np.all((P >= lo - rho) & (P <= hi + rho), axis=1)
        sub = np.zeros(len(P), bool)
        for a in range(0, G, 8):
            l2, h2 = q[a:a + 8].min(0), q[a:a + 8].max(0)
            sub |= np.all((P >= l2 - rho) & (P <= h2 + rho), axis=1)
        dmin = np.sqrt(((P[:, None, :] - q[None, :, :]) ** 2).sum(2).min(1))
        ball = dmin <= rho
        ok = (rk <= rho).mean()
        out += [inbox.sum(), sub.sum(), ball.sum(), ok]
    # oracle radius: 1.05 * the largest k-th distance of the group
    rho = 1.05 * rk.max()
    cand = np.array(tree.query_ball_point((lo + hi) / 2, np.linalg.norm((hi - lo) / 2 + rho) + 1e-9))
    P = sp[cand]
    out += [np.all((P >= lo - rho) & (P <= hi + rho), axis=1).sum(), rk.max() / diag, np.median(rk) / diag]
    rows.append(out)
R = np.array(rows, float)
names = ['ideal union', 'box .4', 'subbox .4', 'balls .4', 'okfrac .4', 'box .5', 'subbox .5', 'balls .5', 'okfrac .5', 'box oracle-rho', 'rkmax/diag', 'rkmed/diag']
for i, nme in enumerate(names):
    c = R[:, i]
    print(f"{nme:16s} mean {c.mean():8.2f}  median {np.median(c):8.2f}  p90 {np.quantile(c, .9):8.2f}  p99 {np.quantile(c, .99):8.2f}")
