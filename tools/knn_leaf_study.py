"""CPU study of the KNN staging geometry (no GPU): for Morton vs Hilbert ordering and a radius rule,
how many cell groups does a warp stage and what fraction of queries is certified complete."""
import sys, time
import numpy as np
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
from open3dpypro.synthetic import lidar_like
from scipy.spatial import cKDTree

def spread3(v):
    v = v.astype(np.uint64) & np.uint64(0x1fffff)
    v = (v | (v << np.uint64(32))) & np.uint64(0x1f00000000ffff)
    v = (v | (v << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
    v = (v | (v << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
    v = (v | (v << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
    v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
    return v

def hilbert3(c, bits):
    """Skilling's transpose algorithm: axes -> Hilbert index (vectorised)."""
    X = [c[:, 0].astype(np.uint64).copy(), c[:, 1].astype(np.uint64).copy(), c[:, 2].astype(np.uint64).copy()]
    M = np.uint64(1) << np.uint64(bits - 1)
    Q = M
    while Q > np.uint64(1):
        P = Q - np.uint64(1)
        for i in range(3):
            sel = (X[i] & Q) != 0
            X[0] = np.where(sel, X[0] ^ P, X[0])
            t = (X[0] ^ X[i]) & P
            t = np.where(sel, np.uint64(0), t)
            X[0] ^= t; X[i] ^= t
        Q >>= np.uint64(1)
    for i in range(1, 3):
        X[i] ^= X[i - 1]
    t = np.zeros_like(X[0])
    Q = M
    while Q > np.uint64(1):
        t = np.where((X[2] & Q) != 0, t ^ (Q - np.uint64(1)), t)
        Q >>= np.uint64(1)
    for i in range(3):
        X[i] ^= t
    # interleave: X[0] is the most significant
    return (spread3(X[0]) << np.uint64(2)) | (spread3(X[1]) << np.uint64(1)) | spread3(X[2])

def study(pts, order_kind, rho_rule, K=30, RMAX=24, nsample=1500, seed=0, G=32):
    n = len(pts)
    lg = int(np.ceil(np.log2(n)))
    bits = min(21, max(8, (lg + 1) // 2 + 4))
    mn, mx = pts.min(0), pts.max(0)
    ext = (mx - mn).max()
    cells = np.minimum(((pts - mn) * ((1 << bits) / (ext * 1.000000001))).astype(np.int64), (1 << bits) - 1)
    if order_kind == 'morton':
        key = spread3(cells[:, 0]) | (spread3(cells[:, 1]) << np.uint64(1)) | (spread3(cells[:, 2]) << np.uint64(2))
    else:
        key = hilbert3(cells, bits)
    order = np.argsort(key, kind='stable')
    sp = pts[order]
    nleaf = (n + G - 1) // G
    pad = nleaf * G - n
    spp = np.concatenate([sp, np.repeat(sp[-1:], pad, 0)], 0).reshape(nleaf, G, 3)
    lo, hi = spp.min(1), spp.max(1)
    tree = cKDTree(sp)
    rng = np.random.default_rng(seed)
    sel = rng.choice(nleaf - 1, size=min(nsample, nleaf - 1), replace=False)
    dk = tree.query(spp[sel].reshape(-1, 3), k=K)[0][:, -1].reshape(len(sel), G)
    ext3 = hi[sel] - lo[sel]
    diag = np.linalg.norm(ext3, axis=1)
    es = np.sort(ext3, axis=1)[:, ::-1]
    if rho_rule[0] == 'diag':
        rho = rho_rule[1] * diag
    else:
        r2d = np.sqrt(K * es[:, 0] * es[:, 1] / (G * np.pi))
        r3d = np.cbrt(3 * K * es[:, 0] * es[:, 1] * es[:, 2] / (4 * np.pi * G))
        rho = rho_rule[1] * np.maximum(r2d, r3d)
    nn = np.empty(len(sel), np.int64); nl = np.empty(len(sel), np.int64); okfrac = np.empty(len(sel))
    for j, s in enumerate(sel):
        g = np.maximum(np.maximum(lo - hi[s], lo[s] - hi), 0.0)
        gap = np.sqrt((g * g).sum(1))
        gap[s] = np.inf
        within = np.sort(gap[gap <= rho[j]])
        nn[j] = len(within)
        if len(within) > RMAX - 1:
            rho1 = within[RMAX - 1] * 0.999999   # nearest excluded
            nl[j] = RMAX
        else:
            rho1 = rho[j]; nl[j] = 1 + len(within)
        okfrac[j] = (dk[j] <= rho1).mean()
    print(f"{order_kind:8s} G={G} rule={rho_rule} RMAX={RMAX}: nn mean {nn.mean():.1f} p50 {np.median(nn):.0f} p90 {np.quantile(nn,0.9):.0f} | nl mean {nl.mean():.1f} | "
          f"certified-complete {okfrac.mean():.4f} | dk/diag med {np.median(dk/diag[:,None]):.3f} p99 {np.quantile(dk/diag[:,None],0.99):.3f}")
    return nl.mean(), okfrac.mean()

if __name__ == '__main__':
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    pts = lidar_like(N, 1235)
    for kind in (('morton', 'hilbert') if len(sys.argv) <= 2 else ()):
        for rule in (('diag', 0.5), ('diag', 0.7), ('dens', 1.25), ('dens', 1.6)):
            for RMAX in (24, 32):
                study(pts, kind, rule, RMAX=RMAX)


def study_points(pts, order_kind, rho_scale, K=30, CAP=512, nsample=400, seed=0, G=32):
    """staging by POINT filter: candidates = points inside the home box expanded by rho (max-norm)."""
    n = len(pts)
    lg = int(np.ceil(np.log2(n)))
    bits = min(21, max(8, (lg + 1) // 2 + 4))
    mn, mx = pts.min(0), pts.max(0)
    ext = (mx - mn).max()
    cells = np.minimum(((pts - mn) * ((1 << bits) / (ext * 1.000000001))).astype(np.int64), (1 << bits) - 1)
    if order_kind == 'morton':
        key = spread3(cells[:, 0]) | (spread3(cells[:, 1]) << np.uint64(1)) | (spread3(cells[:, 2]) << np.uint64(2))
    else:
        key = hilbert3(cells, bits)
    order = np.argsort(key, kind='stable')
    sp = pts[order]
    nleaf = (n + G - 1) // G
    pad = nleaf * G - n
    spp = np.concatenate([sp, np.repeat(sp[-1:], pad, 0)], 0).reshape(nleaf, G, 3)
    lo, hi = spp.min(1), spp.max(1)
    tree = cKDTree(sp)
    rng = np.random.default_rng(seed)
    sel = rng.choice(nleaf - 1, size=min(nsample, nleaf - 1), replace=False)
    dk = tree.query(spp[sel].reshape(-1, 3), k=K)[0][:, -1].reshape(len(sel), G)
    diag = np.linalg.norm(hi[sel] - lo[sel], axis=1)
    nc = np.empty(len(sel), np.int64); ok = np.empty(len(sel)); nleaves = np.empty(len(sel), np.int64)
    for j, s in enumerate(sel):
        rho = rho_scale * diag[j]
        cen = 0.5 * (lo[s] + hi[s]); half = 0.5 * (hi[s] - lo[s])
        while True:
            cand = np.asarray(tree.query_ball_point(cen, half.max() + rho, p=np.inf), dtype=np.int64)
            q = sp[cand]
            c = int(((q >= lo[s] - rho) & (q <= hi[s] + rho)).all(1).sum())
            if c <= CAP or rho < 0.02 * diag[j]:
                if c > CAP: rho = 0.0
                break
            rho *= 0.8
        g = np.maximum(np.maximum(lo - hi[s], lo[s] - hi), 0.0)
        nleaves[j] = int((np.sqrt((g * g).sum(1)) <= rho).sum())
        nc[j] = c
        ok[j] = (dk[j] <= rho).mean()
    print(f"{order_kind:8s} point-filter rho={rho_scale} CAP={CAP}: ncand mean {nc.mean():.0f} p50 {np.median(nc):.0f} p90 {np.quantile(nc,0.9):.0f} p99 {np.quantile(nc,0.99):.0f} "
          f"rows mean {np.ceil(np.minimum(nc,CAP)/32).mean():.1f} overflow {np.mean(nc>CAP):.3f} | leaves scanned mean {nleaves.mean():.1f} p90 {np.quantile(nleaves,0.9):.0f} | certified-complete {ok.mean():.4f}")


if __name__ == '__main__' and len(sys.argv) > 2 and sys.argv[2] == 'points':
    for kind in ('morton', 'hilbert'):
        for rs in (0.4, 0.5, 0.6):
            study_points(pts, kind, rs); sys.stdout.flush()
