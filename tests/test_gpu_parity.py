"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.

Bars (BASELINE.json north_star): voxel keys / rows, KNN index sets under (d2, index) order and
RANSAC inlier sets for a given hypothesis sequence are bit-exact; centroids within 1e-5 relative
(they are in fact bit-exact here); normals within 1e-4 rad after sign alignment.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like


def _angle(a, b):
    c = np.abs(np.einsum("ij,ij->i", a, b)) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))
    return np.arccos(np.clip(c, -1.0, 1.0))


def _rand_T(rng):
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    T = np.eye(4)
    T[:3, :3] = q
    T[:3, 3] = rng.normal(size=3)
    return T


# ------------------------------------------------------------------ state / transform
def test_upload_download_roundtrip(ctx, bunny):
    ctx.upload(nat.ATTR_POINTS, bunny)
    assert ctx.size() == len(bunny)
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), bunny)
    ctx.upload(nat.ATTR_POINTS, bunny.astype(np.float32))
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), bunny)
    mn, mx = ctx.bounds()
    assert np.array_equal(mn, bunny.min(0)) and np.array_equal(mx, bunny.max(0))


@pytest.mark.parametrize("n", [1, 2, 3, 255, 256, 257, 35947])
def test_transform_bit_exact(ctx, orc, bunny, n):
    rng = np.random.default_rng(n)
    pts = bunny[:n] if n <= len(bunny) else rng.normal(size=(n, 3))
    nrm = rng.normal(size=(n, 3))
    T = _rand_T(rng)
    T[3] = [0.01, -0.02, 0.03, 1.5]  # projective row: exercises the divide by w
    ctx.upload(nat.ATTR_POINTS, pts)
    ctx.upload(nat.ATTR_NORMALS, nrm)
    ctx.transform(T)
    ep, en = orc.transform(pts, T, nrm)
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), ep)
    assert np.array_equal(ctx.download(nat.ATTR_NORMALS), en)
    ctx.drop(nat.ATTR_NORMALS)


def test_transform_affine_fast_path_and_non_finite(ctx, orc):
    """affine T: the division by w == 1.0 is skipped (r / 1.0 == r bit for bit); a non-finite coordinate makes w NaN and
    takes the division, so the result is the oracle's NaN pattern too"""
    rng = np.random.default_rng(17)
    pts = rng.normal(scale=50.0, size=(1001, 3))
    pts[5] = [np.inf, 1.0, 2.0]
    pts[77] = [0.5, np.nan, -3.0]
    pts[500] = [1e308, -1e308, 1e308]
    T = _rand_T(rng)
    ctx.upload(nat.ATTR_POINTS, pts)
    ctx.transform(T)
    with np.errstate(all="ignore"):
        ep, _ = orc.transform(pts, T)
    got = ctx.download(nat.ATTR_POINTS)
    assert np.array_equal(got, ep, equal_nan=True)
    assert np.array_equal(np.isnan(got), np.isnan(ep))


def test_transform_points_leaves_normals(ctx, orc, bunny):
    """o3dpp_transform_points: PlaneNormalize rotates pcd[:, :3] only (reference processors.py:725-744)"""
    rng = np.random.default_rng(18)
    nrm = rng.normal(size=bunny.shape)
    T = _rand_T(rng)
    ctx.upload(nat.ATTR_POINTS, bunny)
    ctx.upload(nat.ATTR_NORMALS, nrm)
    ctx.transform_points(T)
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), orc.transform(bunny, T)[0])
    assert np.array_equal(ctx.download(nat.ATTR_NORMALS), nrm)
    rows = np.empty((len(bunny), 6))
    ctx.download_rows6(rows)
    assert np.array_equal(rows[:, :3], orc.transform(bunny, T)[0]) and np.array_equal(rows[:, 3:], nrm)
    r32 = np.empty((len(bunny), 6), np.float32)
    ctx.download_rows6(r32)
    assert np.array_equal(r32, rows.astype(np.float32))
    ctx.drop(nat.ATTR_NORMALS)


def test_fp64_peak_is_plausible(ctx):
    tf = ctx.fp64_peak()
    assert 5.0 < tf < 200.0          # B200: ~34 TFLOP/s of FP64 FMA measured


# ------------------------------------------------------------------ voxel
def _check_voxel(ctx, orc, pts, vs, normals=None, colors=None):
    ctx.upload(nat.ATTR_POINTS, pts)
    if normals is not None:
        ctx.upload(nat.ATTR_NORMALS, normals)
    if colors is not None:
        ctx.upload(nat.ATTR_COLORS, colors)
    V = ctx.voxel(vs)
    ref = orc.voxel_trace(pts, vs, normals, colors)
    assert V == ref["V"]
    fields = ["rep_asc", "rep", "count", "centroid", "cubic_id", "order", "seg"]
    if normals is not None:
        fields.append("normal")
    if colors is not None:
        fields.append("color")
    got = ctx.voxel_fetch(V, len(pts), fields)
    assert np.array_equal(got["seg"], ref["seg"])
    assert np.array_equal(got["order"], ref["order"])
    assert np.array_equal(got["rep"], ref["rep"])
    assert np.array_equal(got["rep_asc"], np.sort(ref["rep"]))
    assert np.array_equal(got["count"], ref["count"])
    assert np.array_equal(got["cubic_id"], ref["cubic_id"])
    assert np.array_equal(got["centroid"], ref["centroid"])  # same summation order -> bit exact
    if normals is not None:
        assert np.array_equal(got["normal"], ref["normal"])
    if colors is not None:
        assert np.array_equal(got["color"], ref["color"])
    ctx.drop(nat.ATTR_NORMALS)
    ctx.drop(nat.ATTR_COLORS)
    return V


@pytest.mark.parametrize("vs,V", [(0.002, 15930), (0.005, 3017), (0.01, 751)])
def test_voxel_bunny(ctx, orc, bunny, vs, V):
    rng = np.random.default_rng(0)
    got = _check_voxel(ctx, orc, bunny, vs, rng.normal(size=bunny.shape), rng.uniform(size=bunny.shape))
    assert got == V


def test_voxel_edge_cases(ctx, orc):
    rng = np.random.default_rng(5)
    _check_voxel(ctx, orc, rng.normal(size=(1, 3)), 0.1)
    _check_voxel(ctx, orc, np.repeat(rng.normal(size=(1, 3)), 100, 0), 0.1)      # all in one voxel
    _check_voxel(ctx, orc, rng.uniform(-1, 1, size=(5000, 3)), 1e-7)             # every point alone, wide keys
    _check_voxel(ctx, orc, rng.uniform(-1, 1, size=(5000, 3)), 1e-4)
    _check_voxel(ctx, orc, rng.uniform(-1, 1, size=(4097, 3)), 0.25)
    pts = rng.uniform(-1, 1, size=(100, 3))
    ctx.upload(nat.ATTR_POINTS, pts)
    with pytest.raises(nat.O3dppError, match="voxel_size <= 0"):
        ctx.voxel(0.0)
    with pytest.raises(nat.O3dppError, match="too small"):
        ctx.voxel(1e-12)


def test_voxel_select(ctx, orc, bunny):
    ctx.upload(nat.ATTR_POINTS, bunny)
    V = ctx.voxel(0.005)
    ref = orc.voxel_trace(bunny, 0.005)
    ctx.voxel_select()
    assert ctx.size() == V
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), bunny[np.sort(ref["rep"])])


def test_voxel_lidar_1m(ctx, orc):
    pts = lidar_like(1_000_000, 1234)
    _check_voxel(ctx, orc, pts, 0.1)


# ------------------------------------------------------------------ KNN
def _check_knn(ctx, orc, pts, k, mode=0, radius=0.0):
    ctx.upload(nat.ATTR_POINTS, pts)
    idx, d2, cnt = ctx.knn(mode, k, radius)
    eidx, ed2, ecnt = orc.knn(pts, k, mode=mode, radius=radius)
    assert np.array_equal(cnt, ecnt)
    bad = np.nonzero((idx != eidx).any(1))[0]
    assert bad.size == 0, f"{bad.size} rows differ, first {bad[:5]}: {idx[bad[:1]]} vs {eidx[bad[:1]]}"
    assert np.array_equal(d2, ed2)


def test_knn_bunny_k30(ctx, orc, bunny):
    _check_knn(ctx, orc, bunny, 30)


@pytest.mark.parametrize("k", [1, 2, 16, 31, 32, 33, 50, 64, 65, 100, 128])
def test_knn_k_sweep(ctx, orc, bunny, k):
    _check_knn(ctx, orc, bunny[::7], k)


def test_knn_hybrid(ctx, orc, bunny):
    _check_knn(ctx, orc, bunny, 30, mode=2, radius=0.004)
    _check_knn(ctx, orc, bunny, 30, mode=2, radius=0.0005)
    _check_knn(ctx, orc, bunny, 30, mode=2, radius=10.0)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 29, 30, 31, 32, 33, 63, 64, 65, 1000, 1025])
def test_knn_small_clouds(ctx, orc, n):
    rng = np.random.default_rng(n)
    _check_knn(ctx, orc, rng.normal(size=(n, 3)), 30)


def test_knn_ties_and_duplicates(ctx, orc):
    rng = np.random.default_rng(11)
    # integer lattice: massive exact d2 ties, index order decides
    g = np.stack(np.meshgrid(np.arange(12), np.arange(12), np.arange(12), indexing="ij"), -1).reshape(-1, 3).astype(np.float64)
    _check_knn(ctx, orc, g[rng.permutation(len(g))], 30)
    # duplicated points
    p = rng.normal(size=(500, 3))
    p = np.concatenate([p, p[:200], p[:50]], 0)
    _check_knn(ctx, orc, p[rng.permutation(len(p))], 30)
    # all points identical
    _check_knn(ctx, orc, np.ones((100, 3)), 30)


def test_knn_lidar_200k(ctx, orc):
    _check_knn(ctx, orc, lidar_like(200_000, 77), 30)


# ------------------------------------------------------------------ normals
def _check_normals(ctx, orc, pts, mode, k, radius, old=None, exact=True):
    ctx.upload(nat.ATTR_POINTS, pts)
    if old is not None:
        ctx.upload(nat.ATTR_NORMALS, old)
    else:
        ctx.drop(nat.ATTR_NORMALS)
    ctx.estimate_normals(mode, k, radius)
    got = ctx.download(nat.ATTR_NORMALS)
    ref, cov = orc.estimate_normals(pts, mode={0: "knn", 1: "radius", 2: "hybrid"}[mode], k=k, radius=radius,
                                    old_normals=old, want_cov=True)
    # conditioning of every 3x3 problem: the closed form loses the direction when the two smallest
    # eigenvalues coincide (collinear neighbourhoods), so the angle bar applies where the gap is real
    Cm = np.empty((len(pts), 3, 3))
    Cm[:, 0, 0], Cm[:, 0, 1], Cm[:, 0, 2] = cov[:, 0], cov[:, 1], cov[:, 2]
    Cm[:, 1, 0], Cm[:, 1, 1], Cm[:, 1, 2] = cov[:, 1], cov[:, 3], cov[:, 4]
    Cm[:, 2, 0], Cm[:, 2, 1], Cm[:, 2, 2] = cov[:, 2], cov[:, 4], cov[:, 5]
    lam = np.linalg.eigvalsh(Cm)
    scale = np.maximum(np.abs(lam).max(1), 1e-300)
    well = (lam[:, 1] - lam[:, 0]) / scale > 1e-6
    ang = _angle(got, ref)
    assert np.isfinite(got).all()
    assert np.abs(np.linalg.norm(got, axis=1) - 1.0).max() < 1e-9
    assert ang[well].max(initial=0.0) < 1e-4, f"max angle {ang[well].max()}"
    # everywhere: the returned direction is as good a minimiser of n^T C n as the oracle's
    ray_got = np.einsum("ni,nij,nj->n", got, Cm, got)
    ray_ref = np.einsum("ni,nij,nj->n", ref, Cm, ref)
    assert ((ray_got - ray_ref) / scale).max() < 1e-9
    if old is not None:  # sign must follow the old normals exactly as the oracle does
        assert (np.einsum("ij,ij->i", got, ref)[well] > 0).all()
    if exact:
        # same neighbour order, same cumulant order, same solver: only libm acos/cos may differ
        assert np.abs(got[well] - ref[well]).max(initial=0.0) < 1e-7
    ctx.drop(nat.ATTR_NORMALS)
    return well.mean()


def test_normals_bunny_knn(ctx, orc, bunny):
    _check_normals(ctx, orc, bunny, 0, 30, 0.0)


def test_normals_bunny_hybrid(ctx, orc, bunny):
    _check_normals(ctx, orc, bunny, 2, 30, 0.01)   # test_mesh.py:17-19 of the reference
    _check_normals(ctx, orc, bunny, 2, 30, 0.002)


def test_normals_bunny_radius(ctx, orc, bunny):
    _check_normals(ctx, orc, bunny, 1, 0, 0.004, exact=False)


def test_normals_sign_alignment(ctx, orc, bunny):
    rng = np.random.default_rng(3)
    _check_normals(ctx, orc, bunny, 0, 30, 0.0, old=rng.normal(size=bunny.shape))


def test_normals_degenerate(ctx, orc):
    rng = np.random.default_rng(4)
    _check_normals(ctx, orc, rng.normal(size=(2, 3)), 0, 30, 0.0)             # < 3 neighbours -> (0,0,1)
    _check_normals(ctx, orc, np.ones((50, 3)), 0, 30, 0.0)                    # zero covariance
    line = np.outer(np.arange(100.0), [1.0, 0, 0])
    _check_normals(ctx, orc, line, 0, 30, 0.0, exact=False)                   # rank-1 covariance
    plane = np.concatenate([rng.uniform(size=(400, 2)), np.zeros((400, 1))], 1)
    _check_normals(ctx, orc, plane, 0, 30, 0.0)


def test_normals_lidar_200k(ctx, orc):
    _check_normals(ctx, orc, lidar_like(200_000, 78), 0, 30, 0.0)
    _check_normals(ctx, orc, lidar_like(200_000, 78), 2, 30, 0.5)


def _adversarial(n_each=40_000, seed=21, offset=0.0):
    """what breaks a grid: uniform volume + duplicated surface points + tight clusters far apart + an exact lattice
    (d2 ties) + a few far outliers, shuffled"""
    rng = np.random.default_rng(seed)
    vol = rng.uniform(-20, 20, size=(n_each, 3))
    surf = np.concatenate([rng.uniform(-30, 30, size=(n_each // 4, 2)), np.zeros((n_each // 4, 1))], 1)
    surf = np.repeat(surf, 4, axis=0)                                                   # every point four times
    cen = rng.uniform(-50, 50, size=(40, 3))
    clus = cen[rng.integers(0, 40, n_each)] + rng.normal(scale=1e-3, size=(n_each, 3))  # 1 mm clusters, 100 m apart
    g = np.arange(34, dtype=np.float64) * 0.25
    lat = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3) + 60.0       # 39 304 lattice points
    far = rng.uniform(-1000, 1000, size=(50, 3))
    pts = np.concatenate([vol, surf, clus, lat, far], 0) + offset
    return pts[rng.permutation(len(pts))]


def test_knn_and_normals_adversarial(ctx, orc):
    pts = _adversarial()
    _check_knn(ctx, orc, pts[::4], 30)                 # index sets incl. exact ties, list kernel
    _check_normals(ctx, orc, pts, 0, 30, 0.0, exact=False)           # tier chain
    _check_normals(ctx, orc, pts, 2, 30, 0.3, exact=False)
    _check_normals(ctx, orc, _adversarial(10_000, 22, offset=5000.0), 0, 30, 0.0, exact=False)   # large coordinates


# ------------------------------------------------------------------ segment_plane
def _check_segment(ctx, orc, pts, thr, rn, H, seed=7, prob=0.99999999):
    ctx.upload(nat.ATTR_POINTS, pts)
    plane, inl, dbg = ctx.segment_plane(thr, rn, H, prob, sampler=nat.SAMPLER_PHILOX, seed=seed, debug=True)
    samples = orc.sample_philox(seed, len(pts), rn, H)
    assert np.array_equal(ctx.fetch_samples(H, rn), samples)          # device Philox == CPU replay
    eplane, einl, edbg = orc.segment_plane(pts, thr, rn, H, samples, prob=prob, debug=True)
    assert np.array_equal(dbg["counts"], edbg["counts"])               # incl. -1 (early exit) and -2 (degenerate)
    assert dbg["best_iter"] == edbg["best_iter"]
    assert np.array_equal(dbg["best_plane"], edbg["best_plane"])       # plane models bit-exact
    assert np.array_equal(inl, einl)
    sc = np.maximum(np.abs(edbg["errs"]), 1e-300)
    assert (np.abs(dbg["errs"] - edbg["errs"]) / sc).max() < 1e-10
    assert np.abs(plane - eplane).max() < 1e-9
    return dbg


def test_segment_bunny(ctx, orc, bunny):
    _check_segment(ctx, orc, bunny, 0.01, 3, 450)
    _check_segment(ctx, orc, bunny, 0.005, 3, 100, seed=99)


def test_segment_ransac_n(ctx, orc, bunny):
    _check_segment(ctx, orc, bunny, 0.01, 5, 200, seed=3)


def test_segment_early_exit_lidar(ctx, orc):
    pts = lidar_like(300_000, 1234)
    dbg = _check_segment(ctx, orc, pts, 0.05, 3, 1000, seed=1)
    assert (dbg["counts"] == -1).any()      # the ground plane is found early and the rest is skipped


def test_segment_samplers(ctx, orc, bunny):
    ctx.upload(nat.ATTR_POINTS, bunny)
    H = 64
    s = orc.sample_mt19937(5, len(bunny), 3, H)
    ctx.segment_plane(0.01, 3, H, sampler=nat.SAMPLER_MT19937, seed=5)
    assert np.array_equal(ctx.fetch_samples(H, 3), s)
    p1, i1 = ctx.segment_plane(0.01, 3, H, samples=s)
    p2, i2 = orc.segment_plane(bunny, 0.01, 3, H, s)
    assert np.array_equal(i1, i2) and np.abs(p1 - p2).max() < 1e-9


def test_segment_errors_and_degenerate(ctx, orc):
    rng = np.random.default_rng(1)
    pts = rng.normal(size=(100, 3))
    ctx.upload(nat.ATTR_POINTS, pts)
    with pytest.raises(nat.O3dppError, match="ransac_n"):
        ctx.segment_plane(0.1, 2, 10)
    with pytest.raises(nat.O3dppError, match="Probability"):
        ctx.segment_plane(0.1, 3, 10, probability=0.0)
    ctx.upload(nat.ATTR_POINTS, pts[:2])
    with pytest.raises(nat.O3dppError, match="at least"):
        ctx.segment_plane(0.1, 3, 10)
    # collinear points: every hypothesis is degenerate -> zero plane, no inliers
    line = np.outer(np.arange(50.0), [1.0, 2.0, 3.0])
    _check_segment(ctx, orc, line, 0.1, 3, 20)
    # perfect plane: fitness 1 -> break after the first hypothesis
    flat = np.concatenate([rng.uniform(size=(1000, 2)), np.zeros((1000, 1))], 1)
    _check_segment(ctx, orc, flat, 0.01, 3, 50)


def test_select(ctx, bunny):
    ctx.upload(nat.ATTR_POINTS, bunny)
    idx = np.array([5, 3, 3, 100, 35946], np.int32)
    ctx.select(idx)
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), bunny[np.unique(idx)])
    ctx.upload(nat.ATTR_POINTS, bunny)
    ctx.select(idx, invert=True)
    m = np.ones(len(bunny), bool)
    m[idx] = False
    assert np.array_equal(ctx.download(nat.ATTR_POINTS), bunny[m])
