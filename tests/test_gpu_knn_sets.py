"""Bit-exact neighbour index SETS of the estimate_normals tier chain (knn_lane_kernel -> knn_range_kernel -> knn_exact_kernel)
against the oracle's KD-tree, the call this replaces being pcd.estimate_normals(param) at reference PointCloud.py:68-73.

north_star: "KNN index sets (ties broken by index) ... bit-exact".  o3dpp_knn (the sorted-list API) runs another kernel,
so these tests read the sets the chain itself selected (O3DPP_OPT_KEEP_NEIGHBORS + o3dpp_fetch_neighbors) and compare
them, per query, with orc.knn under the (d2, index) rule: on the bunny (incl. the 137 points with an exact d2 tie between
neighbours 30 and 31), an integer lattice, duplicates, the adversarial mixture, a 200 k LiDAR-like scan, in KNN and hybrid
mode (also with fewer than k points inside the radius), with forced shrink / window paths, and 2 000 sampled queries of the
10 M-point scan of config C3."""
import numpy as np
import pytest

from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like, SEED_C3
from test_gpu_parity import _adversarial

pytestmark = pytest.mark.gpu

K = 30


@pytest.fixture()
def kctx(ctx):
    ctx.set_option(nat.OPT_KEEP_NEIGHBORS, 1)
    yield ctx
    for o in (nat.OPT_KEEP_NEIGHBORS, nat.OPT_KNN_GSPLIT, nat.OPT_KNN_MAXSHRINK, nat.OPT_KNN_RHO, nat.OPT_KNN_SHRINK):
        ctx.set_option(o, -1 if o == nat.OPT_KNN_MAXSHRINK else 0)


def _expected_sets(orc, pts, k, mode, radius, queries=None):
    eidx, _, ecnt = orc.knn(pts, k, queries=None if queries is None else pts[queries], mode=mode, radius=radius)
    big = np.iinfo(np.int32).max
    e = np.where(np.arange(k)[None, :] < ecnt[:, None], eidx, big)
    e.sort(axis=1)
    e[e == big] = -1
    return e, ecnt


def _check_sets(ctx, orc, pts, k=K, mode=0, radius=0.0, queries=None, want_tiers=()):
    ctx.upload(nat.ATTR_POINTS, pts)
    ctx.drop(nat.ATTR_NORMALS)
    ctx.estimate_normals(mode, k, radius)
    idx, cnt, tier = ctx.fetch_neighbors(k, queries)
    e, ecnt = _expected_sets(orc, pts, k, mode, radius, queries)
    assert np.array_equal(cnt, ecnt)
    bad = np.nonzero((idx != e).any(1))[0]
    assert bad.size == 0, f"{bad.size} of {len(e)} neighbour sets differ (tiers {np.bincount(tier[bad], minlength=4)}), first {bad[:3]}: {idx[bad[:1]]} vs {e[bad[:1]]}"
    assert ((tier >= 1) & (tier <= 3)).all()
    for t in want_tiers:
        assert (tier == t).any(), f"tier {t} answered no query of this case"
    return np.bincount(tier, minlength=4)


def test_sets_bunny_knn(kctx, orc, bunny):
    _check_sets(kctx, orc, bunny, want_tiers=(1, 2))


def test_sets_bunny_hybrid(kctx, orc, bunny):
    _check_sets(kctx, orc, bunny, mode=2, radius=0.01)
    _check_sets(kctx, orc, bunny, mode=2, radius=0.002)      # fewer than k inside the radius for most points: take-all path
    _check_sets(kctx, orc, bunny, mode=2, radius=0.0005)
    _check_sets(kctx, orc, bunny, mode=2, radius=10.0)


@pytest.mark.parametrize("k", [1, 2, 3, 8, 16, 29, 31, 32])
def test_sets_k_sweep(kctx, orc, bunny, k):
    _check_sets(kctx, orc, bunny[::3], k=k)


@pytest.mark.parametrize("n", [1, 2, 5, 29, 30, 31, 32, 33, 63, 64, 65, 1000, 1025])
def test_sets_small_clouds(kctx, orc, n):
    _check_sets(kctx, orc, np.random.default_rng(n).normal(size=(n, 3)))


def test_sets_ties_and_duplicates(kctx, orc):
    rng = np.random.default_rng(11)
    g = np.stack(np.meshgrid(np.arange(12), np.arange(12), np.arange(12), indexing="ij"), -1).reshape(-1, 3).astype(np.float64)
    _check_sets(kctx, orc, g[rng.permutation(len(g))])                       # integer lattice: massive exact ties
    p = rng.normal(size=(500, 3))
    p = np.concatenate([p, p[:200], p[:50]], 0)
    _check_sets(kctx, orc, p[rng.permutation(len(p))])                       # duplicated points
    _check_sets(kctx, orc, np.ones((100, 3)))                                # all identical


def test_sets_adversarial(kctx, orc):
    pts = _adversarial()
    _check_sets(kctx, orc, pts, want_tiers=(1, 2, 3))
    _check_sets(kctx, orc, pts, mode=2, radius=0.3)
    _check_sets(kctx, orc, _adversarial(10_000, 22, offset=5000.0))          # large coordinates: wide FP32 error margins


def test_sets_lidar_200k(kctx, orc):
    pts = lidar_like(200_000, 77)
    _check_sets(kctx, orc, pts, want_tiers=(1, 2, 3))
    _check_sets(kctx, orc, pts, mode=2, radius=0.5, want_tiers=(1, 2))


def test_sets_forced_paths(kctx, orc):
    """the rarely taken branches of tier 1: staging attempts abandoned early (small gsplit), parts halved instead of
    shrunk (max_shrink < 12), wide and narrow staging radii"""
    pts = lidar_like(120_000, 79)
    kctx.set_option(nat.OPT_KNN_GSPLIT, 16)
    _check_sets(kctx, orc, pts)
    kctx.set_option(nat.OPT_KNN_GSPLIT, 0)
    kctx.set_option(nat.OPT_KNN_MAXSHRINK, 1)
    _check_sets(kctx, orc, pts)
    kctx.set_option(nat.OPT_KNN_MAXSHRINK, 0)
    _check_sets(kctx, orc, pts, mode=2, radius=0.4)
    kctx.set_option(nat.OPT_KNN_MAXSHRINK, -1)
    for rho in (0.2, 0.8):
        kctx.set_option(nat.OPT_KNN_RHO, rho)
        _check_sets(kctx, orc, pts)
    kctx.set_option(nat.OPT_KNN_RHO, 0)
    # curve keys of 48 bits do not leave room for the index in the key: the (key, index) pair sort instead of the packed one
    kctx.set_option(nat.OPT_KNN_BITS, 16)
    try:
        _check_sets(kctx, orc, pts)
    finally:
        kctx.set_option(nat.OPT_KNN_BITS, 0)


@pytest.mark.parametrize("mode,radius", [(0, 0.0), (2, 0.5)])
def test_sets_c3_sample(kctx, orc, mode, radius):
    """config C3: 2 000 sampled queries inside the whole 10 M-point scan, KNN and hybrid"""
    scan = lidar_like(10_000_000, SEED_C3)
    qi = np.random.default_rng(5).choice(len(scan), 2000, replace=False)
    counts = _check_sets(kctx, orc, scan, mode=mode, radius=radius, queries=qi)
    assert counts[1] > 1500          # most of the sample is answered by the certified FP32 selection


@pytest.mark.parametrize("mode,radius", [(0, 0.0), (2, 0.5)])
def test_query_first_answers_only_the_first_points(kctx, orc, mode, radius):
    """O3DPP_OPT_QUERY_FIRST = m (second pass of the slab-sharded pure-KNN scan): the first m points get exactly the
    normals and neighbour sets of the full call, whichever tier answers them; the others are not queried"""
    pts = lidar_like(150_000, 81)
    kctx.upload(nat.ATTR_POINTS, pts)
    kctx.drop(nat.ATTR_NORMALS)
    kctx.estimate_normals(mode, K, radius)
    full = np.empty((len(pts), 3))
    kctx.download(nat.ATTR_NORMALS, full)
    try:
        for m in (1, 777, 40_000, len(pts)):
            kctx.drop(nat.ATTR_NORMALS)
            kctx.set_option(nat.OPT_QUERY_FIRST, m)
            kctx.estimate_normals(mode, K, radius)
            got = np.empty((len(pts), 3))
            kctx.download(nat.ATTR_NORMALS, got)
            assert np.array_equal(got[:m], full[:m]), m
            qi = np.unique(np.random.default_rng(m).integers(0, m, min(m, 1500)))
            idx, cnt, tier = kctx.fetch_neighbors(K, qi)
            e, ecnt = _expected_sets(orc, pts, K, mode, radius, qi)
            assert np.array_equal(cnt, ecnt) and np.array_equal(idx, e)
            assert ((tier >= 1) & (tier <= 3)).all()
    finally:
        kctx.set_option(nat.OPT_QUERY_FIRST, 0)
