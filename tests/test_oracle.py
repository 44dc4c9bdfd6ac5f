"""CPU suite, part 1: the oracle (oracle/o3d_oracle.c) against the committed goldens, against
published known-answer vectors of its RNGs, and against independent implementations (scipy
cKDTree, numpy eigh, plain numpy scoring).  No GPU, no product code."""
import os

import numpy as np
import pytest
from scipy.spatial import cKDTree

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bunny_goldens.npz"))
T_FIXED = GOLD["transform_T"]


# ------------------------------------------------------------------ goldens
@pytest.mark.parametrize("vs,tag,V", [(0.002, "0p002", 15930), (0.005, "0p005", 3017), (0.01, "0p01", 751)])
def test_voxel_goldens(orc, bunny, vs, tag, V):
    r = orc.voxel_trace(bunny, vs)
    assert r["V"] == V == int(GOLD[f"vox_{tag}_V"])            # SURVEY.md §8(c) expects exactly these counts
    assert np.array_equal(np.sort(r["rep"]), GOLD[f"vox_{tag}_rep_sorted"])
    assert int((r["count"].astype(np.int64) ** 2).sum()) == int(GOLD[f"vox_{tag}_count_sum_sq"])
    assert np.allclose(r["centroid"].sum(0), GOLD[f"vox_{tag}_centroid_sum"], rtol=1e-12)


def test_voxel_against_numpy(orc, bunny):
    """independent restatement: floor((p - min) / vs) keys, largest index per key"""
    vs = 0.005
    key = np.floor((bunny - bunny.min(0)) / vs).astype(np.int64)
    order = np.lexsort((np.arange(len(bunny)), key[:, 2], key[:, 1], key[:, 0]))
    k = key[order]
    last = np.r_[np.any(k[1:] != k[:-1], axis=1), True]
    rep = order[last]
    r = orc.voxel_trace(bunny, vs)
    assert np.array_equal(rep, r["rep"])
    assert np.array_equal(k[last], r["keys"])
    cnt = np.diff(np.r_[-1, np.nonzero(last)[0]])
    assert np.array_equal(cnt, r["count"])
    # octant table: largest index per octant
    ref = (bunny - bunny.min(0)) / vs
    octant = ((ref - np.floor(ref)) >= 0.5).astype(np.int64) @ np.array([1, 2, 4])
    v = 17
    members = order[r["seg"][v]:r["seg"][v + 1]]
    exp = np.full(8, -1)
    for i in members:
        exp[octant[i]] = max(exp[octant[i]], i)
    assert np.array_equal(exp, r["cubic_id"][v])


def test_knn_goldens_and_ckdtree(orc, bunny):
    rows = GOLD["knn_rows"]
    idx, d2, cnt = orc.knn(bunny, 30, queries=bunny[rows])
    assert np.array_equal(idx, GOLD["knn_idx"]) and np.array_equal(d2, GOLD["knn_d2"])
    assert (cnt == 30).all()
    # independent: same d2 multiset as scipy (its distances are sqrt'ed doubles)
    dd, ii = cKDTree(bunny).query(bunny[rows], k=30)
    assert np.allclose(np.sqrt(d2), dd, rtol=1e-12, atol=0)
    # index sets agree wherever the 30th/31st distances do not tie
    d31 = cKDTree(bunny).query(bunny[rows], k=31)[0]
    clear = d31[:, 30] > d31[:, 29] * (1 + 1e-12)
    same = np.array([set(a) == set(b) for a, b in zip(idx, ii)])
    assert same[clear].all()
    # order inside a row is (d2, index)
    for r in range(0, len(rows), 37):
        key = list(zip(d2[r], idx[r]))
        assert key == sorted(key)


def test_knn_tree_equals_brute(orc):
    rng = np.random.default_rng(5)
    p = np.round(rng.normal(size=(3000, 3)), 2)          # coarse grid: many exact ties
    a = orc.knn(p, 30)
    b = orc.knn(p, 30, brute=True)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    h1 = orc.knn(p, 30, mode=2, radius=0.3)
    h2 = orc.knn(p, 30, mode=2, radius=0.3, brute=True)
    assert np.array_equal(h1[0], h2[0]) and np.array_equal(h1[2], h2[2])


def test_boundary_tie_count(orc, bunny):
    """SURVEY.md §3.2 measured 137 bunny points whose 30th and 31st neighbours tie exactly"""
    d31 = orc.knn(bunny, 31)[1]
    assert int((d31[:, 29] == d31[:, 30]).sum()) == 137 == int(GOLD["knn_boundary_tie_count"])


def test_normals_goldens_and_eigh(orc, bunny):
    rows = GOLD["knn_rows"]
    nrm, cov = orc.estimate_normals(bunny, "knn", 30, want_cov=True)
    assert np.allclose(nrm[rows], GOLD["normals_rows"], atol=1e-12)
    nh = orc.estimate_normals(bunny, "hybrid", 30, 0.01)
    assert np.allclose(nh[rows], GOLD["normals_hybrid_rows"], atol=1e-12)
    # independent: covariance from the golden index lists, eigenvector from numpy
    C = np.empty((len(rows), 3, 3))
    for j, r in enumerate(rows):
        q = bunny[GOLD["knn_idx"][j]]
        m = q.mean(0)
        C[j] = (q[:, :, None] * q[:, None, :]).mean(0) - np.outer(m, m)
    got = cov[rows]
    assert np.allclose(C[:, 0, 0], got[:, 0], atol=1e-15) and np.allclose(C[:, 1, 2], got[:, 4], atol=1e-15)
    w, v = np.linalg.eigh(C)
    well = (w[:, 1] - w[:, 0]) / w[:, 2] > 1e-6
    cosang = np.abs(np.einsum("ij,ij->i", v[:, :, 0], nrm[rows]))
    assert np.arccos(np.clip(cosang[well], -1, 1)).max() < 1e-6     # SURVEY measured 2.1e-8 rad on bunny


def test_fast_eigen_special_cases(orc):
    assert np.array_equal(orc.fast_eigen3x3([0, 0, 0, 0, 0, 0])[0], [0, 0, 0])          # max coeff 0 -> zero vector
    assert np.array_equal(orc.fast_eigen3x3([3, 0, 0, 2, 0, 1])[0], [0, 0, 1])          # diagonal: smallest axis
    assert np.array_equal(orc.fast_eigen3x3([1, 0, 0, 2, 0, 3])[0], [1, 0, 0])
    assert np.array_equal(orc.fast_eigen3x3([1, 0, 0, 1, 0, 1])[0], [0, 0, 1])          # identity -> +z


# ------------------------------------------------------------------ RNG known answers
def test_philox_known_answer(orc):
    """Random123 kat_vectors: philox4x32-10"""
    assert [hex(v) for v in orc.philox(0, 0, 0, 0, 0)] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    k = (0xffffffff << 32) | 0xffffffff
    assert [hex(v) for v in orc.philox(k, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff)] == \
        ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    k = (0xa4093822 << 32) | 0x299f31d0 if False else (0x299f31d0 << 32) | 0xa4093822
    assert [hex(v) for v in orc.philox(k, 0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344)] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_mt19937_known_answer(orc):
    """ISO C++ [rand.predef]: the 10000th invocation of a default-constructed mt19937 is 4123659995"""
    assert orc.mt19937_nth(5489, 9999) == 4123659995      # nth is 0-based
    s = orc.sample_mt19937(42, 35947, 3, 8)
    assert np.array_equal(s, GOLD["mt19937_samples_head"])
    assert (np.sort(s, 1)[:, 1:] != np.sort(s, 1)[:, :-1]).all()      # distinct inside a sample


# ------------------------------------------------------------------ RANSAC
def test_segment_goldens_and_numpy(orc, bunny):
    samples = orc.sample_philox(5, len(bunny), 3, 450)
    assert np.array_equal(samples[:16], GOLD["seg_samples_head"])
    plane, inl, dbg = orc.segment_plane(bunny, 0.01, 3, 450, samples, debug=True)
    assert np.array_equal(inl, GOLD["seg_inliers"]) and dbg["best_iter"] == int(GOLD["seg_best_iter"])
    assert np.array_equal(dbg["counts"], GOLD["seg_counts"])
    assert np.allclose(plane, GOLD["seg_plane"], atol=1e-13)
    # independent scoring of every hypothesis with numpy
    p0, p1, p2 = bunny[samples[:, 0]], bunny[samples[:, 1]], bunny[samples[:, 2]]
    n = np.cross(p1 - p0, p2 - p0)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    d = -np.einsum("ij,ij->i", n, p0)
    cnt = np.array([(np.abs(bunny @ n[h] + d[h]) < 0.01).sum() for h in range(0, 450, 15)])
    assert np.array_equal(cnt, dbg["counts"][::15])
    # Open3D's serial rule: first strictly better fitness wins (ties by rmse)
    assert dbg["counts"][dbg["best_iter"]] == dbg["counts"].max()
    # inliers are those of the pre-refit winner, ascending
    bp = dbg["best_plane"]
    assert np.array_equal(inl, np.nonzero(np.abs(bunny @ bp[:3] + bp[3]) < 0.01)[0])
    # refit (Open3D GetPlaneFromPoints = regression along the dominant axis, not total least squares):
    # close to the SVD normal, through the centroid exactly
    c = bunny[inl].mean(0)
    vt = np.linalg.svd(bunny[inl] - c)[2][2]
    assert abs(vt @ plane[:3]) > 0.999 and abs(plane[:3] @ c + plane[3]) < 1e-12


def test_segment_early_exit_rule(orc):
    """a dominant plane: break_iteration = log(1-p)/log(1-fitness^3) cuts the sequence short"""
    rng = np.random.default_rng(1)
    p = np.concatenate([np.c_[rng.uniform(-1, 1, (9000, 2)), rng.normal(0, 1e-3, 9000)], rng.uniform(-1, 1, (1000, 3))])
    samples = orc.sample_philox(3, len(p), 3, 1000)
    _, inl, dbg = orc.segment_plane(p, 0.01, 3, 1000, samples, debug=True)
    scored = (dbg["counts"] >= 0).sum()
    assert 0 < scored < 200 and (dbg["counts"][scored + 5:] == -1).all()
    assert len(inl) >= 9000


def test_segment_errors(orc, bunny):
    with pytest.raises(RuntimeError, match="ransac_n"):
        orc.segment_plane(bunny, 0.01, 2, 10, np.zeros((10, 2), np.int64))
    with pytest.raises(RuntimeError, match="at least"):
        orc.segment_plane(bunny[:2], 0.01, 3, 10, np.zeros((10, 3), np.int64))


# ------------------------------------------------------------------ transform
def test_transform_golden_and_numpy(orc, bunny):
    tp, _ = orc.transform(bunny[:64], T_FIXED)
    assert np.array_equal(tp, GOLD["transform_head"])
    h = np.c_[bunny[:64], np.ones(64)] @ T_FIXED.T
    assert np.allclose(tp, h[:, :3] / h[:, 3:], rtol=1e-13, atol=1e-15)
    P = T_FIXED.copy()
    P[3] = [0.1, 0.0, 0.2, 2.0]                      # projective row: divide by w
    tp2, tn2 = orc.transform(bunny[:64], P, normals=bunny[:64])
    h = np.c_[bunny[:64], np.ones(64)] @ P.T
    assert np.allclose(tp2, h[:, :3] / h[:, 3:], rtol=1e-14)
    assert np.allclose(tn2, bunny[:64] @ P[:3, :3].T, rtol=1e-14)   # normals: no translation, no renormalise
