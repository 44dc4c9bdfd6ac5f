"""Parity at BASELINE.json's full size (config C3: 10 M-point scan, estimate_normals(k = 30) + segment_plane(0.05, 3, 1000)).
The oracle cannot run the whole scan in seconds, so the checks are the size-independent ones:
  * normals: for 2 000 sampled points the oracle's KD-tree gives the exact 30 nearest neighbours inside the FULL 10 M
    cloud; covariance (Open3D's cumulant formula) and FastEigen3x3 of those sets must give the library's normal within
    1e-4 rad; every normal of the scan is finite and unit length;
  * k-th distance bounds left by the tier chain: never below the true k-th d2, and equal to it (FP32 rounding) where the
    query was answered;
  * segment_plane: the device Philox stream equals the CPU replay, the winning model re-scored by the oracle over all
    10 M points has exactly the returned inlier set, and the early exit stops at the same batch as a second run."""
import numpy as np
import pytest

from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like, SEED_C3

pytestmark = pytest.mark.gpu

N = 10_000_000
K = 30


@pytest.fixture(scope="module")
def scan():
    return lidar_like(N, SEED_C3)


def _covariance6(nb):
    """Open3D ComputeCovariance (cumulants / n, then E[ab] - E[a]E[b]) of one neighbour set: [c00 c01 c02 c11 c12 c22]"""
    c = np.zeros(9)
    for x, y, z in nb:                     # the reference's summation order: neighbours by increasing distance
        c += (x, y, z, x * x, x * y, x * z, y * y, y * z, z * z)
    c /= len(nb)
    return np.array([c[3] - c[0] * c[0], c[4] - c[0] * c[1], c[5] - c[0] * c[2],
                     c[6] - c[1] * c[1], c[7] - c[1] * c[2], c[8] - c[2] * c[2]])


def test_c3_normals_full_size(ctx, orc, scan):
    ctx.upload(nat.ATTR_POINTS, scan)
    ctx.drop(nat.ATTR_NORMALS)
    ctx.estimate_normals(nat.SEARCH_KNN, K, 0.0)
    nrm = ctx.download(nat.ATTR_NORMALS)
    assert np.isfinite(nrm).all()
    assert np.abs(np.einsum("ij,ij->i", nrm, nrm) - 1.0).max() < 1e-12
    bound = ctx.kth_bound()
    qi = np.random.default_rng(5).choice(N, 2000, replace=False)
    idx, d2, cnt = orc.knn(scan, K, queries=scan[qi])
    assert (cnt == K).all()
    cov = np.stack([_covariance6(scan[idx[j]]) for j in range(len(qi))])
    ref = orc.fast_eigen3x3(cov)
    ref /= np.linalg.norm(ref, axis=1, keepdims=True)
    cosang = np.abs(np.einsum("ij,ij->i", ref, nrm[qi]))
    ang = np.arccos(np.clip(cosang, -1.0, 1.0))
    assert ang.max() < 1e-4, (float(ang.max()), int((ang >= 1e-4).sum()))     # north_star tolerance: 1e-4 rad
    kth = d2[:, K - 1]
    b = bound[qi].astype(np.float64)
    assert (b >= kth * (1 - 1e-6)).all()                  # a valid upper bound everywhere ...
    assert np.median(b / np.maximum(kth, 1e-300)) < 1.001      # ... and tight: the k-th d2 plus the FP32 error margin


def test_c3_segment_plane_full_size(ctx, orc, scan):
    ctx.upload(nat.ATTR_POINTS, scan)
    plane, inl, dbg = ctx.segment_plane(0.05, 3, 1000, seed=1, debug=True)
    assert np.array_equal(ctx.fetch_samples(1000, 3), orc.sample_philox(1, N, 3, 1000))
    bi = dbg["best_iter"]
    # the winning hypothesis, rebuilt and re-scored by the oracle over the whole scan
    samples = orc.sample_philox(1, N, 3, 1000)
    model = orc.plane_model(scan, samples[bi])
    assert np.array_equal(model, dbg["best_plane"])
    count, err = orc.score_plane(scan, model, 0.05)
    assert count == dbg["counts"][bi] == len(inl)
    assert abs(err - dbg["errs"][bi]) <= 1e-10 * abs(err)
    dist = np.abs(scan @ model[:3] + model[3])
    odd = np.setxor1d(np.nonzero(dist < 0.05)[0], inl)             # numpy's dot may round the last bit differently
    assert odd.size == 0 or (np.abs(dist[odd] - 0.05) < 1e-12).all()
    # hypotheses scored before the early exit: none beats the winner under Open3D's rule (fitness, then rmse)
    scored = dbg["counts"] >= 0
    assert scored.sum() >= bi + 1 and scored.sum() < 1000          # the ground plane ends the search early
    assert dbg["counts"][scored].max() == dbg["counts"][bi]
    # refit plane: unit normal, and the inliers' centroid lies on it
    assert abs(np.linalg.norm(plane[:3]) - 1.0) < 1e-12
    assert abs(scan[inl].mean(0) @ plane[:3] + plane[3]) < 1e-9
    plane2, inl2 = ctx.segment_plane(0.05, 3, 1000, seed=1)
    assert np.array_equal(plane2, plane) and np.array_equal(inl2, inl)     # run-to-run reproducible
