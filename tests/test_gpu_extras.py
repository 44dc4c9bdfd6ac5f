"""Rows f2 / f3 of SURVEY.md §8: the KNN consumers and the z-depth splat, against the oracle's KNN and numpy."""
import numpy as np
import pytest


def _stat_outlier_oracle(orc, pts, k, ratio):
    """Open3D RemoveStatisticalOutliers restated on the oracle's exact (d2, index)-ordered KNN lists"""
    idx, d2, cnt = orc.knn(pts, k)
    avg = np.full(len(pts), -1.0)
    for i in range(len(pts)):
        s = 0.0
        for j in range(cnt[i]):
            s += float(np.sqrt(d2[i, j]))
        if cnt[i]:
            avg[i] = s / cnt[i]
    valid = avg > 0
    nv = int((avg >= 0).sum())
    mean = np.where(valid, avg, 0).sum() / nv
    std = np.sqrt(np.where(valid, (avg - mean) ** 2, 0).sum() / (nv - 1))
    thr = mean + ratio * std
    return avg, thr, np.nonzero(valid & (avg < thr))[0]


@pytest.mark.gpu
def test_knn_mean_distance_and_statistical_outlier(orc, ctx, bunny):
    from open3dpypro import _native as nat
    from open3dpypro.PointCloud import PointCloud
    pts = bunny[::3].copy()
    pts[::97] += 0.02                                    # a few outliers
    avg, thr, keep = _stat_outlier_oracle(orc, pts, 20, 2.0)
    ctx.upload(nat.ATTR_POINTS, pts)
    got = ctx.knn_mean_distance(20)
    assert np.array_equal(got, avg)                      # same d2, same summation order: bit-exact
    pc, idx = PointCloud(pts).remove_statistical_outlier(20, 2.0)
    border = np.abs(avg - thr) < 1e-12 * thr
    assert np.array_equal(np.setdiff1d(idx, np.nonzero(border)[0]), np.setdiff1d(keep, np.nonzero(border)[0]))
    assert pc.size() == len(idx) and np.array_equal(pc.get_points(), pts[idx])
    assert 0 < len(idx) < len(pts)


@pytest.mark.gpu
def test_zsplat_matches_numpy_last_write_wins(ctx):
    from open3dpypro import _native as nat
    from open3dpypro.synthetic import lidar_like
    pts32 = lidar_like(300_000, 8).astype(np.float32)
    g = 128
    x, y, z = pts32[:, 0], pts32[:, 1], pts32[:, 2]
    x_min, x_max, y_min, y_max = x.min(), x.max(), y.min(), y.max()
    xi = np.clip(((x - x_min) / (x_max - x_min) * (g - 1)).astype(np.int32), 0, g - 1)
    yi = np.clip(((y - y_min) / (y_max - y_min) * (g - 1)).astype(np.int32), 0, g - 1)
    ref = np.full((g, g), np.nan, dtype=np.float32)
    ref[yi, xi] = z                                      # numpy keeps the last point of a cell
    last_ref = np.full((g, g), -1, np.int32)
    last_ref[yi, xi] = np.arange(len(z), dtype=np.int32)
    ctx.upload(nat.ATTR_POINTS, pts32)
    zg, last = ctx.zsplat(g, float(x_min), float(x_max), float(y_min), float(y_max))
    assert np.array_equal(last, last_ref)
    assert np.array_equal(np.isnan(zg), np.isnan(ref)) and np.array_equal(zg[~np.isnan(zg)], ref[~np.isnan(ref)])


@pytest.mark.gpu
def test_zdepth_viewer_and_outlier_blocks():
    from open3dpypro.processors import Processors
    from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
    from open3dpypro.synthetic import lidar_like
    pts = lidar_like(50_000, 3).astype(np.float32)
    v = Processors.ZDepthViewer(grid_size=64, z_min=-1.0, z_max=10.0)
    out = v.forward_raw([pts])
    assert out[0] is pts and v._img_data[0].shape == (64, 64, 3) and v._img_data[0].dtype == np.uint8
    assert v._img_data[0][:, :, 1].max() > 0
    r = Processors.RemoveStatisticalOutlier(nb_neighbors=10, std_ratio=1.0)
    kept = r.forward_raw([pts])[0]
    assert 0 < len(kept) < len(pts) and kept.shape[1] == 3
    sel = Processors.RadiusSelection(radius=20.0).forward_raw([pts])[0]
    assert np.all(np.linalg.norm(sel, axis=1) <= 20.0) and len(sel) == int((np.linalg.norm(pts[:, :3].astype(np.float64), axis=1) <= 20.0).sum())


def test_single_query_kdtree_matches_brute_force():
    from open3dpypro.PointCloud import PointCloud
    rng = np.random.default_rng(0)
    pts = rng.random((2000, 3))
    pc = PointCloud(pts)
    d2 = ((pts - pts[17]) ** 2).sum(1)
    k, idx, dd = pc.get_points_by_knn(17, 25)
    assert k == 25 and idx[0] == 17 and np.array_equal(np.sort(d2)[:25], np.sort(dd))
    k, idx, dd = pc.get_points_radius(17, 0.2)
    assert k == int((d2 < 0.04).sum()) and np.all(np.diff(dd) >= 0)
    assert PointCloud(pts).select_by_radius(0.5).size() == int((np.linalg.norm(pts, axis=1) <= 0.5).sum())
    img = PointCloud(pts).to_2D_Img(0.05)[0]
    assert img.ndim == 2 and img.max() == 255
