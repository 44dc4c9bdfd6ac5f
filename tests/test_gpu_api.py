"""The reference-facing surface (open3dpypro.PointCloud and the processor blocks) on the GPU,
checked against the oracle.  These read like the tests the reference would have for this path."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from open3dpypro.PointCloud import PointCloud, KDTreeSearchParamHybrid, KDTreeSearchParamKNN, KDTreeSearchParamRadius
from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
from open3dpypro.processors import Processors, PointCloudMatProcessors
from open3dpypro.generator import ArrayFrameGenerator
from open3dpypro.synthetic import lidar_like


def _cos(a, b):
    return np.abs(np.einsum("ij,ij->i", a, b))


def test_config1_bunny_chain(orc, bunny):
    """BASELINE config 1: voxel_down_sample + estimate_normals(k=30) + segment_plane on bunny"""
    pc = PointCloud(bunny)
    down = pc.voxel_down_sample(0.005)
    ref = orc.voxel_trace(bunny, 0.005)
    keep = np.sort(ref["rep"])
    assert down.size() == 3017
    assert np.array_equal(down.get_points(), bunny[keep])
    down.estimate_normals()                                     # default KNN(30)
    rn = orc.estimate_normals(bunny[keep], "knn", 30)
    assert np.quantile(_cos(down.get_normals(), rn), 0.01) > 1 - 1e-8
    plane, inl = down.segment_plane(0.01, 3, 450, seed=5)
    eplane, einl = orc.segment_plane(bunny[keep], 0.01, 3, 450, orc.sample_philox(5, len(keep), 3, 450))
    assert np.array_equal(inl, einl)
    assert np.allclose(plane, eplane, atol=1e-9)
    assert isinstance(plane, list) and len(plane) == 4


def test_trace_outputs(orc, bunny):
    rng = np.random.default_rng(0)
    pc = PointCloud(bunny, rgb=rng.uniform(size=bunny.shape), normals=rng.normal(size=bunny.shape),
                    intensity=rng.uniform(size=(len(bunny), 1)))
    sub, idxmat, vec = pc.voxel_down_sample_and_trace(0.01)
    ref = orc.voxel_trace(bunny, 0.01)
    assert idxmat.dtype == np.int32 and idxmat.shape == (751, 8)
    assert np.array_equal(idxmat, ref["cubic_id"])
    assert len(vec) == 751
    for v in (0, 17, 750):
        assert np.array_equal(vec[v], ref["order"][ref["seg"][v]:ref["seg"][v + 1]])
    keep = np.sort(ref["rep"])
    assert np.array_equal(sub.get_points(), bunny[keep])
    assert np.array_equal(sub.get_colors(), pc.get_colors()[keep])
    assert np.array_equal(sub.get_normals(), pc.get_normals()[keep])
    assert np.array_equal(sub.get_intensity(), pc.get_intensity()[keep])
    # Open3D's own first return value (centroids / mean colour / normalised normal sum)
    cent, _, _ = pc.pcd.voxel_down_sample_and_trace(0.01)
    refc = orc.voxel_trace(bunny, 0.01, pc.get_normals(), pc.get_colors())
    assert np.array_equal(cent.points, refc["centroid"])
    assert np.array_equal(cent.colors, refc["color"])
    assert np.array_equal(cent.normals, refc["normal"])


def test_search_params(orc, bunny):
    sub = bunny[::5]
    for param, kw in [(KDTreeSearchParamKNN(16), dict(mode="knn", k=16)),
                      (KDTreeSearchParamHybrid(0.01, 30), dict(mode="hybrid", k=30, radius=0.01)),
                      (KDTreeSearchParamRadius(0.01), dict(mode="radius", radius=0.01))]:
        pc = PointCloud(sub)
        assert pc.estimate_normals(param=param) is pc
        assert pc.has_normals()
        rn = orc.estimate_normals(sub, **kw)
        assert np.quantile(_cos(pc.get_normals(), rn), 0.02) > 1 - 1e-8


def test_views_stay_valid_and_mutable(bunny):
    pc = PointCloud(bunny)
    view = pc.get_points()
    T = np.eye(4)
    T[0, 3] = 1.0
    assert pc.transform(T) is pc
    assert view is pc.get_points()                 # transform updated the same host buffer
    assert np.allclose(view[:, 0], bunny[:, 0] + 1.0)
    view[:, 0] -= 1.0                              # in-place edits are seen by the next call
    pc.transform(T)
    assert np.allclose(pc.get_points()[:, 0], bunny[:, 0] + 1.0)


def test_transform_with_normals(orc, bunny):
    rng = np.random.default_rng(1)
    nrm = rng.normal(size=bunny.shape)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    T = np.eye(4)
    T[:3, :3] = q
    T[:3, 3] = [0.1, 0.2, 0.3]
    pc = PointCloud(bunny, normals=nrm).transform(T)
    ep, en = orc.transform(bunny, T, nrm)
    assert np.array_equal(pc.get_points(), ep) and np.array_equal(pc.get_normals(), en)


def test_errors_match_open3d(bunny):
    pc = PointCloud(bunny)
    with pytest.raises(RuntimeError, match="voxel_size <= 0"):
        pc.voxel_down_sample(0)
    with pytest.raises(RuntimeError, match="ransac_n"):
        pc.segment_plane(0.01, 2, 10)
    with pytest.raises(RuntimeError, match="at least"):
        PointCloud(bunny[:2]).segment_plane(0.01, 3, 10)


def test_seg_planes_and_rotate_to_plane():
    rng = np.random.default_rng(2)
    a = np.c_[rng.uniform(-5, 5, (4000, 2)), rng.normal(0, 0.002, 4000)]
    b = np.c_[rng.uniform(-5, 5, 2500), rng.normal(0, 0.002, 2500) + 2.0, rng.uniform(0, 4, 2500)]
    noise = rng.uniform(-5, 5, (300, 3))
    pc = PointCloud(np.r_[a, b, noise])
    pc.pcd.seed = 11
    planes, pcds, aabbs = pc.seg_planes(thickness=0.02, num_iterations=300, minPointsRatio=0.1)
    assert len(planes) == 2 and len(pcds) == 3
    assert abs(abs(planes[0][2]) - 1.0) < 1e-3 and abs(abs(planes[1][1]) - 1.0) < 1e-3
    assert pcds[0].size() >= 4000 and pcds[1].size() >= 2400
    flat, T = pcds[0].rotate_to_plane(planes[0])
    assert np.abs(flat.get_points()[:, 2]).max() < 0.03


def test_seg_planes_equals_oracle_peel(orc):
    """device-resident peel == the oracle's segment_plane applied to the shrinking host cloud, pass by pass"""
    pts = lidar_like(40000, 9)
    pc = PointCloud(pts, rgb=np.random.default_rng(0).random((len(pts), 3)))
    pc.pcd.seed = 100
    planes, pcds, _ = pc.seg_planes(thickness=0.05, num_iterations=200, minPointsRatio=0.3)
    cur, col = pts, pc.get_colors()
    for i, plane in enumerate(planes):
        eplane, einl = orc.segment_plane(cur, 0.05, 3, 200, orc.sample_philox(100 + i, len(cur), 3, 200))
        assert np.abs(np.asarray(plane) - eplane).max() < 1e-9
        assert np.array_equal(pcds[i].get_points(), cur[einl]) and np.array_equal(pcds[i].get_colors(), col[einl])
        keep = np.ones(len(cur), bool)
        keep[einl] = False
        cur, col = cur[keep], col[keep]
    assert len(planes) >= 1 and np.array_equal(pcds[-1].get_points(), cur)


def test_resident_mode(orc):
    pts = lidar_like(30000, 5)
    pc = PointCloud(pts)
    with pc.pcd.resident():
        pc.estimate_normals()
        plane, inl = pc.segment_plane(0.05, 3, 100, seed=2)
    rn = orc.estimate_normals(pts, "knn", 30)
    assert np.quantile(_cos(pc.get_normals(), rn), 0.01) > 1 - 1e-8
    assert np.array_equal(inl, orc.segment_plane(pts, 0.05, 3, 100, orc.sample_philox(2, len(pts), 3, 100))[1])


def test_blocks_pipeline(orc):
    """VoxelDownsample -> PlaneDetection -> PlaneNormalize -> CPUNormals over generated frames,
    after a JSON round trip of the block list (reference processors.py:1045-1052)."""
    frames = np.stack([lidar_like(20000, 2000 + f, np.float32) for f in range(3)], 0)
    gen = ArrayFrameGenerator(sources=[frames], shape_types=[ShapeType.XYZ])
    det = Processors.PlaneDetection(distance_threshold=0.05, num_iterations=200, alpha=1.0, seed=4)
    pipes = [Processors.VoxelDownsample(voxel_size=0.2), det,
             Processors.PlaneNormalize(detection_uuid=det.uuid), Processors.CPUNormals()]
    pipes = PointCloudMatProcessors.loads(PointCloudMatProcessors.dumps(pipes))
    n = 0
    meta = {}
    for mats in gen:
        first = n == 0
        outs, meta = PointCloudMatProcessors.run_once(mats, meta, pipes, validate=first)
        n += 1
        out = outs[0].data()
        assert outs[0].info.shape_type == ShapeType.XYZN and out.shape[1] == 6
        plane = np.asarray(meta[pipes[1].uuid][0])
        assert abs(plane[2]) > 0.99 and plane[3] >= -1e-9          # ground plane, flipped towards the sensor
        # after normalisation the ground sits at z ~ 0 with normals ~ +-z
        ground = np.abs(out[:, 2]) < 0.05
        assert ground.mean() > 0.3
        assert np.median(np.abs(out[ground, 5])) > 0.9
    assert n == 2   # the generator consumes its first frame while sizing the mats (reference :44)
    # block results against the oracle for the last frame
    src = frames[2].astype(np.float64)
    vox = Processors.VoxelDownsample(voxel_size=0.2)
    down = vox.forward_raw([frames[2]])[0]
    ref = orc.voxel_trace(src, 0.2)
    assert np.array_equal(down, frames[2][ref["rep"]])


def test_download_async_overlaps_and_matches():
    """o3dpp_download_async: the normals travel on the copy stream while segment_plane runs; same bytes as download"""
    from open3dpypro import _native as nat
    pts = lidar_like(200_000, 14)
    c = nat.Context(0)
    c.upload(nat.ATTR_POINTS, pts)
    c.estimate_normals(nat.SEARCH_KNN, 30, 0.0)
    buf = nat.PinnedArray((len(pts), 3), np.float64)
    c.download_async(nat.ATTR_NORMALS, buf.array)
    plane, inl = c.segment_plane(0.05, 3, 100, seed=1)
    c.synchronize()
    assert np.array_equal(buf.array, c.download(nat.ATTR_NORMALS)) and len(inl) > 0
    buf.free()
    c.close()
