"""Block-level parity: the device blocks against the oracle plus the reference's own
host-side lines, restated here in numpy next to their citation.

  CPUNormals      reference processors.py:243-249  np.hstack([pcd, PointCloudBase(pcd[:, :3]).estimate_normals().get_normals()])
  PlaneDetection  reference processors.py:633-651 (segment_plane + flip towards the sensor), :688-699 (optional pre-voxel,
                  EMA best = best (1 - alpha) + plane alpha, meta[uuid])
  PlaneNormalize  reference processors.py:709-759 (Rodrigues rotation and rigid transform in the MAT's dtype, float32)
and the frame residency of the device blocks (one upload per frame per run_once), CUDA-tensor mats, the bounded voxel
call (PointCloud.py:339-340 passes min_bound / max_bound) and two clouds interleaved inside resident()."""
import numpy as np
import pytest

from open3dpypro import _native as nat
from open3dpypro.PointCloud import PointCloud
from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
from open3dpypro.processors import Processors, PointCloudMatProcessors
from open3dpypro.synthetic import lidar_like

pytestmark = pytest.mark.gpu


def _angle(a, b):
    c = np.abs(np.einsum("ij,ij->i", a, b)) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))
    return np.arccos(np.clip(c, -1.0, 1.0))


def _well(cov):
    Cm = np.empty((len(cov), 3, 3))
    Cm[:, 0, 0], Cm[:, 0, 1], Cm[:, 0, 2] = cov[:, 0], cov[:, 1], cov[:, 2]
    Cm[:, 1, 0], Cm[:, 1, 1], Cm[:, 1, 2] = cov[:, 1], cov[:, 3], cov[:, 4]
    Cm[:, 2, 0], Cm[:, 2, 1], Cm[:, 2, 2] = cov[:, 2], cov[:, 4], cov[:, 5]
    lam = np.linalg.eigvalsh(Cm)
    return (lam[:, 1] - lam[:, 0]) / np.maximum(np.abs(lam).max(1), 1e-300) > 1e-6


# ------------------------------------------------------------------ a8 CPUNormals
@pytest.mark.parametrize("extra", [0, 1])
def test_cpunormals_block_equals_oracle(orc, extra):
    frame = lidar_like(40_000, 2001, np.float32)
    if extra:
        frame = np.hstack([frame, np.random.default_rng(0).uniform(size=(len(frame), 1)).astype(np.float32)])   # XYZi
    blk = Processors.CPUNormals()
    mat = PointCloudMat(shape_type=ShapeType.XYZi if extra else ShapeType.XYZ).build(frame)
    outs, _ = blk.validate([mat], {})
    out = outs[0].data()
    assert out.dtype == np.float64 and out.shape == (len(frame), frame.shape[1] + 3)        # np.hstack([pcd, ns]) promotes to float64
    assert np.array_equal(out[:, :frame.shape[1]], frame.astype(np.float64))
    ref, cov = orc.estimate_normals(frame[:, :3].astype(np.float64), "knn", 30, want_cov=True)
    well = _well(cov)
    assert well.mean() > 0.99
    assert _angle(out[:, -3:], ref)[well].max() < 1e-4
    assert np.abs(np.linalg.norm(out[:, -3:], axis=1) - 1).max() < 1e-9
    # hybrid variant of the block (radius > 0 -> KDTreeSearchParamHybrid(radius, knn))
    hyb = Processors.CPUNormals(radius=0.5).forward_raw([frame])[0]
    refh, covh = orc.estimate_normals(frame[:, :3].astype(np.float64), "hybrid", 30, radius=0.5, want_cov=True)
    assert _angle(hyb[:, -3:], refh)[_well(covh)].max() < 1e-4


# ------------------------------------------------------------------ a10 PlaneDetection
def _ref_plane(orc, xyz, thr, H, seed):
    """reference processors.py:637-650: segment_plane(ransac_n=3) then flip so that the normal looks at the sensor origin"""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    plane, _ = orc.segment_plane(xyz, thr, 3, H, orc.sample_philox(seed, len(xyz), 3, H))
    plane = np.asarray(plane)
    n, d = plane[:3], plane[3]
    if np.dot(n, np.zeros(3) - (-d * n)) < 0:
        plane = -plane
    return plane


@pytest.mark.parametrize("variant", ["plain", "zband", "prevoxel"])
def test_plane_detection_block_ema_flip_zband_voxel(orc, variant):
    frames = [lidar_like(30_000, 2100 + f, np.float32) for f in range(3)]
    for k, f in enumerate(frames):
        f[:, 2] += 1.5 if k == 1 else -1.5   # sensor above the ground (d > 0), once below it (d < 0: the flip branch)
    alpha, thr, H, seed = 0.3, 0.05, 200, 11
    kw = dict(distance_threshold=thr, num_iterations=H, alpha=alpha, seed=seed)
    if variant == "zband":
        kw.update(z_lower=-2.5, z_upper=2.5)
    if variant == "prevoxel":
        kw.update(voxel_size=0.2)
    det = Processors.PlaneDetection(**kw)
    best = np.zeros(4)
    meta = {}
    mat = PointCloudMat(shape_type=ShapeType.XYZ).build(frames[0])
    flipped = 0
    for call, frame in enumerate(frames):
        mat.unsafe_update_mat(frame)
        outs, meta = PointCloudMatProcessors.run_once([mat], meta, [det], validate=(call == 0))
        xyz = frame.astype(np.float64)
        if variant == "zband":
            xyz = xyz[(-2.5 < xyz[:, 2]) & (xyz[:, 2] < 2.5)]
        if variant == "prevoxel":
            rep = orc.voxel_trace(xyz, 0.2)["rep"]
            xyz = xyz[rep]                                       # pcd[idxmat.max(1)], rows in ascending voxel key
            assert np.array_equal(outs[0].data(), frame[rep])    # the block hands the down-sampled mat on (reference :691-692)
        plane = _ref_plane(orc, xyz, thr, H, seed + call)
        raw, _ = orc.segment_plane(xyz, thr, 3, H, orc.sample_philox(seed + call, len(xyz), 3, H))
        flipped += int(np.dot(plane, raw) < 0)
        best = best * (1.0 - alpha) + plane * alpha              # reference :697
        assert np.allclose(meta[det.uuid][0], best, rtol=0, atol=1e-9), (call, meta[det.uuid][0], best)
        assert plane[3] >= 0
    assert 0.0 < np.linalg.norm(best[:3]) < 1.0                  # alpha != 1: three EMA steps from zero
    assert flipped == 1                                          # the flip branch was exercised (frame 1)


# ------------------------------------------------------------------ a11 PlaneNormalize
def _reference_rotate_to_plane(pcd, plane):
    """reference processors.py:709-744 with numpy as the mat backend, in the dtype of the mat"""
    dt = pcd.dtype
    a, b, c, d = plane
    normal = np.array([a, b, c], dtype=dt)
    z_axis = np.array([0, 0, 1], dtype=dt)
    va = normal / np.linalg.norm(normal)
    vb = z_axis / np.linalg.norm(z_axis)
    v = np.cross(va, vb)
    if np.linalg.norm(v) < 1e-6:
        R = np.eye(3, dtype=dt)
    else:
        cc = np.dot(va, vb)
        s = np.linalg.norm(v)
        kmat = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=dt)
        R = np.eye(3, dtype=dt) + kmat + np.matmul(kmat, kmat) * ((1 - cc) / (s ** 2))
    point_on_plane = -d * normal / np.dot(normal, normal)
    t = -np.matmul(R, point_on_plane)
    T = np.eye(4, dtype=dt)
    T[:3, :3] = R
    T[:3, 3] = t
    homo = np.hstack([pcd[:, :3], np.ones((len(pcd), 1), dtype=dt)])
    return np.matmul(T, homo.T).T[:, :3], T


@pytest.mark.parametrize("dtype,cols", [(np.float32, 3), (np.float32, 6), (np.float64, 4)])
def test_plane_normalize_block_equals_reference_formula(dtype, cols):
    rng = np.random.default_rng(3)
    frame = lidar_like(50_000, 2200, np.float32).astype(dtype)
    if cols > 3:
        frame = np.hstack([frame, rng.uniform(size=(len(frame), cols - 3)).astype(dtype)])
    plane = [0.02, -0.03, 0.999, 1.4]
    blk = Processors.PlaneNormalize(detection_uuid="det")
    out = blk.forward_raw([frame], meta={"det": [plane]})[0]
    exp_xyz, T = _reference_rotate_to_plane(frame, plane)
    assert out.dtype == frame.dtype and out.shape == frame.shape
    assert np.array_equal(np.asarray(blk.forward_T[0], dtype=dtype), T)         # same formula, same dtype, same bits
    scale = np.abs(exp_xyz).max()
    assert np.abs(out[:, :3] - exp_xyz).max() <= 1e-5 * scale                    # bar: 1e-5 relative (BLAS sums in another order)
    if dtype == np.float32:                                                      # and in fact within float32 rounding of the exact product
        exact = (T.astype(np.float64) @ np.hstack([frame[:, :3].astype(np.float64), np.ones((len(frame), 1))]).T).T[:, :3]
        assert np.abs(out[:, :3] - exact.astype(np.float32)).max() <= 2e-7 * scale
    assert np.array_equal(out[:, 3:], frame[:, 3:])
    # a plane already facing +z: identity rotation branch (reference :713-714)
    out2 = blk.forward_raw([frame], meta={"det": [[0.0, 0.0, 1.0, 0.5]]})[0]
    assert np.abs(out2[:, 2] - (frame[:, 2] + dtype(0.5))).max() <= 1e-6 * max(1.0, np.abs(frame[:, 2]).max())


# ------------------------------------------------------------------ residency across the blocks of one run_once
def test_one_upload_per_frame_through_three_blocks(orc, monkeypatch):
    frame = lidar_like(30_000, 2300, np.float32)
    det = Processors.PlaneDetection(distance_threshold=0.05, num_iterations=100, alpha=1.0, seed=5)
    pipes = [Processors.CPUNormals(), det, Processors.PlaneNormalize(detection_uuid=det.uuid)]
    mat = PointCloudMat(shape_type=ShapeType.XYZ).build(frame)
    PointCloudMatProcessors.run_once([mat], {}, pipes, validate=True)
    uploads = []
    real = nat.Context.upload
    monkeypatch.setattr(nat.Context, "upload", lambda self, attr, arr: (uploads.append((attr, arr.dtype, arr.shape)), real(self, attr, arr))[1])
    outs, meta = PointCloudMatProcessors.run_once([mat], {}, pipes)
    assert uploads == [(nat.ATTR_POINTS, np.dtype(np.float32), (len(frame), 3))]         # once, as float32
    # the chained result equals the blocks run one by one from host arrays (fresh meta: every block uploads)
    monkeypatch.undo()
    a = Processors.CPUNormals().forward_raw([frame], meta={})[0]
    plane = _ref_plane(orc, frame.astype(np.float64), 0.05, 100, 5 + 1)                  # second call of the block: seed + 1
    assert np.allclose(meta[det.uuid][0], plane, atol=1e-9)
    b = Processors.PlaneNormalize(detection_uuid="d").forward_raw([a], meta={"d": meta[det.uuid]})[0]
    assert np.array_equal(outs[0].data(), b)


def test_several_mats_per_call_side_by_side(monkeypatch):
    """run_once over a LIST of mats (one per sensor in the reference): every mat has its own library context and host thread;
    the results equal the mats run one at a time, each mat is uploaded once per call, and a second call reuses the contexts"""
    frames = [lidar_like(30_000 + 1000 * k, 2500 + k, np.float32) for k in range(3)]
    def pipes_():
        det = Processors.PlaneDetection(distance_threshold=0.05, num_iterations=100, alpha=1.0, seed=5)
        return [Processors.CPUNormals(), det, Processors.PlaneNormalize(detection_uuid=det.uuid)], det
    single = []
    for f in frames:
        pipes, det = pipes_()
        mat = PointCloudMat(shape_type=ShapeType.XYZ).build(f)
        outs, meta = PointCloudMatProcessors.run_once([mat], {}, pipes, validate=True)
        single.append((outs[0].data().copy(), list(meta[det.uuid][0])))
    pipes, det = pipes_()
    mats = [PointCloudMat(shape_type=ShapeType.XYZ).build(f) for f in frames]
    uploads = []
    real = nat.Context.upload
    monkeypatch.setattr(nat.Context, "upload", lambda self, attr, arr: (uploads.append((id(self), arr.shape)), real(self, attr, arr))[1])
    outs, meta = PointCloudMatProcessors.run_once(mats, {}, pipes, validate=True)
    assert sorted(sh for _, sh in uploads) == sorted((len(f), 3) for f in frames)      # one upload per mat
    assert len({c for c, _ in uploads}) == len(frames)                                  # a context per mat
    for k in range(len(frames)):
        assert np.array_equal(outs[k].data(), single[k][0])
        assert meta[det.uuid][k] == single[k][1]
    monkeypatch.undo()
    seq = [Processors.CPUNormals(parallel_mats=False).forward_raw([f], meta={})[0] for f in frames]
    par = Processors.CPUNormals().forward_raw(frames, meta={})
    for a, b in zip(seq, par):
        assert np.array_equal(a, b)


def test_cuda_tensor_mats_stay_on_the_device():
    torch = pytest.importorskip("torch")
    frame = lidar_like(30_000, 2400, np.float32)
    t = torch.from_numpy(frame).cuda()
    host = Processors.CPUNormals().forward_raw([frame], meta={})[0]
    dev = Processors.CPUNormals().forward_raw([t], meta={})[0]
    assert dev.is_cuda and dev.dtype == torch.float64 and np.array_equal(dev.cpu().numpy(), host)
    plane = [0.02, -0.03, 0.999, 1.4]
    h2 = Processors.PlaneNormalize(detection_uuid="d").forward_raw([frame], meta={"d": [plane]})[0]
    d2 = Processors.PlaneNormalize(detection_uuid="d").forward_raw([t], meta={"d": [plane]})[0]
    assert d2.is_cuda and d2.dtype == torch.float32 and np.array_equal(d2.cpu().numpy(), h2)
    v1 = Processors.VoxelDownsample(voxel_size=0.3).forward_raw([frame])[0]
    v2 = Processors.VoxelDownsample(voxel_size=0.3).forward_raw([t])[0]
    assert v2.is_cuda and np.array_equal(v2.cpu().numpy(), v1)
    mat = PointCloudMat(shape_type=ShapeType.XYZ).build(t)                               # tensor mats pass validation of the device blocks
    outs, _ = Processors.CPUNormals().validate([mat], {})
    assert outs[0].data().is_cuda


# ------------------------------------------------------------------ ADVICE: two clouds interleaved inside resident()
def test_resident_cloud_survives_work_on_another_cloud(orc):
    a, b = lidar_like(20_000, 31), lidar_like(15_000, 32)
    pa, pb = PointCloud(a), PointCloud(b)
    with pa.pcd.resident():
        pa.estimate_normals()
        pb.estimate_normals()                                    # another cloud, same device, in between
        plane_b, inl_b = pb.segment_plane(0.05, 3, 50, seed=3)
        plane_a, inl_a = pa.segment_plane(0.05, 3, 50, seed=2)
        T = np.eye(4); T[:3, 3] = [1.0, -2.0, 0.5]
        pa.transform(T)
    assert np.array_equal(inl_a, orc.segment_plane(a, 0.05, 3, 50, orc.sample_philox(2, len(a), 3, 50))[1])
    assert np.array_equal(inl_b, orc.segment_plane(b, 0.05, 3, 50, orc.sample_philox(3, len(b), 3, 50))[1])
    assert np.array_equal(pa.get_points(), orc.transform(a, T)[0])
    na = orc.estimate_normals(a, "knn", 30)
    assert np.quantile(np.abs(np.einsum("ij,ij->i", pa.get_normals(), na)), 0.01) > 1 - 1e-8
    assert np.array_equal(pb.get_points(), b)


# ------------------------------------------------------------------ bounded voxel (PointCloud.py:339-340 passes the bounds)
def test_voxel_bounded_origin(ctx, orc, bunny):
    rng = np.random.default_rng(8)
    cases = [(bunny, 0.005, bunny.min(0) - [0.0123, 0.0, 0.0071], bunny.max(0) + 0.01),       # origin below the cloud
             (bunny, 0.005, bunny.min(0) + [0.02, 0.01, 0.0], None),                           # origin inside: negative indices
             (lidar_like(200_000, 9), 0.1, np.array([-150.0, -150.0, -20.0]), np.array([150.0, 150.0, 30.0])),
             (rng.uniform(-1, 1, size=(5000, 3)), 1e-7, np.array([-3.0, -3.0, -3.0]), np.array([3.0, 3.0, 3.0]))]   # wide keys
    for pts, vs, mn, mx in cases:
        ctx.upload(nat.ATTR_POINTS, pts)
        V = ctx.voxel(vs, mn, mx)
        ref = orc.voxel_trace(pts, vs, min_bound=mn, max_bound=mx)
        got = ctx.voxel_fetch(V, len(pts), ["rep", "count", "cubic_id", "order", "seg", "centroid"])
        assert V == ref["V"]
        for k in ("rep", "count", "cubic_id", "order", "seg", "centroid"):
            assert np.array_equal(got[k], ref[k]), k
    # a shifted origin really changes the grid
    assert orc.voxel_trace(bunny, 0.005)["V"] != orc.voxel_trace(bunny, 0.005, min_bound=bunny.min(0) - 0.0023)["V"]
    with pytest.raises(nat.O3dppError, match="too small"):
        ctx.voxel(1e-12, bunny.min(0), bunny.max(0))
    # the Open3D-level call forwards the bounds (ADVICE: they used to be ignored)
    pc = PointCloud(bunny)
    mn = bunny.min(0) - 0.0023
    _, idxmat, _ = pc.pcd.voxel_down_sample_and_trace(0.005, min_bound=mn, max_bound=bunny.max(0))
    assert np.array_equal(idxmat, orc.voxel_trace(bunny, 0.005, min_bound=mn, max_bound=bunny.max(0))["cubic_id"])


def test_upload_from_page_locked_shared_memory_segment():
    """a named shared-memory frame segment (reference shmIO.py:281-296) page-locked with o3dpp_host_register"""
    from open3dpypro.shmIO import NumpyFloat32SharedMemoryStreamIO as IO
    frame = lidar_like(20_000, 41, np.float32)
    w = IO.writer("o3dpp_test:pinned", frame.shape)
    r = IO.reader("o3dpp_test:pinned", frame.shape, pin=True)
    try:
        w.write(frame)
        view = r.read(copy=False)
        c = nat.Context(0)
        c.upload(nat.ATTR_POINTS, view)
        assert np.array_equal(c.download(nat.ATTR_POINTS), frame.astype(np.float64))
        c.close()
    finally:
        r.close()
        w.close()
