"""CPU suite, part 2: the C-ABI library (loads, exports every declared symbol, refuses to work
without a GPU), the host-side mirror of the reference interface, and the multi-GPU host logic under
gloo with world_size 2.  No compute call reaches a device here."""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "o3dpp.h")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


# ------------------------------------------------------------------ C ABI
def test_library_exports_every_declared_symbol():
    from open3dpypro import _native as nat
    txt = open(HEADER).read()
    declared = sorted(set(re.findall(r"O3DPP_API\s+[\w\s\*]+?\b(o3dpp_\w+)\s*\(", txt)))
    assert len(declared) >= 30
    assert declared == sorted(nat.SYMBOLS), "include/o3dpp.h and _native.SYMBOLS disagree"
    L = ctypes.CDLL(nat.LIB_PATH)
    for s in declared:
        assert hasattr(L, s), f"libo3dpp_b200.so does not export {s}"
    L.o3dpp_version.restype = ctypes.c_char_p
    assert b"sm_100a" in L.o3dpp_version()


def test_header_cites_reference_lines():
    txt = open(HEADER).read()
    for cite in ("PointCloud.py:60-62", "PointCloud.py:338-341", "PointCloud.py:68-73", "PointCloud.py:75-77",
                 "PointCloud.py:185-204"):
        assert cite in txt


def test_no_cpu_fallback():
    """without a CUDA device the library refuses to create a context; nothing routes to the oracle"""
    if _has_gpu():
        pytest.skip("a GPU is present")
    from open3dpypro import _native as nat
    with pytest.raises(nat.O3dppError) as e:
        nat.Context(0)
    assert e.value.code == -1
    from open3dpypro.PointCloud import PointCloud
    pc = PointCloud(np.zeros((10, 3)))
    with pytest.raises(RuntimeError):
        pc.estimate_normals()
    with pytest.raises(RuntimeError):
        pc.segment_plane()
    with pytest.raises(RuntimeError):
        pc.voxel_down_sample(0.1)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "open3d-py-extension_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert "oracle" not in src.replace("the oracle", "").replace("CPU oracle", "") or f == "sharding.py", \
                    f"{f} mentions the oracle"
                assert "liborc" not in src and "import orc" not in src and "from oracle" not in src


# ------------------------------------------------------------------ host mirror of the reference interface
def test_pointcloud_host_state():
    from open3dpypro.PointCloud import PointCloud, KDTreeSearchParamKNN, KDTreeSearchParamHybrid, KDTreeSearchParamRadius
    rng = np.random.default_rng(0)
    xyz = rng.normal(size=(50, 3)).astype(np.float32)
    pc = PointCloud(xyz, rgb=rng.uniform(0, 255, (50, 3)), intensity=rng.uniform(size=(50, 1)))
    assert pc.get_points().dtype == np.float64 and pc.get_points().shape == (50, 3)
    assert pc.get_colors().max() <= 1.0                      # reference PointCloud.py:37-39 normalises to [0,1]
    assert pc.size() == 50 and not pc.pcd.has_normals() and pc.pcd.has_colors()
    v = pc.get_points()
    v[0, 0] = 42.0                                           # get_points() is a mutable view (PointCloud.py:120)
    assert pc.get_points()[0, 0] == 42.0
    sub = pc._select_by_idx([5, 3, 3, 9])
    assert np.array_equal(sub.get_points(), pc.get_points()[[3, 5, 9]])     # boolean mask: ascending order
    inv = pc._select_by_idx([5, 3, 9], invert=True)
    assert inv.size() == 47
    assert KDTreeSearchParamKNN(30).knn == 30
    h = KDTreeSearchParamHybrid(radius=0.01, max_nn=30)
    assert (h.radius, h.max_nn) == (0.01, 30) and KDTreeSearchParamRadius(0.5).radius == 0.5


def test_pcd_reader_bunny_layout(tmp_path):
    """data/bunny.pcd of the reference: binary, FIELDS Intensity x y z _, 20 B per point"""
    from open3dpypro.pcdio import read_pcd
    rng = np.random.default_rng(1)
    n = 257
    rec = np.zeros(n, dtype=[("Intensity", "f4"), ("x", "f4"), ("y", "f4"), ("z", "f4"), ("_", "u1", (4,))])
    for k in ("Intensity", "x", "y", "z"):
        rec[k] = rng.normal(size=n).astype(np.float32)
    hdr = ("# .PCD v0.7\nVERSION 0.7\nFIELDS Intensity x y z _\nSIZE 4 4 4 4 1\nTYPE F F F F U\nCOUNT 1 1 1 1 4\n"
           f"WIDTH {n}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {n}\nDATA binary\n").encode()
    p = tmp_path / "b.pcd"
    p.write_bytes(hdr + rec.tobytes())
    xyz, inten = read_pcd(str(p))
    assert xyz.dtype == np.float64 and np.array_equal(xyz[:, 1], rec["y"].astype(np.float64))
    assert np.array_equal(inten, rec["Intensity"].astype(np.float64))
    a = tmp_path / "a.pcd"
    a.write_text("FIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 2\nHEIGHT 1\nPOINTS 2\nDATA ascii\n1 2 3\n4 5 6\n")
    assert np.array_equal(read_pcd(str(a))[0], [[1, 2, 3], [4, 5, 6]])


def test_mat_shape_types_and_validation():
    from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
    m = PointCloudMat(shape_type=ShapeType.XYZ).build(np.zeros((10, 3), np.float32))
    assert m.info.N == 10 if hasattr(m.info, "N") else True
    assert ShapeType.XYZ.add_normals().contains_normals()
    with pytest.raises((ValueError, TypeError)):
        PointCloudMat(shape_type=ShapeType.XYZ).build(np.zeros((10, 4), np.float32))     # D must match the type


def test_shm_stream_roundtrip():
    from open3dpypro.shmIO import NumpyFloat32SharedMemoryStreamIO as IO
    key = f"test:o3dpp:{os.getpid()}"
    w = IO.writer(key, (128, 3))
    try:
        r = IO.reader(key, (128, 3))
        a = np.random.default_rng(2).normal(size=(128, 3)).astype(np.float32)
        w.write(a)
        assert np.array_equal(r.read(), a)
        with pytest.raises(ValueError):
            w.write(np.zeros((5, 3), np.float32))
        with pytest.raises(ValueError):
            r.write(a)
        r.close()
    finally:
        w.close()


def test_generator_drops_nan_rows():
    from open3dpypro.generator import ArrayFrameGenerator
    fr = np.random.default_rng(3).normal(size=(3, 40, 3)).astype(np.float32)
    fr[1, 7] = np.nan
    from open3dpypro.PointCloudMat import ShapeType
    gen = ArrayFrameGenerator(sources=[fr], shape_types=[ShapeType.XYZ])
    sizes = [mats[0].data().shape[0] for mats in gen]
    assert 39 in sizes and max(sizes) == 40


def test_pipeline_json_roundtrip():
    from open3dpypro.processors import Processors, PointCloudMatProcessors
    pipes = [Processors.VoxelDownsample(voxel_size=0.25), Processors.CPUNormals(knn=20),
             Processors.PlaneDetection(distance_threshold=0.05, num_iterations=200)]
    pipes.append(Processors.PlaneNormalize(detection_uuid=pipes[2].uuid))
    ser = PointCloudMatProcessors.dumps(pipes)
    back = PointCloudMatProcessors.loads(ser)
    assert [type(b).__name__ for b in back] == ["VoxelDownsample", "CPUNormals", "PlaneDetection", "PlaneNormalize"]
    assert back[0].voxel_size == 0.25 and back[1].knn == 20 and back[2].num_iterations == 200
    assert [b.uuid for b in back] == [p.uuid for p in pipes]        # uuid prefix = class name (processors.py:1050)


# ------------------------------------------------------------------ sharding (single process)
def test_frame_round_robin():
    from open3dpypro.sharding import frames_of_rank
    F, G = 256, 8
    seen = sorted(f for r in range(G) for f in frames_of_rank(F, r, G))
    assert seen == list(range(F))
    assert frames_of_rank(10, 3, 4) == [3, 7] and frames_of_rank(5, 0, 1) == [0, 1, 2, 3, 4]
    with pytest.raises(ValueError):
        frames_of_rank(5, 4, 4)


def test_slab_cuts_single_process():
    from open3dpypro.sharding import slab_cuts, slab_of, halo_mask
    from open3dpypro.synthetic import lidar_like
    p = lidar_like(200_000, 9)
    axis, cuts = slab_cuts(p, 4, 0.1)
    owner = slab_of(p, axis, cuts)
    share = np.bincount(owner, minlength=4) / len(p)
    assert np.abs(share - 0.25).max() < 0.02                       # near-equal point counts
    mn = p.min(0)
    k = (cuts[1:-1] - mn[axis]) / 0.1
    assert np.abs(k - np.round(k)).max() < 1e-6                    # cuts lie on voxel planes
    vox = np.floor((p[:, axis] - mn[axis]) / 0.1)
    for r in range(3):                                             # so no voxel straddles two slabs
        assert vox[owner == r].max() < vox[owner == r + 1].min()
    h = halo_mask(p, axis, cuts, 1, 0.5)
    assert h.any() and (owner[h] != 1).all()
    x = p[h, axis]
    assert ((x >= cuts[1] - 0.5) & (x < cuts[2] + 0.5)).all()


# ------------------------------------------------------------------ gloo, world_size 2
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
    import torch.distributed as dist
    from open3dpypro import sharding as sh
    from open3dpypro.synthetic import lidar_like
    from oracle import orc
    from scipy.spatial import cKDTree
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        # frames: every rank "detects" a plane per owned frame; results gathered in frame order
        F = 7
        mine = sh.frames_of_rank(F, rank, world)
        local = np.array([[f, 10 * f, rank, 1.0] for f in mine], dtype=np.float64).reshape(-1, 4)
        allp = sh.gather_frame_results(local, F, rank, world)
        assert np.array_equal(allp[:, 0], np.arange(F)) and np.array_equal(allp[:, 2], np.arange(F) % world)
        # slabs: each rank starts with an arbitrary half of the scan
        pts = lidar_like(60_000, 21)
        part = pts[rank::world]
        axis, cuts = sh.slab_cuts(part, world, 0.1)
        ref_axis, ref_cuts = sh.slab_cuts(pts, world, 0.1) if False else (axis, cuts)
        owner = sh.slab_of(pts, axis, cuts)
        owned = pts[owner == rank]
        radius = 0.5
        halo = sh.exchange_halo(owned, axis, cuts, rank, world, radius)
        exp = pts[sh.halo_mask(pts, axis, cuts, rank, radius)]
        assert len(halo) == len(exp) and np.array_equal(np.sort(halo, 0), np.sort(exp, 0))
        # with the halo, every owned query sees exactly its global radius neighbourhood
        loc = np.concatenate([owned, halo], 0)
        cg = cKDTree(pts).query_ball_point(owned[::50], radius, return_length=True)
        cl = cKDTree(loc).query_ball_point(owned[::50], radius, return_length=True)
        assert np.array_equal(cg, cl)
        # segment_plane across slabs: per-hypothesis counts add up to the global counts
        samples = orc.sample_philox(4, len(pts), 3, 16)
        counts = np.empty(16, np.int64)
        errs = np.empty(16)
        for h in range(16):
            pl = orc.plane_model(pts, samples[h])
            counts[h], errs[h] = orc.score_plane(owned, pl, 0.05)
        gc, ge = sh.allreduce_scores(counts, errs)
        for h in (0, 7, 15):
            c, e = orc.score_plane(pts, orc.plane_model(pts, samples[h]), 0.05)
            assert gc[h] == c and abs(ge[h] - e) <= 1e-9 * max(1.0, abs(e))
        # voxel_down_sample across slabs: global bounds, ownership by voxel index, rows of all ranks in
        # the scan's ascending key order == the oracle's rows (the per-rank voxel call is emulated by the oracle here; the
        # device call is exercised at world 2 in tests/test_gpu_sharded.py)
        gmn, gmx = sh.global_bounds(part, world)
        assert np.array_equal(gmn, pts.min(0)) and np.array_equal(gmx, pts.max(0))
        vaxis, vcuts = sh.slab_cuts(part, world, 0.1, min_bound=gmn)
        mine_v = sh.redistribute(part, vaxis, vcuts, rank, world, grid=(gmn, 0.1))
        assert np.all(sh.slab_of(mine_v, vaxis, vcuts, (gmn, 0.1)) == rank)
        loc = orc.voxel_trace(mine_v, 0.1, min_bound=gmn, max_bound=gmx)
        rows = sh.gather_voxel_rows(dict(V=loc["V"], keys=loc["keys"].astype(np.int64), count=loc["count"], centroid=loc["centroid"]), rank, world)
        ref = orc.voxel_trace(pts, 0.1)
        assert rows["V_total"] == ref["V"] and np.array_equal(rows["keys"], ref["keys"]) and np.array_equal(rows["count"], ref["count"])
        assert np.abs(rows["centroid"] - ref["centroid"]).max() <= 1e-12 * np.abs(ref["centroid"]).max()
        q.put((rank, "ok", [float(c) for c in cuts]))
    except Exception as ex:   # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
    finally:
        dist.destroy_process_group()


class _OracleCtx:
    """stand-in for nat.Context in the CPU tests of sharding.slab_estimate_normals: the oracle answers what the device would
    (test infrastructure only).  Honours O3DPP_OPT_QUERY_FIRST the way the library does: only the first m points are answered."""

    def __init__(self):
        self.pts = np.empty((0, 3))
        self.nrm = None
        self.kth = None
        self.query_first = 0
        self.calls = []

    def upload(self, attr, arr):
        self.pts = np.ascontiguousarray(arr, np.float64).reshape(-1, 3)
        self.nrm = None

    def drop(self, attr):
        self.nrm = None

    def size(self):
        return len(self.pts)

    def set_option(self, opt, value):
        from open3dpypro import _native as nat
        assert opt == nat.OPT_QUERY_FIRST
        self.query_first = int(value)

    def estimate_normals(self, mode, knn, radius):
        from oracle import orc
        n = len(self.pts)
        m = self.query_first if 0 < self.query_first < n else n
        self.calls.append((n, m))
        full = orc.estimate_normals(self.pts, "knn" if mode == 0 else "hybrid", knn, radius)
        self.nrm = np.full((n, 3), np.nan)
        self.nrm[:m] = full[:m]                                  # the others are not queried
        idx, d2, cnt = orc.knn(self.pts, knn, mode=mode, radius=radius)
        kth = np.where(cnt >= knn, d2[:, knn - 1], np.inf)
        self.kth = np.nextafter(kth.astype(np.float32), np.float32(np.inf))     # an upper bound, as the device publishes

    def kth_bound(self):
        return self.kth

    def download(self, attr, out):
        out[...] = self.nrm
        return out


def _gloo_knn_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
    import torch.distributed as dist
    from open3dpypro import sharding as sh, _native as nat
    from open3dpypro.synthetic import lidar_like
    from oracle import orc
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        pts = lidar_like(6000, 91)
        part = pts[rank::world]
        axis, cuts = sh.slab_cuts(part, world, 0.1)
        owned = sh.redistribute(part, axis, cuts, rank, world)
        ctx = _OracleCtx()
        nrm, info = sh.slab_estimate_normals(ctx, owned, axis, cuts, rank, world, nat.SEARCH_KNN, 30)
        # the second pass queried only the unsafe points, placed first, among the strips + halo
        assert info["passes"] == 2 and info["halo_points"] > 0
        n_second, m_second = ctx.calls[1]
        assert m_second == info["second_pass_queries"] < n_second == info["second_pass_points"]
        assert ctx.query_first == 0 and len(ctx.pts) == len(owned)          # option reset, owned points left in the context
        # every owned point got the normal of the whole-scan call (same neighbour sets; direction up to rounding)
        whole = orc.estimate_normals(pts, "knn", 30)
        from scipy.spatial import cKDTree
        dist0, where = cKDTree(pts).query(owned)                            # the scan row of every owned point
        assert dist0.max() == 0.0
        cosang = np.abs(np.einsum("ij,ij->i", nrm, whole[where]))
        assert not np.isnan(nrm).any()
        assert cosang.min() > 1 - 1e-9, float(cosang.min())
        # hybrid search: one halo exchange of width r, one pass, exact
        ctx2 = _OracleCtx()
        nrm2, info2 = sh.slab_estimate_normals(ctx2, owned, axis, cuts, rank, world, nat.SEARCH_HYBRID, 30, 0.5)
        assert info2["passes"] == 1 and info2["halo_width"] == 0.5 and len(ctx2.calls) == 1
        whole2 = orc.estimate_normals(pts, "hybrid", 30, 0.5)
        assert np.abs(np.einsum("ij,ij->i", nrm2, whole2[where])).min() > 1 - 1e-9
        q.put((rank, "ok", (len(owned), m_second, n_second)))
    except Exception:   # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_pure_knn_slabs_two_passes():
    """sharding.slab_estimate_normals, pure KNN, world 2 over gloo with the oracle standing in for the device: first pass over
    the owned points, bound all-reduce, halo exchange, second pass that queries only the unsafe points (O3DPP_OPT_QUERY_FIRST),
    result equal to the whole-scan normals"""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_knn_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), [r[1] for r in res]
    assert sum(r[2][0] for r in res) == 6000


def test_gloo_world2_frames_and_slabs():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(30)
    assert all(r[1] == "ok" for r in res), res
    assert res[0][2] == res[1][2]                                 # identical cut planes on both ranks


def test_hilbert_table_is_what_the_generator_writes():
    """csrc/hilbert_lut.cuh (the curve of the neighbour grid as a state machine) is generated and checked against the
    bit-twiddling Hilbert index for every bit count by tools/gen_hilbert_lut.py; the committed header must not drift"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_hilbert_lut", os.path.join(ROOT, "tools", "gen_hilbert_lut.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.main(check_only=True)


def test_first_native_call_may_be_the_shared_context():
    """nat.context() as the very first call into the library (what __graft_entry__.smoke() does): loading the library takes the
    module lock, so context() must not hold it while it constructs.  Fresh interpreter, bounded: a deadlock fails the test
    instead of hanging the suite.  Without a GPU the call raises O3dppError (no CPU fallback)."""
    import subprocess
    code = ("import sys; sys.path.insert(0, %r); from open3dpypro import _native as nat\n"
            "try:\n    c = nat.context(0); print('context', c.size())\n"
            "except nat.O3dppError as e:\n    print('raised', e.code)\n") % os.path.join(ROOT, "open3d-py-extension_b200")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-500:]
    assert r.stdout.startswith(("raised", "context")), r.stdout
