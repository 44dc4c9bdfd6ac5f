"""The probe SURVEY.md §7 asks for: the oracle (and, on a GPU box, the CUDA path) against REAL Open3D, driven through the
reference's own wrappers (reference open3dpypro/PointCloud.py:60-62 transform, :68-73 estimate_normals, :75-77
segment_plane, :338-341 voxel_down_sample_and_trace).

Auto-skips unless `import open3d` works (also tried with baseline/_ref on the path).  Open3D 0.19.0 is an un-vendored
binary wheel that is installed neither here nor on the GPU box, so this file normally SKIPS; it is the one test that can
move parity from "pinned by the restatement only" to "pinned by Open3D" the day a wheel is reachable.  Where Open3D is
implementation-defined the comparison is the one SURVEY.md §3 prescribes: voxel rows as sets keyed by the
representative, KNN sets off the boundary-tie rows, normals up to sign, RANSAC through a fixed mt19937 seed."""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_REF = os.path.join(ROOT, "baseline", "_ref")
if os.path.isdir(_REF) and _REF not in sys.path:
    sys.path.append(_REF)
o3d = pytest.importorskip("open3d", reason="Open3D 0.19.0 wheel not installed (no network): parity stays pinned by the oracle only")


def _reference_wrapper():
    """the reference's PointCloud module under an alias (this repo's package has the same name), or None"""
    for base in (os.environ.get("O3DPP_REFERENCE"), "/root/reference", _REF):
        path = os.path.join(base, "open3dpypro", "PointCloud.py") if base else None
        if path and os.path.exists(path):
            spec = importlib.util.spec_from_file_location("_ref_open3dpypro_PointCloud", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    return None


@pytest.fixture(scope="module")
def ref_cloud(bunny):
    mod = _reference_wrapper()

    def make(pts):
        if mod is not None:
            return mod.PointCloud(np.asarray(pts, np.float64))
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(np.asarray(pts, np.float64))

        class Thin:        # the four delegations of PointCloud.py, verbatim in behaviour
            pcd = pc
            def transform(self, T): pc.transform(T); return self
            def estimate_normals(self, method="default", param=o3d.geometry.KDTreeSearchParamKNN(30)): pc.estimate_normals(param); return self
            def segment_plane(self, thickness=0.01, ransac_n=3, num_iterations=450): return pc.segment_plane(thickness, ransac_n, num_iterations)
            def get_points(self): return np.asarray(pc.points)
            def get_normals(self): return np.asarray(pc.normals)
        return Thin()
    return make


def test_transform_vs_open3d(orc, bunny, ref_cloud):
    rng = np.random.default_rng(0)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    T = np.eye(4); T[:3, :3] = q; T[:3, 3] = rng.normal(size=3); T[3] = [0.01, -0.02, 0.03, 1.5]
    got = ref_cloud(bunny).transform(T).get_points()
    exp, _ = orc.transform(bunny, T)
    assert np.array_equal(got, exp), "oracle's ((T0 x + T1 y) + T2 z) + T3 order differs from Open3D's Eigen product"


@pytest.mark.parametrize("vs,V", [(0.002, 15930), (0.005, 3017), (0.01, 751)])
def test_voxel_vs_open3d(orc, bunny, vs, V):
    pc = o3d.geometry.PointCloud(); pc.points = o3d.utility.Vector3dVector(bunny)
    down, idxmat, vec = pc.voxel_down_sample_and_trace(vs, pc.get_min_bound(), pc.get_max_bound(), False)   # PointCloud.py:339-340
    idxmat = np.asarray(idxmat)
    ref = orc.voxel_trace(bunny, vs)
    assert len(idxmat) == ref["V"] == V
    a = np.argsort(idxmat.max(1)); b = np.argsort(ref["rep"])            # rows keyed by the representative index
    assert np.array_equal(idxmat.max(1)[a], ref["rep"][b])
    assert np.array_equal(idxmat[a], ref["cubic_id"][b])
    cen = np.asarray(down.points)
    assert np.abs(cen[a] - ref["centroid"][b]).max() <= 1e-5 * np.abs(ref["centroid"]).max()


def test_knn_and_normals_vs_open3d(orc, bunny, ref_cloud):
    pc = o3d.geometry.PointCloud(); pc.points = o3d.utility.Vector3dVector(bunny)
    tree = o3d.geometry.KDTreeFlann(pc)
    eidx, ed2, _ = orc.knn(bunny, 31)
    tie = ed2[:, 29] == ed2[:, 30]                    # Open3D's winner among exact boundary ties is traversal-defined
    assert tie.sum() == 137
    rows = np.random.default_rng(1).choice(np.nonzero(~tie)[0], 3000, replace=False)
    for i in rows:
        k, idx, d2 = tree.search_knn_vector_3d(bunny[i], 30)
        assert k == 30 and set(idx) == set(eidx[i, :30].tolist())
        assert np.array_equal(np.asarray(d2), ed2[i, :30])
    n_o3d = ref_cloud(bunny).estimate_normals().get_normals()
    n_orc = orc.estimate_normals(bunny, "knn", 30)
    ang = np.arccos(np.clip(np.abs(np.einsum("ij,ij->i", n_o3d, n_orc)), -1, 1))
    assert ang[~tie].max() < 1e-4


def test_segment_plane_vs_open3d(orc, bunny):
    """fixed mt19937 seed: sample sequence, winner, inlier set and refit plane"""
    pc = o3d.geometry.PointCloud(); pc.points = o3d.utility.Vector3dVector(bunny)
    o3d.utility.random.seed(5)
    os.environ["OMP_NUM_THREADS"] = "1"              # the deterministic definition is the single-thread replay
    plane, inl = pc.segment_plane(0.01, 3, 450)
    samples = orc.sample_mt19937(5, len(bunny), 3, 450)
    eplane, einl = orc.segment_plane(bunny, 0.01, 3, 450, samples)
    assert np.array_equal(np.asarray(inl), einl)
    assert np.abs(np.asarray(plane) - eplane).max() < 1e-9


@pytest.mark.gpu
def test_cuda_path_vs_open3d(bunny, ref_cloud):
    from open3dpypro.PointCloud import PointCloud
    mine = PointCloud(bunny).estimate_normals().get_normals()
    theirs = ref_cloud(bunny).estimate_normals().get_normals()
    ang = np.arccos(np.clip(np.abs(np.einsum("ij,ij->i", mine, theirs)), -1, 1))
    assert np.quantile(ang, 0.995) < 1e-4            # all but the boundary-tie rows
