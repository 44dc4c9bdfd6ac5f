"""Stand-ins for the handful of `open3d.geometry` / `open3d.utility` names the reference's
PointCloud.py touches (reference PointCloud.py:24-47, :60-77, :94-126, :338-341), backed by the
device library instead of the Open3D wheel.

`DeviceCloud` plays the role of `o3d.geometry.PointCloud` (the object the reference keeps in
`self.pcd`): float64 `(n,3)` host arrays `points / normals / colors` that stay mutable views, and
the Open3D method names the reference forwards to.  Every geometry method uploads the current
host arrays (they may have been edited in place, exactly as with Open3D's numpy views), runs the
CUDA kernels and writes the results back.  Inside `with cloud.resident():` the cloud stays on
the GPU between calls and host arrays are refreshed only when they are read.
"""
import contextlib
import functools
import os

import numpy as np

from . import _native as nat


class KDTreeSearchParamKNN:
    """o3d.geometry.KDTreeSearchParamKNN(knn=30)"""

    def __init__(self, knn=30):
        self.knn = int(knn)

    def __repr__(self):
        return f"KDTreeSearchParamKNN(knn={self.knn})"


class KDTreeSearchParamRadius:
    """o3d.geometry.KDTreeSearchParamRadius(radius)"""

    def __init__(self, radius):
        self.radius = float(radius)

    def __repr__(self):
        return f"KDTreeSearchParamRadius(radius={self.radius})"


class KDTreeSearchParamHybrid:
    """o3d.geometry.KDTreeSearchParamHybrid(radius, max_nn)"""

    def __init__(self, radius, max_nn):
        self.radius = float(radius)
        self.max_nn = int(max_nn)

    def __repr__(self):
        return f"KDTreeSearchParamHybrid(radius={self.radius}, max_nn={self.max_nn})"


def search_args(param):
    """(mode, knn, radius) from one of the classes above or from a real open3d search param
    (duck typed on the attribute names Open3D uses)."""
    if param is None:
        return nat.SEARCH_KNN, 30, 0.0
    name = type(param).__name__
    has_r, has_k, has_m = hasattr(param, "radius"), hasattr(param, "knn"), hasattr(param, "max_nn")
    if "Hybrid" in name or (has_r and has_m):
        return nat.SEARCH_HYBRID, int(param.max_nn), float(param.radius)
    if "Radius" in name or (has_r and not has_k):
        return nat.SEARCH_RADIUS, 0, float(param.radius)
    if has_k:
        return nat.SEARCH_KNN, int(param.knn), 0.0
    raise TypeError(f"unsupported search parameter {param!r}")


def Vector3dVector(a):
    """o3d.utility.Vector3dVector: a float64 C-contiguous (n,3) copy."""
    a = np.array(a, dtype=np.float64, order="C", copy=True)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError("expected an (n,3) array")
    return a


_EMPTY = np.empty((0, 3), np.float64)


class TraceList:
    """Lazy stand-in for the `List[IntVector]` third result of voxel_down_sample_and_trace:
    row v lists the original indices of voxel row v in ascending order."""

    def __init__(self, order, seg):
        self._order, self._seg = order, seg

    def __len__(self):
        return len(self._seg) - 1

    def __getitem__(self, v):
        if isinstance(v, slice):
            return [self[i] for i in range(*v.indices(len(self)))]
        if v < 0:
            v += len(self)
        return self._order[self._seg[v]:self._seg[v + 1]]

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def _locked(fn):
    """one geometry call at a time per device context: upload -> kernels -> download of one cloud must not interleave
    with another thread's on the same o3dpp_ctx (the C ABI's rule: calls on one ctx must not overlap)"""
    @functools.wraps(fn)
    def wrapper(self, *a, **kw):
        with nat.context_lock(self._device):
            return fn(self, *a, **kw)
    return wrapper


class DeviceCloud:
    def __init__(self, points=None, device=None):
        self._points = _EMPTY
        self._normals = _EMPTY
        self._colors = _EMPTY
        self._device = int(os.environ.get("O3DPP_DEVICE", "0")) if device is None else int(device)
        self._resident = False       # inside resident(): device copy is authoritative
        self._own_ctx = None         # inside resident(): a context of its own (the shared one serves every other cloud)
        self._stale = set()          # host attributes to refresh from the device (resident mode)
        self.seed = None             # RANSAC seed; None draws one from numpy's global RNG
        if points is not None:
            self.points = points

    # ---- host state (Open3D semantics: assignment copies, reads are mutable views)
    def _get(self, name, attr):
        if attr in self._stale:
            ctx = self._ctx()
            setattr(self, name, ctx.download(attr) if ctx.has(attr) else _EMPTY)
            self._stale.discard(attr)
        return getattr(self, name)

    points = property(lambda s: s._get("_points", nat.ATTR_POINTS))
    normals = property(lambda s: s._get("_normals", nat.ATTR_NORMALS))
    colors = property(lambda s: s._get("_colors", nat.ATTR_COLORS))

    @points.setter
    def points(self, a):
        self._points = Vector3dVector(a)
        self._stale.discard(nat.ATTR_POINTS)
        if self._resident:
            self._ctx().upload(nat.ATTR_POINTS, self._points)

    @normals.setter
    def normals(self, a):
        self._normals = Vector3dVector(a)
        self._stale.discard(nat.ATTR_NORMALS)
        if self._resident:
            self._push_attr(nat.ATTR_NORMALS, self._normals)

    @colors.setter
    def colors(self, a):
        self._colors = Vector3dVector(a)
        self._stale.discard(nat.ATTR_COLORS)
        if self._resident:
            self._push_attr(nat.ATTR_COLORS, self._colors)

    def has_points(self):
        return len(self.points) > 0

    def has_normals(self):
        return len(self.points) > 0 and len(self.normals) == len(self.points)

    def has_colors(self):
        return len(self.points) > 0 and len(self.colors) == len(self.points)

    def is_empty(self):
        return not self.has_points()

    def get_center(self):
        return self.points.mean(0) if self.has_points() else np.zeros(3)

    def get_min_bound(self):
        return self.points.min(0) if self.has_points() else np.zeros(3)

    def get_max_bound(self):
        return self.points.max(0) if self.has_points() else np.zeros(3)

    def paint_uniform_color(self, color):
        self.colors = np.tile(np.asarray(color, np.float64).reshape(1, 3), (len(self.points), 1))
        return self

    # ---- device plumbing
    def _ctx(self):
        return self._own_ctx if self._own_ctx is not None else nat.context(self._device)

    def _push_attr(self, attr, arr):
        ctx = self._ctx()
        if len(arr) == len(self._points) and len(arr) > 0:
            ctx.upload(attr, arr)
        else:
            ctx.drop(attr)

    def _push(self, normals=True, colors=False):
        """make the device copy current (no-op in resident mode)"""
        ctx = self._ctx()
        if self._resident:
            return ctx
        ctx.upload(nat.ATTR_POINTS, self._points)
        self._push_attr(nat.ATTR_NORMALS, self._normals if normals else _EMPTY)
        self._push_attr(nat.ATTR_COLORS, self._colors if colors else _EMPTY)
        return ctx

    def _pull(self, ctx, *attrs):
        """take results back: at once, or lazily in resident mode"""
        for a in attrs:
            if self._resident:
                self._stale.add(a)
            else:
                name = {nat.ATTR_POINTS: "_points", nat.ATTR_NORMALS: "_normals", nat.ATTR_COLORS: "_colors"}[a]
                cur = getattr(self, name)
                n = ctx.size()
                if cur.shape == (n, 3) and cur.flags.writeable:
                    ctx.download(a, cur)           # keep existing views valid, as Open3D does
                else:
                    setattr(self, name, ctx.download(a))

    @contextlib.contextmanager
    def resident(self):
        """Keep the cloud on the GPU across several calls (no host round trip per call).  The cloud gets a context of
        its own for the duration: the process-wide context of the device keeps serving every other DeviceCloud and block,
        so work on another cloud inside the `with` cannot replace this cloud's device copy."""
        self._own_ctx = nat.Context(self._device)
        try:
            self._push(normals=True, colors=True)
            self._resident = True
            yield self
        finally:
            try:
                for a in list(self._stale):
                    self._get({nat.ATTR_POINTS: "_points", nat.ATTR_NORMALS: "_normals", nat.ATTR_COLORS: "_colors"}[a], a)
            finally:
                self._resident = False
                self._stale.clear()
                ctx, self._own_ctx = self._own_ctx, None
                ctx.close()

    # ---- the hot path (Open3D method names)
    @_locked
    def transform(self, T):
        T = np.asarray(T, dtype=np.float64)
        if T.shape != (4, 4):
            raise ValueError("transformation must be 4x4")
        if not self.has_points():
            return self
        ctx = self._push(normals=True)
        ctx.transform(T)
        self._pull(ctx, nat.ATTR_POINTS, *([nat.ATTR_NORMALS] if ctx.has(nat.ATTR_NORMALS) else []))
        return self

    def translate(self, t, relative=True):
        t = np.asarray(t, np.float64).reshape(3)
        if not relative:
            t = t - self.get_center()
        if self.has_points():
            self.points[...] = self.points + t
            if self._resident:
                self._ctx().upload(nat.ATTR_POINTS, self._points)
        return self

    def rotate(self, R, center=None):
        R = np.asarray(R, np.float64).reshape(3, 3)
        c = self.get_center() if center is None else np.asarray(center, np.float64).reshape(3)
        T = np.eye(4)
        T[:3, :3] = R
        T[:3, 3] = c - R @ c
        return self.transform(T)

    @_locked
    def estimate_normals(self, search_param=None, fast_normal_computation=True):
        if not fast_normal_computation:
            raise NotImplementedError("only fast_normal_computation=True (Open3D's default) is implemented")
        mode, knn, radius = search_args(search_param)
        if not self.has_points():
            return self
        ctx = self._push(normals=True)
        ctx.estimate_normals(mode, knn, radius)
        self._pull(ctx, nat.ATTR_NORMALS)
        return self

    @_locked
    def segment_plane(self, distance_threshold, ransac_n, num_iterations, probability=0.99999999, seed=None):
        """-> (plane float64[4], ascending inlier indices).  Open3D returns the indices as a
        list[int]; here it is an int32 ndarray (list(...) it if a list is needed)."""
        if seed is None:
            seed = self.seed
        if seed is None:
            seed = int(np.random.randint(0, 2**31 - 1))   # the reference never seeds Open3D either
        ctx = self._push(normals=False)
        if ctx.size() != len(self._points) and not self._resident:
            raise RuntimeError("device cloud out of sync")
        plane, inl = ctx.segment_plane(float(distance_threshold), int(ransac_n), int(num_iterations),
                                       float(probability), sampler=nat.SAMPLER_PHILOX, seed=int(seed))
        return plane, inl

    @_locked
    def voxel_down_sample_and_trace(self, voxel_size, min_bound=None, max_bound=None, approximate_class=False):
        """-> (DeviceCloud of voxel centroids, int32 (V,8) cubic-id matrix, TraceList).
        Rows are in ascending (ix,iy,iz); Open3D's are in unordered_map order.  min_bound is the grid
        origin and max_bound enters the size check, as in Open3D's trace variant; None = the cloud's own
        bounds, which is what the reference's only call site passes (PointCloud.py:339-340)."""
        ctx = self._push(normals=True, colors=True)
        V = ctx.voxel(float(voxel_size), min_bound, max_bound)
        n = ctx.size()
        fields = ["centroid", "cubic_id", "order", "seg"]
        hn, hc = ctx.has(nat.ATTR_NORMALS), ctx.has(nat.ATTR_COLORS)
        if hn:
            fields.append("normal")
        if hc:
            fields.append("color")
        r = ctx.voxel_fetch(V, n, fields)
        out = DeviceCloud(device=self._device)
        out._points = r["centroid"]
        if hn:
            out._normals = r["normal"]
        if hc:
            out._colors = r["color"]
        return out, r["cubic_id"], TraceList(r["order"], r["seg"])

    def voxel_down_sample(self, voxel_size):
        return self.voxel_down_sample_and_trace(voxel_size)[0]

    @_locked
    def voxel_representatives(self, voxel_size, min_bound=None, max_bound=None):
        """ascending `idxmat.max(1)`: the original indices reference PointCloud.py:341 keeps"""
        ctx = self._push(normals=False)
        V = ctx.voxel(float(voxel_size), min_bound, max_bound)
        return ctx.voxel_fetch(V, ctx.size(), ["rep_asc"])["rep_asc"]

    @_locked
    def search_knn(self, knn=30, radius=None):
        """batched KDTreeFlann query of every point: (idx (n,knn) int32, d2, counts)"""
        ctx = self._push(normals=False)
        mode = nat.SEARCH_KNN if radius is None else nat.SEARCH_HYBRID
        return ctx.knn(mode, int(knn), 0.0 if radius is None else float(radius))

    # ---- the callers either side of the path
    @_locked
    def remove_statistical_outlier(self, nb_neighbors, std_ratio, print_progress=False):
        """-> (DeviceCloud of the kept points, kept indices ascending): Open3D RemoveStatisticalOutliers.
        The per-point mean neighbour distances come from the device (o3dpp_knn_mean_distance); the cloud
        mean, deviation and threshold are three scalars over those n numbers."""
        if nb_neighbors < 1 or std_ratio <= 0:
            raise RuntimeError("Illegal input parameters, the number of neighbors and standard deviation ratio must be positive.")
        if not self.has_points():
            return DeviceCloud(device=self._device), np.empty(0, np.int64)
        ctx = self._push(normals=False)
        avg = ctx.knn_mean_distance(int(nb_neighbors))
        valid = avg > 0
        nvalid = int((avg >= 0).sum())
        if nvalid == 0:
            return DeviceCloud(device=self._device), np.empty(0, np.int64)
        cloud_mean = float(np.where(valid, avg, 0.0).sum()) / nvalid
        sq = float(np.where(valid, (avg - cloud_mean) ** 2, 0.0).sum())
        std = (sq / (nvalid - 1)) ** 0.5 if nvalid > 1 else 0.0
        thr = cloud_mean + float(std_ratio) * std
        idx = np.nonzero(valid & (avg < thr))[0]
        out = DeviceCloud(device=self._device)
        out._points = self._points[idx]
        if self.has_normals():
            out._normals = self._normals[idx]
        if self.has_colors():
            out._colors = self._colors[idx]
        return out, idx

    @_locked
    def zsplat(self, grid_size, x_min=None, x_max=None, y_min=None, y_max=None):
        """ZDepthViewer's scatter (reference processors.py:876-894) on the device: float32 (grid, grid) z of the
        last point per cell (NaN where empty) and that point's index.  Bounds default to the float32 min / max."""
        p = self._points
        f = lambda v, d: float(np.float32(d)) if v is None else float(v)
        ctx = self._push(normals=False)
        return ctx.zsplat(int(grid_size), f(x_min, p[:, 0].min()), f(x_max, p[:, 0].max()), f(y_min, p[:, 1].min()),
                          f(y_max, p[:, 1].max()))


class KDTreeFlann:
    """o3d.geometry.KDTreeFlann over a DeviceCloud for the single-query calls of the reference
    (PointCloud.py:148-163).  One query is a host job (an O(n) numpy pass, results ordered by (d2, index) as
    the device searches order theirs); the batched searches are DeviceCloud.search_knn / o3dpp_knn."""

    def __init__(self, cloud):
        self._pts = np.asarray(cloud.points, dtype=np.float64)

    def _d2(self, q):
        d = self._pts - np.asarray(q, dtype=np.float64).reshape(1, 3)
        return (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]

    def search_knn_vector_3d(self, query, knn):
        d2 = self._d2(query)
        k = min(int(knn), len(d2))
        order = np.lexsort((np.arange(len(d2)), d2))[:k]
        return k, order.astype(np.int32), d2[order]

    def search_radius_vector_3d(self, query, radius):
        d2 = self._d2(query)
        hit = np.nonzero(d2 < float(radius) * float(radius))[0]
        order = hit[np.lexsort((hit, d2[hit]))]
        return len(order), order.astype(np.int32), d2[order]

    def search_hybrid_vector_3d(self, query, radius, max_nn):
        k, idx, d2 = self.search_radius_vector_3d(query, radius)
        k = min(k, int(max_nn))
        return k, idx[:k], d2[:k]
