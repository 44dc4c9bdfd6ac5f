"""open3dpypro.PointCloud on the B200 library.

Drop-in for the hot-path surface of the reference class chain (reference open3dpypro/PointCloud.py):
constructor / getters / setters (:10-49, :94-126), `transform` (:60-62), `estimate_normals`
(:68-73), `segment_plane` (:75-77), `_select_by_idx` (:185-204), `voxel_down_sample_and_trace` /
`voxel_down_sample` (:338-341, :361-362) and their direct callers `rotate_to_plane` (:931-939) and
`seg_planes` (:941-985).  Signatures, return shapes and error behaviour follow the reference;
`self.pcd` is a `geometry.DeviceCloud` instead of an `o3d.geometry.PointCloud`.

Left out on purpose: GL windows (`draw`), LAS/E57/image codecs, cv2/sklearn
image-space helpers.
"""
import uuid as _uuid

import numpy as np

from .geometry import (DeviceCloud, KDTreeFlann, KDTreeSearchParamHybrid, KDTreeSearchParamKNN,  # noqa: F401
                       KDTreeSearchParamRadius)


def _given(a):
    return a is not None and len(a) > 0 and np.prod(np.shape(a)) > 0


class PointCloudBase:
    def __init__(self, xyz=None, rgb=None, normals=None, intensity=None, labels=None,
                 row_index=None, column_index=None, e57=None):
        self.e57 = None
        self.intensity, self.labels, self.row_index, self.column_index = [], [], [], []
        if isinstance(xyz, DeviceCloud):
            self.pcd = xyz
        else:
            self.pcd = DeviceCloud()
            if _given(intensity):
                self.set_intensity(intensity)
            if _given(row_index):
                self.row_index = row_index
            if _given(column_index):
                self.column_index = column_index
            if _given(xyz):
                self.set_points(xyz)
            if _given(normals):
                self.set_normals(normals)
            if _given(rgb):
                rgb = np.asarray(rgb)
                if rgb.max() > 1.0 or rgb.min() < 0.0:      # reference :37-39: stretch to [0,1]
                    rgb = rgb * 1.0
                    rgb = (rgb - rgb.min()) / (rgb.max() - rgb.min())
                self.set_rgb(rgb)
        self.pcd_tree = None
        self.scan_No = -1
        self.uuid = _uuid.uuid4()

    # forwarded Open3D conveniences (reference :43-47 binds them once; forwarding survives clear())
    def get_center(self):
        return self.pcd.get_center()

    def has_colors(self):
        return self.pcd.has_colors()

    def has_points(self):
        return self.pcd.has_points()

    def is_empty(self):
        return self.pcd.is_empty()

    def rotate(self, R, center=None):
        return self.pcd.rotate(R, center)

    def clear(self):
        self.pcd = DeviceCloud()

    def size(self):
        return len(self.get_points())

    # ---- hot path
    def transform(self, T):
        self.pcd.transform(T)
        return self

    def translate(self, t):
        self.pcd.translate(t)
        return self

    def estimate_normals(self, method="default", param=None):
        if method == "poisson":
            raise NotImplementedError("orient_normals_consistent_tangent_plane is outside the accelerated path")
        self.pcd.estimate_normals(KDTreeSearchParamKNN(30) if param is None else param)
        return self

    def segment_plane(self, thickness=0.01, ransac_n=3, num_iterations=450, distance_threshold=None,
                      probability=0.99999999, seed=None):
        """-> ([a,b,c,d], inlier indices).  `distance_threshold` is accepted as an alias of
        `thickness` because reference :965 calls the wrapper with that keyword."""
        if distance_threshold is not None:
            thickness = distance_threshold
        plane, inliers = self.pcd.segment_plane(thickness, ransac_n, num_iterations, probability, seed)
        return [float(v) for v in plane], inliers

    # ---- attributes
    def has_rgb(self):
        return self.has_points() and self.size() == len(self.pcd.colors)

    def has_normals(self):
        return self.has_points() and self.pcd.has_normals()

    def has_intensity(self):
        return self.has_points() and self.size() == len(self.intensity)

    def has_labels(self):
        return self.has_points() and self.size() == len(self.labels)

    def has_col_row(self):
        return self.has_points() and self.size() == len(self.column_index) == len(self.row_index)

    def set_rgb(self, colors):
        assert np.shape(colors)[1] == 3, 'points shape must be (n,3)'
        self.pcd.colors = colors
        return self

    def set_points(self, points):
        assert np.shape(points)[1] == 3, 'points shape must be (n,3)'
        self.pcd.points = points
        return self

    def set_normals(self, normals):
        assert np.shape(normals)[1] == 3, 'normals shape must be (n,3)'
        self.pcd.normals = normals
        return self

    def set_intensity(self, intensity):
        assert np.shape(intensity)[1] == 1, 'intensity shape must be (n,1)'
        self.intensity = intensity
        return self

    def set_labels(self, labels):
        assert np.shape(labels)[1] == 1, 'labels shape must be (n,1)'
        self.labels = labels
        return self

    def get_points(self):
        return self.pcd.points

    def get_colors(self):
        return self.pcd.colors

    def get_normals(self):
        return self.pcd.normals

    def get_intensity(self):
        return self.intensity

    def get_labels(self):
        return self.labels

    def get_col_row(self):
        return (self.column_index, self.row_index)

    def set_uniform_label(self, label):
        self.labels = np.full((self.size(), 1), label, dtype=int)
        return self

    def set_uniform_intensity(self, intensity):
        return self.set_intensity(np.ones((self.size(), 1), dtype=int) * intensity)

    def get_aabb(self):
        return self.pcd.get_min_bound(), self.pcd.get_max_bound()

    # single-query KD-tree calls (reference :148-163)
    def calc_KDtree(self):
        self.pcd_tree = KDTreeFlann(self.pcd)
        return self.pcd_tree

    def get_KDtree(self):
        if self.pcd_tree is None:
            self.calc_KDtree()
        return self.pcd_tree

    def get_points_by_knn(self, point_idx, max_nn=1000000):
        k, idx, d2 = self.get_KDtree().search_knn_vector_3d(self.pcd.points[point_idx], max_nn)
        return k, idx, d2

    def get_points_radius(self, point_idx, search_radius=30.0):
        k, idx, d2 = self.get_KDtree().search_radius_vector_3d(self.pcd.points[point_idx], search_radius)
        return k, idx, d2

    def read_pcd(self, filename, **_ignored):
        from .pcdio import read_pcd
        xyz, intensity = read_pcd(filename)
        out = self.__class__(xyz)
        if intensity is not None:
            out.set_intensity(intensity.reshape(-1, 1))
        return out


class PointCloudSelections(PointCloudBase):
    def clone(self, invert=False):
        return self._select_by_idx(np.arange(self.size()), invert=invert)

    def _select_by_idx(self, indices, invert=False):
        """Boolean-mask gather: the result is in ascending original index whatever the order of
        `indices` (reference :185-204)."""
        keep = np.full(self.size(), bool(invert))
        keep[np.asarray(indices, dtype=np.int64)] = not invert
        out = self.__class__()
        if self.has_points():
            out.set_points(self.get_points()[keep])
        if self.has_colors():
            out.set_rgb(self.get_colors()[keep])
        if self.has_normals():
            out.set_normals(self.get_normals()[keep])
        if self.has_intensity():
            out.intensity = np.asarray(self.intensity)[keep].copy()
        if self.has_labels():
            out.labels = np.asarray(self.labels)[keep].copy()
        return out

    def _bool2index(self, bools, invert=False):
        b = np.asarray(bools)
        return np.where(~b if invert else b)[0]

    def select_by_bool(self, bools, invert=False):
        return self._select_by_idx(self._bool2index(bools), invert=invert)

    def get_index_by_plane(self, model, thickness=0.03, invert=False):
        a, b, c, d = model
        s = (self.get_points() @ np.asarray([a, b, c], dtype=np.float64) + d) / (a * a + b * b + c * c) ** 0.5
        hit = (s > thickness[0]) & (s < thickness[1]) if isinstance(thickness, tuple) else np.abs(s) < thickness
        return self._bool2index(hit, invert=invert)

    def select_by_plane(self, model, thickness=0.03, invert=False):
        return self._select_by_idx(self.get_index_by_plane(model, thickness), invert=invert)

    def get_index_by_XYZ(self, comfunc=lambda Xs, Ys, Zs: np.ones(len(Xs), dtype=bool), invert=False):
        p = self.get_points()
        return self._bool2index(comfunc(p[:, 0], p[:, 1], p[:, 2]), invert=invert)

    def select_by_XYZ(self, comfunc=lambda Xs, Ys, Zs: np.ones(len(Xs), dtype=bool), invert=False):
        return self._select_by_idx(self.get_index_by_XYZ(comfunc), invert=invert)

    def get_index_by_radius(self, r, invert=False):
        return self.get_index_by_XYZ(lambda Xs, Ys, Zs: (Xs**2 + Ys**2 + Zs**2)**0.5 <= r, invert=invert)

    def select_by_radius(self, r, invert=False):
        return self._select_by_idx(self.get_index_by_radius(r), invert=invert)

    def get_index_by_aabb(self, aabb_min, aabb_max, invert=False):
        p = self.get_points()
        inside = np.all((p >= np.asarray(aabb_min).reshape(1, 3)) & (p <= np.asarray(aabb_max).reshape(1, 3)), axis=1)
        return self._bool2index(inside, invert=invert)

    def select_by_aabb(self, aabb_min, aabb_max, invert=False):
        return self._select_by_idx(self.get_index_by_aabb(aabb_min, aabb_max), invert=invert)

    def select_by_topN(self, n):
        return self._select_by_idx(np.arange(self.size())[:n])


class PointCloudUtility(PointCloudSelections):
    def paint_uniform_color(self, color=(1., 0., 0.)):
        self.pcd.paint_uniform_color(color)
        return self

    def centralize(self):
        return self.translate(-self.get_center())

    def voxel_down_sample_and_trace(self, voxel_size):
        """-> (PointCloud of one original point per voxel, ascending index; int32 (V,8) idxmat;
        per-voxel index lists).  As in the reference (:338-341) Open3D's centroids are dropped in
        favour of `_select_by_idx(idxmat.max(1))`.  idxmat / lists rows: ascending voxel key."""
        _cent, idxmat, vec = self.pcd.voxel_down_sample_and_trace(voxel_size)
        return self._select_by_idx(idxmat.max(1)), idxmat, vec

    def voxel_down_sample(self, voxel_size):
        """Same cloud as voxel_down_sample_and_trace(...)[0], without fetching the trace."""
        return self._select_by_idx(self.pcd.voxel_representatives(voxel_size))

    def random_down_sample(self, down_sample_ratio=0.1):
        if down_sample_ratio <= 0 or down_sample_ratio > 1:
            return self.__class__()
        idx = np.random.permutation(self.size())[:int(self.size() * down_sample_ratio)]
        return self._select_by_idx(idx)

    def uniform_down_sample(self, down_sample_ratio=0.1):
        if down_sample_ratio <= 0 or down_sample_ratio > 1:
            return self.__class__()
        return self._select_by_idx(np.arange(self.size())[::int(1 / down_sample_ratio)])

    def remove_statistical_outlier(self, nb_neighbors=20, std_ratio=2.0, print_progress=False):
        """-> (PointCloud of the kept points, their indices): reference :370-372"""
        idx = self.pcd.remove_statistical_outlier(nb_neighbors=nb_neighbors, std_ratio=std_ratio)[1]
        return self._select_by_idx(idx), idx

    def distance2plane(self, plane):
        a, b, c, d = plane
        return (self.get_points() @ np.asarray([a, b, c], dtype=np.float64) + d) / (a * a + b * b + c * c) ** 0.5

    def project2plane(self, plane):
        plane = np.asarray(plane, dtype=np.float64)
        return self.__class__(self.get_points() - self.distance2plane(plane).reshape(-1, 1) * plane[:3].reshape(1, 3))


class PointCloudAdvanceIO(PointCloudUtility):
    """LAS / E57 / image codecs of the reference are out of scope."""


class PointCloud(PointCloudAdvanceIO):
    def rotation_matrix_from_vectors(self, vec1, vec2):
        a = np.asarray(vec1, np.float64).reshape(3)
        b = np.asarray(vec2, np.float64).reshape(3)
        a, b = a / np.linalg.norm(a), b / np.linalg.norm(b)
        v = np.cross(a, b)
        if not v.any():
            return np.eye(3)
        K = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
        return np.eye(3) + K + K @ K * ((1 - a @ b) / (v @ v))

    def rotate_to_plane(self, plane):
        a, b, c, d = plane
        n = np.array([a, b, c], dtype=np.float64)
        T = np.eye(4)
        T[2, 3] = d
        T[:3, :3] = self.rotation_matrix_from_vectors(n / np.linalg.norm(n), (0, 0, 1))
        self.transform(T)
        return self, T

    rotate_by_normal = rotate_to_plane

    def to_2D_Img(self, resolution=0.05, color=False):
        """Top-down occupancy / colour image of the cloud (reference :785-823): pixel = round((p - min) / resolution),
        later points overwrite earlier ones.  -> (img, minx, T, miny, invT, resolution, idx2img)"""
        usecolor = bool(color and self.has_colors())
        if color and not usecolor:
            print('has no rgb data, cannot use color!')
        p = self.get_points()
        minx, miny = p[:, 0].min(), p[:, 1].min()
        ir = 1 / resolution
        T = np.diag([ir, ir, ir, 1.0]) @ np.array([[1., 0., 0., -minx], [0., 1., 0., -miny], [0., 0., 1., 0.], [0., 0., 0., 1.]])
        px = ((T[:3, :3] @ p.T).T + T[:3, 3]).round().astype(int)
        ix, iy = px[:, 0], px[:, 1]
        sizex, sizey = ix.max(), iy.max()
        ix = np.clip(ix, 0, sizex - 1)
        iy = np.clip(iy, 0, sizey - 1)
        if usecolor:
            img = np.zeros((sizey, sizex, 3), np.uint8)
            img[iy, ix] = (self.get_colors() * 255).astype(np.uint8)
        else:
            img = np.zeros((sizey, sizex), np.uint8)
            img[iy, ix] = 255
        return img, minx, T, miny, np.linalg.inv(T), resolution, list(zip(ix, iy))

    def seg_planes(self, thickness=0.01, ransac_n=3, num_iterations=450, top_n=10e9, minPointsRatio=0.1):
        """Peel planes until fewer than `minPointsRatio` of the points remain (reference :941-985).
        The reference's loop dies on its first pass (keyword mismatch swallowed by `except`); this one runs
        as its docstring describes.  The cloud stays on the GPU across the peels: every pass is
        o3dpp_segment_plane + o3dpp_select_inliers(invert) on the device copy, only the inlier indices
        (4 B each) come back; the host keeps the list of surviving original indices to cut the attribute
        arrays.  Pass i uses RANSAC seed `self.pcd.seed + i` (a random base seed if none is set)."""
        from . import _native as nat
        planes, pcds, aabbs = [], [], []
        n0 = self.size()
        if n0 < ransac_n:
            return planes, pcds, aabbs
        ctx = nat.context(self.pcd._device)
        ctx.upload(nat.ATTR_POINTS, self.get_points())
        ctx.drop(nat.ATTR_NORMALS)
        ctx.drop(nat.ATTR_COLORS)
        alive = np.arange(n0)
        base = self.pcd.seed if self.pcd.seed is not None else int(np.random.randint(0, 2**31 - 1))
        while len(alive) / n0 > minPointsRatio:
            try:
                plane, inl = ctx.segment_plane(float(thickness), int(ransac_n), int(num_iterations),
                                               seed=int(base) + len(planes))
            except Exception as e:      # same containment as the reference
                print(e)
                break
            if len(inl) == 0:
                break
            planes.append(plane)
            part = self._select_by_idx(alive[inl])
            pcds.append(part)
            aabbs.append(part.get_aabb())
            ctx.select_inliers(invert=True)
            alive = np.delete(alive, inl)
            if len(planes) > top_n:
                break
        rest = self._select_by_idx(alive) if len(alive) else self.__class__()
        pcds.append(rest)
        aabbs.append(rest.get_aabb())
        return planes, pcds, aabbs
