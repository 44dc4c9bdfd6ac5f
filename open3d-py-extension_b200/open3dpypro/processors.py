"""Pipeline blocks of the hot path, re-pointed at the B200 library.

Same block names, fields and `forward_raw` contracts as the reference's `Processors` namespace
(open3dpypro/processors.py): `CPUNormals` (:228-249), `VoxelDownsample` (:418-474, cpu model),
`PlaneDetection` (:502-699, cpu model incl. the d>=0 flip and the EMA) and `PlaneNormalize`
(:701-759), plus the runner `PointCloudMatProcessors` (:1043-1098) with its JSON round trip keyed on
the class name in `uuid`.  The names keep "CPU" for drop-in compatibility; the work runs on the GPU
selected round-robin per input mat (`cuda:{i % num_gpus}`, reference :206).
"""
import json
import multiprocessing
import os
from typing import List, Optional

import numpy as np

from . import _native as nat
from .PointCloudMat import PointCloudMat, PointCloudMatInfo, PointCloudMatProcessor, ShapeType

logger = print


def _xyz(pcd):
    """contiguous (N,3) view/copy of the position columns, dtype kept (f32 uploads as f32)"""
    a = pcd[:, :3]
    if a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float64)
    return np.ascontiguousarray(a)


class _DeviceBlock(PointCloudMatProcessor):
    """common part of the device blocks: one library context per GPU, created lazily in the
    process that runs the block (so blocks survive `run_async`'s JSON round trip and fork)"""
    multi_gpu: int = -1

    def _device(self, i):
        """mat i -> cuda:{i % num_gpus} inside one process (reference :206); a process pinned to one GPU
        (one process per GPU under torchrun: O3DPP_DEVICE = LOCAL_RANK) keeps every mat on that GPU"""
        if not self.num_gpus:
            self.devices_info(gpu=True, multi_gpu=self.multi_gpu)
            if not self.num_gpus:
                raise nat.O3dppError(-1, "no CUDA device: this block has no CPU fallback")
        pin = os.environ.get("O3DPP_DEVICE")
        if pin is not None and pin != "":
            return int(pin)
        return i % len(self.num_devices)

    def _ctx(self, i):
        return nat.context(self._device(i))


class Processors:
    class DoingNothing(PointCloudMatProcessor):
        title: str = 'doing_nothing'

        def validate_pcd(self, idx, pcd):
            pass

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            return pcds_data

    class CPUNormals(_DeviceBlock):
        title: str = 'cpu_calc_normals'
        input_shape_types: List[ShapeType] = []
        knn: int = 30              # the reference block uses PointCloudBase.estimate_normals() defaults
        radius: float = 0.0        # > 0 selects KDTreeSearchParamHybrid(radius, knn)

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()
            self.input_shape_types.append(pcd.shape_type)

        def build_out_mats(self, validated_pcds, converted_raw_pcds):
            self.out_mats = [PointCloudMat(shape_type=old.info.shape_type.add_normals()).build(p)
                             for old, p in zip(validated_pcds, converted_raw_pcds)]
            return self.out_mats

        def forward_raw(self, pcds_data: List[np.ndarray], pcds_info: List[PointCloudMatInfo] = [], meta={}):
            res = []
            for i, pcd in enumerate(pcds_data):
                if i < len(self.input_shape_types) and self.input_shape_types[i].contains_normals():
                    continue
                ctx = self._ctx(i)
                ctx.upload(nat.ATTR_POINTS, _xyz(pcd))
                ctx.drop(nat.ATTR_NORMALS)
                mode = nat.SEARCH_HYBRID if self.radius > 0 else nat.SEARCH_KNN
                ctx.estimate_normals(mode, self.knn, self.radius)
                out = np.empty((len(pcd), pcd.shape[1] + 3), np.float64)      # = np.hstack([pcd, normals]) without the temporaries
                out[:, :pcd.shape[1]] = pcd
                out[:, pcd.shape[1]:] = ctx.download(nat.ATTR_NORMALS)
                res.append(out)
            return res

    class VoxelDownsample(_DeviceBlock):
        title: str = 'voxel_sample'
        voxel_size: float = 0.1

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            res = []
            for i, pcd in enumerate(pcds_data):
                ctx = self._ctx(i)
                ctx.upload(nat.ATTR_POINTS, _xyz(pcd))
                V = ctx.voxel(self.voxel_size)
                rep = ctx.voxel_fetch(V, len(pcd), ["rep"])["rep"]   # rows in ascending voxel key
                res.append(pcd[rep])
            return res

    class PlaneDetection(_DeviceBlock):
        title: str = 'plane_detection'
        distance_threshold: float = 0.01
        arange: bool = False
        best_planes: List[List[float]] = []
        alpha: float = 0.0
        num_iterations: int = 512
        num_iteration_batch: int = 256     # kept for config compatibility; batching is internal
        voxel_size: float = 0.0
        z_lower: Optional[float] = None
        z_upper: Optional[float] = None
        seed: Optional[int] = None
        _calls: int = 0

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()
            self.best_planes.append([0., 0., 0., 0.])

        def detect(self, i, pcd):
            xyz = _xyz(pcd)
            if self.z_lower is not None and self.z_upper is not None:
                xyz = np.ascontiguousarray(xyz[(self.z_lower < xyz[:, 2]) & (xyz[:, 2] < self.z_upper)])
            ctx = self._ctx(i)
            ctx.upload(nat.ATTR_POINTS, xyz)
            seed = self.seed if self.seed is not None else int(np.random.randint(0, 2**31 - 1))
            plane, _ = ctx.segment_plane(self.distance_threshold, 3, self.num_iterations,
                                         seed=seed + self._calls, fetch_inliers=False)
            n, d = plane[:3], plane[3]
            if np.dot(n, np.zeros(3) - (-d * n)) < 0:   # normal towards the sensor origin
                plane = -plane
            return plane

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            if self.voxel_size > 0.0:
                pcds_data = Processors.VoxelDownsample(voxel_size=self.voxel_size, multi_gpu=self.multi_gpu).forward_raw(pcds_data)
            while len(self.best_planes) < len(pcds_data):
                self.best_planes.append([0., 0., 0., 0.])
            for i, pcd in enumerate(pcds_data):
                plane = self.detect(i, pcd)
                self.best_planes[i] = (np.asarray(self.best_planes[i]) * (1.0 - self.alpha) + plane * self.alpha).tolist()
            self._calls += 1
            meta[self.uuid] = self.best_planes
            return pcds_data

    class PlaneNormalize(_DeviceBlock):
        title: str = 'plane_normalize'
        detection_uuid: str
        filter_pcd: bool = False

        def validate_pcd(self, idx, pcd):
            pcd.require_ndarray()

        @staticmethod
        def rotation_matrix_from_vectors(vec1, vec2):
            a = vec1 / np.linalg.norm(vec1)
            b = vec2 / np.linalg.norm(vec2)
            v = np.cross(a, b)
            s = np.linalg.norm(v)
            if s < 1e-6:
                return np.eye(3)
            K = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
            return np.eye(3) + K + K @ K * ((1 - np.dot(a, b)) / (s ** 2))

        @classmethod
        def plane_to_T(cls, plane):
            a, b, c, d = plane
            n = np.array([a, b, c], dtype=np.float64)
            R = cls.rotation_matrix_from_vectors(n, np.array([0., 0., 1.]))
            T = np.eye(4)
            T[:3, :3] = R
            T[:3, 3] = -R @ (-d * n / np.dot(n, n))
            return T

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            models = meta[self.detection_uuid]
            self.forward_T = [None] * len(pcds_data)
            res = []
            for i, pcd in enumerate(pcds_data):
                T = self.plane_to_T(models[i])
                ctx = self._ctx(i)
                ctx.upload(nat.ATTR_POINTS, _xyz(pcd))
                ctx.drop(nat.ATTR_NORMALS)
                ctx.transform(T)
                out = np.empty_like(pcd)                                      # = np.hstack([xyz.astype(dtype), pcd[:, 3:]])
                out[:, :3] = ctx.download(nat.ATTR_POINTS)
                out[:, 3:] = pcd[:, 3:]
                res.append(out)
                self.forward_T[i] = T.tolist()
            return res


    class RadiusSelection(PointCloudMatProcessor):
        """keep the points within `radius` of the sensor origin (reference :372-416, cpu model)"""
        title: str = 'radius_selection'
        radius: float = 5.0

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            res = []
            for pcd in pcds_data:
                xyz = pcd[:, :3].astype(np.float64)
                res.append(xyz[(xyz[:, 0]**2 + xyz[:, 1]**2 + xyz[:, 2]**2)**0.5 <= self.radius])   # the reference returns xyz only
            return res

    class RemoveStatisticalOutlier(_DeviceBlock):
        """reference :486-500; the neighbour distances come from the device KNN"""
        title: str = 'remove_statistical_outlier'
        nb_neighbors: int = 20
        std_ratio: float = 2.0
        print_progress: bool = False

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            from .geometry import DeviceCloud
            res = []
            for i, pcd in enumerate(pcds_data):
                cloud = DeviceCloud(_xyz(pcd), device=self._device(i))
                res.append(pcd[cloud.remove_statistical_outlier(self.nb_neighbors, self.std_ratio)[1]])
            return res

    class ZDepthViewer(_DeviceBlock):
        """Z-depth image of every cloud (reference :838-927).  The scatter of the points into the grid runs on
        the device (o3dpp_zsplat, last point of a cell wins as numpy's fancy assignment does); normalisation and
        colour map act on the grid x grid image.  Headless: the images are kept in `_img_data`, no window is
        opened (cv2.imshow in the reference: unavailable on this box)."""
        title: str = 'z_depth_viewer'
        bg: tuple = (0, 0, 0)
        grid_size: int = 256
        img_size: int = 0
        _img_data: list = []
        inf: float = 9999999999
        x_min: float = inf
        x_max: float = inf
        y_min: float = inf
        y_max: float = inf
        z_min: float = inf
        z_max: float = inf

        def validate_pcd(self, idx, pcd: PointCloudMat):
            pcd.require_ndarray()

        def z_image(self, i, pcd):
            g = self.grid_size
            img = np.zeros((g, g, 3), dtype=np.uint8)
            if len(pcd) == 0:
                return img
            xyz = _xyz(pcd)
            x, y = xyz[:, 0], xyz[:, 1]
            x_min = x.min() if self.x_min == self.inf else self.x_min
            x_max = x.max() if self.x_max == self.inf else self.x_max
            y_min = y.min() if self.y_min == self.inf else self.y_min
            y_max = y.max() if self.y_max == self.inf else self.y_max
            if x_max - x_min < 1e-5 or y_max - y_min < 1e-5:
                return img
            ctx = self._ctx(i)
            ctx.upload(nat.ATTR_POINTS, xyz)
            z_grid, _last = ctx.zsplat(g, x_min, x_max, y_min, y_max)
            z_min = z_grid.min() if self.z_min == self.inf else self.z_min
            z_max = z_grid.max() if self.z_max == self.inf else self.z_max
            if np.isnan(z_grid).any():
                z_grid = np.where(np.isnan(z_grid), z_min, z_grid)
            with np.errstate(invalid='ignore'):
                z_norm = np.zeros_like(z_grid) if z_max - z_min < 1e-5 else (z_grid - z_min) / (z_max - z_min)
                z_img = (z_norm * 255).astype(np.uint8)
            try:
                import cv2
                img = cv2.applyColorMap(z_img, cv2.COLORMAP_JET)
            except ImportError:     # no colour map without OpenCV: grey levels in all three channels
                img = np.repeat(z_img[:, :, None], 3, axis=2)
            img[z_img == 0] = self.bg
            img[:, :, 1] = z_img
            return img

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            self._img_data = [self.z_image(i, pcd) for i, pcd in enumerate(pcds_data)]
            return pcds_data


class PointCloudMatProcessors:
    @staticmethod
    def dumps(pipes: List[PointCloudMatProcessor]):
        return json.dumps([p.model_dump(exclude={'input_mats', 'out_mats', 'meta'}) for p in pipes])

    @staticmethod
    def loads(pipes_json: str) -> List[PointCloudMatProcessor]:
        table = {k: v for k, v in Processors.__dict__.items() if '__' not in k}
        return [table[p["uuid"].split(":")[0]](**p) for p in json.loads(pipes_json)]

    @staticmethod
    def run_once(mats, meta={}, pipes: List[PointCloudMatProcessor] = [], validate=False):
        fn = None
        try:
            for fn in pipes:
                mats, meta = (fn.validate if validate else fn)(mats, meta)
        except Exception as e:
            logger(getattr(fn, 'uuid', '?'), e)
            raise
        return mats, meta

    @staticmethod
    def run(gen, pipes=[], meta={}, validate_once=False):
        if isinstance(pipes, str):
            pipes = PointCloudMatProcessors.loads(pipes)
        for mats in gen:
            PointCloudMatProcessors.run_once(mats, meta, pipes, validate_once)
            if validate_once:
                return

    @staticmethod
    def validate_once(gen, pipes=[]):
        PointCloudMatProcessors.run(gen, pipes, validate_once=True)

    @staticmethod
    def worker(pipes_serialized):
        pipes = PointCloudMatProcessors.loads(pipes_serialized)
        mats, meta = [], {}
        while True:
            for fn in pipes:
                mats, meta = fn(mats, meta)

    @staticmethod
    def run_async(pipes):
        ser = pipes if isinstance(pipes, str) else PointCloudMatProcessors.dumps(pipes)
        # spawn, not fork: a CUDA context does not survive fork
        p = multiprocessing.get_context("spawn").Process(target=PointCloudMatProcessors.worker, args=(ser,))
        p.start()
        return p
