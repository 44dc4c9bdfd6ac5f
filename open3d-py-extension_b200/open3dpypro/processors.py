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
import weakref
from typing import List, Optional

import numpy as np

from . import _native as nat
from .PointCloudMat import PointCloudMat, PointCloudMatInfo, PointCloudMatProcessor, ShapeType

logger = print


try:  # torch tensors are accepted as payload (CUDA tensors stay on the device), never required
    import torch
except Exception:  # pragma: no cover
    torch = None

_RES_KEY = "_o3dpp_resident"      # meta[_RES_KEY][i] = what the previous block of this run_once left on the device for mat i


def _is_cuda_tensor(x):
    return torch is not None and isinstance(x, torch.Tensor) and x.is_cuda


def _xyz(pcd):
    """contiguous (N,3) view/copy of the position columns, dtype kept (f32 uploads as f32)"""
    a = pcd[:, :3]
    if a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float64)
    return np.ascontiguousarray(a)


class _OnDevice:
    """the device copy a block left behind: valid while the array object is the one handed on and nothing else has
    changed the points of that context (Context.generation)"""
    __slots__ = ("ref", "ctx", "generation", "normals_cols")

    def __init__(self, arr, ctx, normals_cols=False):
        self.ref, self.ctx, self.generation = weakref.ref(arr), ctx, ctx.generation
        self.normals_cols = normals_cols      # the device normals are arr[:, 3:6] (arr has exactly six columns)


def _require_array(pcd):
    """device blocks take host ndarrays or torch tensors (CUDA tensors stay on the device)"""
    if not (pcd.is_ndarray() or pcd.is_torch_tensor()):
        raise TypeError(f"Expected np.ndarray or torch.Tensor, got {type(pcd.data())}")


_MAT_POOL = (None, None)


def _mat_pool():
    """threads that work on the mats of one list side by side (library calls release the GIL); one pool per process:
    the threads of a parent's pool do not exist in a forked child (run_async)"""
    global _MAT_POOL
    if _MAT_POOL[0] != os.getpid():
        from concurrent.futures import ThreadPoolExecutor
        _MAT_POOL = (os.getpid(), ThreadPoolExecutor(max_workers=4, thread_name_prefix="o3dpp-mat"))
    return _MAT_POOL[1]


class _DeviceBlock(PointCloudMatProcessor):
    """common part of the device blocks: one library context per GPU, created lazily in the
    process that runs the block (so blocks survive `run_async`'s JSON round trip and fork).

    A frame is uploaded ONCE per run_once: the block that uploads (or changes) the points of mat i records it in
    `meta`, and the next block that receives the very same array object finds the cloud already resident, as the
    reference's blocks find the CUDA tensor NumpyToTorch made (reference processors.py:188-209, :651-663).  float32
    mats travel as float32 (12 B/point); CUDA tensor mats are read and written in place through the *_dev calls."""
    multi_gpu: int = -1
    pinned_outputs: bool = False     # outputs land in two alternating page-locked buffers per mat (full-speed D2H); a result
    _pinned: dict = {}               # is then valid until the block has run twice more, like a shm writer's segment
    parallel_mats: bool = True       # the host mats of one list are worked on side by side, one thread and one library context each
                                     # (the reference loops over them; the results do not depend on each other)

    def _each(self, pcds, fn, meta=None):
        """[fn(i, pcd) for i, pcd in enumerate(pcds)], the host mats concurrently when there are several"""
        if isinstance(meta, dict):
            meta.setdefault(_RES_KEY, {})                      # created once, before any thread records into it
        if len(pcds) < 2 or not self.parallel_mats or any(_is_cuda_tensor(p) for p in pcds):
            return [fn(i, p) for i, p in enumerate(pcds)]
        return list(_mat_pool().map(lambda ip: fn(*ip), list(enumerate(pcds))))

    def _out(self, i, shape, dtype):
        """host array for the output of mat i"""
        if not self.pinned_outputs:
            return np.empty(shape, dtype)
        key = (i, tuple(shape), np.dtype(dtype).str)
        ent = self._pinned.get(key)
        if ent is None:
            ent = self._pinned[key] = [[nat.PinnedArray(shape, dtype), nat.PinnedArray(shape, dtype)], 0]
        ent[1] ^= 1
        return ent[0][ent[1]].array

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

    def _slot(self, i):
        """which of a device's contexts serves mat i: mats that share a GPU get one each.  Slot 0 is the context the
        PointCloud / DeviceCloud API shares under its per-device lock; the blocks, which may run from several threads,
        stay off it"""
        pin = os.environ.get("O3DPP_DEVICE")
        if pin is not None and pin != "":
            return 1 + i
        return 1 + i // max(len(self.num_devices), 1)

    def _ctx(self, i, pcd=None):
        if _is_cuda_tensor(pcd):
            return nat.context(pcd.device.index, 1 + i)
        dev = self._device(i)
        return nat.context(dev, self._slot(i))

    def _resident(self, i, pcd, meta, want_handle=False):
        """-> context of mat i whose points are pcd[:, :3] (uploaded unless the previous block left them there);
        want_handle: -> (context, the record the previous block left, or None after an upload)"""
        ctx = self._ctx(i, pcd)
        h = meta.get(_RES_KEY, {}).get(i) if isinstance(meta, dict) else None
        if h is not None and h.ref() is pcd and h.ctx is ctx and h.generation == ctx.generation and ctx.size() == len(pcd):
            return (ctx, h) if want_handle else ctx
        if _is_cuda_tensor(pcd):
            t = pcd[:, :3]
            if t.dtype not in (torch.float32, torch.float64):
                t = t.double()
            t = t.contiguous()
            torch.cuda.current_stream(pcd.device).synchronize()     # the library reads on its own stream
            ctx.upload_dev(nat.ATTR_POINTS, t.data_ptr(), len(t), nat.F32 if t.dtype == torch.float32 else nat.F64)
        else:
            ctx.upload(nat.ATTR_POINTS, _xyz(pcd))
        ctx.drop(nat.ATTR_NORMALS)
        return (ctx, None) if want_handle else ctx

    @staticmethod
    def _leave(i, arr, ctx, meta, normals_cols=False):
        """record that ctx now holds arr[:, :3] as its points (normals_cols: and arr[:, 3:6] as its normals)"""
        if isinstance(meta, dict):
            meta.setdefault(_RES_KEY, {})[i] = _OnDevice(arr, ctx, normals_cols)


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
            _require_array(pcd)
            self.input_shape_types.append(pcd.shape_type)

        def build_out_mats(self, validated_pcds, converted_raw_pcds):
            self.out_mats = [PointCloudMat(shape_type=old.info.shape_type.add_normals()).build(p)
                             for old, p in zip(validated_pcds, converted_raw_pcds)]
            return self.out_mats

        def forward_raw(self, pcds_data: List[np.ndarray], pcds_info: List[PointCloudMatInfo] = [], meta={}):
            def one(i, pcd):
                if i < len(self.input_shape_types) and self.input_shape_types[i].contains_normals():
                    return None
                ctx = self._resident(i, pcd, meta)
                mode = nat.SEARCH_HYBRID if self.radius > 0 else nat.SEARCH_KNN
                ctx.drop(nat.ATTR_NORMALS)
                ctx.estimate_normals(mode, self.knn, self.radius)
                D = pcd.shape[1]
                if _is_cuda_tensor(pcd):                      # = torch.hstack([pcd, normals]) in float64, on the device
                    out = torch.empty((len(pcd), D + 3), dtype=torch.float64, device=pcd.device)
                    nrm = torch.empty((len(pcd), 3), dtype=torch.float64, device=pcd.device)
                    ctx.download_dev(nat.ATTR_NORMALS, nrm.data_ptr(), nat.F64)
                    out[:, :D] = pcd
                    out[:, D:] = nrm
                else:                                         # = np.hstack([pcd, normals]) (float64) without the temporaries
                    out = self._out(i, (len(pcd), D + 3), np.float64)
                    if D == 3:
                        ctx.download_rows6(out)               # rows assembled on the device, one copy
                    else:
                        out[:, :D] = pcd
                        out[:, D:] = ctx.download(nat.ATTR_NORMALS)
                self._leave(i, out, ctx, meta, normals_cols=(D == 3))
                return out
            return [o for o in self._each(pcds_data, one, meta) if o is not None]

    class VoxelDownsample(_DeviceBlock):
        title: str = 'voxel_sample'
        voxel_size: float = 0.1

        def validate_pcd(self, idx, pcd: PointCloudMat):
            _require_array(pcd)

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            def one(i, pcd):
                ctx = self._resident(i, pcd, meta)
                V = ctx.voxel(self.voxel_size)
                rep = ctx.voxel_fetch(V, len(pcd), ["rep"])["rep"]   # rows in ascending voxel key
                return pcd[torch.as_tensor(rep, device=pcd.device).long()] if _is_cuda_tensor(pcd) else pcd[rep]
            return self._each(pcds_data, one, meta)

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
            _require_array(pcd)
            self.best_planes.append([0., 0., 0., 0.])

        def detect(self, i, pcd, meta={}):
            if self.z_lower is not None and self.z_upper is not None:      # z band: a subset, uploaded on its own
                xyz = pcd[:, :3]
                xyz = xyz[(self.z_lower < xyz[:, 2]) & (xyz[:, 2] < self.z_upper)]
                ctx = self._resident(i, xyz, {})
            else:
                ctx, h = self._resident(i, pcd, meta, want_handle=True)
                self._leave(i, pcd, ctx, meta, normals_cols=bool(h and h.normals_cols))    # segment_plane leaves the cloud as it is
            seed = self.seed if self.seed is not None else int(np.random.randint(0, 2**31 - 1))
            plane, _ = ctx.segment_plane(self.distance_threshold, 3, self.num_iterations,
                                         seed=seed + self._calls, fetch_inliers=False)
            n, d = plane[:3], plane[3]
            if np.dot(n, np.zeros(3) - (-d * n)) < 0:   # normal towards the sensor origin (reference :643-649)
                plane = -plane
            return plane

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            if self.voxel_size > 0.0:
                pcds_data = Processors.VoxelDownsample(voxel_size=self.voxel_size, multi_gpu=self.multi_gpu).forward_raw(pcds_data, meta=meta)
            while len(self.best_planes) < len(pcds_data):
                self.best_planes.append([0., 0., 0., 0.])
            for i, plane in enumerate(self._each(pcds_data, lambda i, pcd: self.detect(i, pcd, meta), meta)):
                self.best_planes[i] = (np.asarray(self.best_planes[i]) * (1.0 - self.alpha) + plane * self.alpha).tolist()
            self._calls += 1
            meta[self.uuid] = self.best_planes
            return pcds_data

    class PlaneNormalize(_DeviceBlock):
        title: str = 'plane_normalize'
        detection_uuid: str
        filter_pcd: bool = False

        def validate_pcd(self, idx, pcd):
            _require_array(pcd)

        @staticmethod
        def rotation_matrix_from_vectors(vec1, vec2):
            """Rodrigues rotation vec1 -> vec2 in the dtype of the inputs (reference :709-723)"""
            dt = vec1.dtype
            a = vec1 / np.linalg.norm(vec1)
            b = vec2 / np.linalg.norm(vec2)
            v = np.cross(a, b)
            if np.linalg.norm(v) < 1e-6:
                return np.eye(3, dtype=dt)
            c = np.dot(a, b)
            s = np.linalg.norm(v)
            kmat = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=dt)
            return np.eye(3, dtype=dt) + kmat + np.matmul(kmat, kmat) * ((1 - c) / (s ** 2))

        @classmethod
        def plane_to_T(cls, plane, dtype=np.float64):
            """the 4x4 of rotate_to_plane (reference :725-739), built in the mat's dtype as the reference builds it"""
            a, b, c, d = plane
            normal = np.array([a, b, c], dtype=dtype)
            R = cls.rotation_matrix_from_vectors(normal, np.array([0, 0, 1], dtype=dtype))
            point_on_plane = -d * normal / np.dot(normal, normal)
            T = np.eye(4, dtype=dtype)
            T[:3, :3] = R
            T[:3, 3] = -np.matmul(R, point_on_plane)
            return T

        def forward_raw(self, pcds_data, pcds_info=[], meta={}):
            models = meta[self.detection_uuid]
            self.forward_T = [None] * len(pcds_data)

            def one(i, pcd):
                cuda = _is_cuda_tensor(pcd)
                dt = np.dtype(str(pcd.dtype).replace("torch.", "")) if cuda else pcd.dtype
                if dt not in (np.float32, np.float64):
                    dt = np.dtype(np.float64)
                T = self.plane_to_T(models[i], dt)
                ctx, h = self._resident(i, pcd, meta, want_handle=True)
                # T holds the mat-dtype values; the product runs in FP64 and is rounded once
                rows6 = not cuda and h is not None and h.normals_cols and pcd.shape[1] == 6 and ctx.has(nat.ATTR_NORMALS)
                if rows6:            # the other three columns are the normals the device still holds: rows assembled there, one copy
                    ctx.transform_points(T.astype(np.float64))
                    out = self._out(i, pcd.shape, dt)
                    ctx.download_rows6(out)
                    self._leave(i, out, ctx, meta, normals_cols=True)
                    self.forward_T[i] = T.tolist()
                    return out
                ctx.drop(nat.ATTR_NORMALS)
                ctx.transform(T.astype(np.float64))
                if cuda:
                    out = torch.empty_like(pcd)
                    xyz = out if pcd.shape[1] == 3 else torch.empty((len(pcd), 3), dtype=pcd.dtype, device=pcd.device)
                    ctx.download_dev(nat.ATTR_POINTS, xyz.data_ptr(), nat.F32 if dt == np.float32 else nat.F64)
                    if pcd.shape[1] > 3:
                        out[:, :3] = xyz
                        out[:, 3:] = pcd[:, 3:]
                else:
                    out = self._out(i, pcd.shape, dt)          # = np.hstack([xyz, pcd[:, 3:]]) in the mat's dtype
                    if pcd.shape[1] == 3:
                        ctx.download_as(nat.ATTR_POINTS, out)
                    else:
                        out[:, :3] = ctx.download_as(nat.ATTR_POINTS, np.empty((len(pcd), 3), dt))
                        out[:, 3:] = pcd[:, 3:]
                self._leave(i, out, ctx, meta)
                self.forward_T[i] = T.tolist()
                return out
            return self._each(pcds_data, one, meta)


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
