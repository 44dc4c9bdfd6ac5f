"""Typed (N,D) point-cloud carrier and the processor plug-in base class.

Mirrors the block-level boundary of the reference (open3dpypro/PointCloudMat.py): `ShapeType`
(:22-38), `PointCloudMatInfo` (:53-119), `PointCloudMat` (:121-267) and `PointCloudMatProcessor`
(:374-545: `validate`, `build_out_mats`, `forward_raw`, `forward`, JSON-able pydantic fields).  The
numpy/torch dispatch table (`MatOps`, :269-370) is not reproduced: blocks here call the device
library, there is one backend.
"""
import enum
import uuid as _uuid
from typing import Any, Dict, List, Literal, Optional, Tuple

import numpy as np
from pydantic import BaseModel, ConfigDict

from .shmIO import NumpyFloat32SharedMemoryStreamIO

try:  # torch tensors are accepted as payload, never required
    import torch
except Exception:  # pragma: no cover
    torch = None


class DeviceType(str, enum.Enum):
    CPU = 'cpu'
    GPU = 'gpu'


class ShapeType(str, enum.Enum):
    XYZ = 'XYZ'
    XYZRGB = 'XYZRGB'
    XYZi = 'XYZi'
    XYZiRGB = 'XYZiRGB'
    XYZRGBi = 'XYZRGBi'
    XYZN = 'XYZN'
    XYZRGBN = 'XYZRGBN'
    XYZiN = 'XYZiN'
    XYZRGBiN = 'XYZRGBiN'

    def contains_normals(self):
        return self.value.endswith('N')

    def add_normals(self):
        return ShapeType(self.value + 'N')

    def dims(self):
        v = self.value
        return 3 + (3 if 'RGB' in v else 0) + (1 if 'i' in v else 0) + (3 if v.endswith('N') else 0)


numpy_pcd_dtype = np.float32


def _is_tensor(x):
    return torch is not None and isinstance(x, torch.Tensor)


class PointCloudMatInfo(BaseModel):
    type: Optional[str] = None
    device: str = ''
    shape_type: Optional[ShapeType] = None
    raw_shape: List[int] = []
    N: int = 0
    uuid: str = ''
    _dtype: Any = None

    @staticmethod
    def numpy_pcd_dtype():
        return numpy_pcd_dtype

    def model_post_init(self, context):
        if not self.uuid:
            self.uuid = f'{self.__class__.__name__}:{_uuid.uuid4()}'

    def build(self, pcd_data):
        if isinstance(pcd_data, np.ndarray):
            self.device = 'cpu'
        elif _is_tensor(pcd_data):
            self.device = str(pcd_data.device)
        else:
            raise TypeError(f"pcd_data must be np.ndarray or torch.Tensor, got {type(pcd_data)}")
        if pcd_data.ndim != 2:
            raise ValueError(f"Point cloud data must be 2D (N, D). Got shape: {tuple(pcd_data.shape)}")
        if self.shape_type is None:
            raise ValueError("shape_type is required")
        want = ShapeType(self.shape_type).dims()
        if pcd_data.shape[1] != want:
            raise ValueError(f"Shape type '{ShapeType(self.shape_type).value}' expects feature dimension {want}, "
                             f"but got {pcd_data.shape[1]}. Full shape: {tuple(pcd_data.shape)}")
        self.type = type(pcd_data).__name__
        self._dtype = pcd_data.dtype
        self.N = int(pcd_data.shape[0])
        self.raw_shape = [int(s) for s in pcd_data.shape]
        return self


class PointCloudMat(BaseModel):
    model_config = ConfigDict(arbitrary_types_allowed=True)
    shape_type: ShapeType
    info: Optional[PointCloudMatInfo] = None
    shmIO_mode: Literal[False, 'writer', 'reader'] = False
    shmIO_writer: Optional[NumpyFloat32SharedMemoryStreamIO.Writer] = None
    shmIO_reader: Optional[NumpyFloat32SharedMemoryStreamIO.Reader] = None
    _pcd_data: Any = None

    @staticmethod
    def random(shape_type, num_points=1000):
        st = ShapeType(shape_type)
        return PointCloudMat(shape_type=st).build(np.random.rand(num_points, st.dims()).astype(numpy_pcd_dtype))

    def model_post_init(self, context):
        if self.shmIO_writer and self.shmIO_reader:
            raise ValueError('PointCloudMat cannot have both shmIO writer and reader.')
        for io in (self.shmIO_writer, self.shmIO_reader):
            if io:
                io.build_buffer()

    def build(self, pcd_data, info=None):
        self.info = info or PointCloudMatInfo(shape_type=self.shape_type).build(pcd_data)
        self._pcd_data = pcd_data
        return self

    def build_shmIO_writer(self):
        self.shmIO_mode = 'writer'
        self.shmIO_writer = NumpyFloat32SharedMemoryStreamIO.writer(self.info.uuid, self.info.raw_shape)
        return self

    def build_shmIO_reader(self, target_mat_info):
        self.shmIO_mode = 'reader'
        self.shmIO_reader = NumpyFloat32SharedMemoryStreamIO.reader(target_mat_info.uuid, target_mat_info.raw_shape)
        return self

    def to_shmIO_reader(self):
        w = self.shmIO_writer
        self.shmIO_mode = 'reader'
        self.shmIO_reader = NumpyFloat32SharedMemoryStreamIO.reader(w.shm_name, w.array_shape)
        self.shmIO_writer = None

    def to_shmIO_writer(self):
        r = self.shmIO_reader
        self.shmIO_mode = 'writer'
        self.shmIO_writer = NumpyFloat32SharedMemoryStreamIO.writer(r.shm_name, r.array_shape)
        self.shmIO_reader = None

    def release(self):
        for io in (self.shmIO_reader, self.shmIO_writer):
            if io:
                io.close()

    def copy(self):
        d = self._pcd_data
        return PointCloudMat(shape_type=self.info.shape_type).build(d.copy() if isinstance(d, np.ndarray) else d.clone())

    clone = copy

    def update_mat(self, pcd_data):
        self.info = PointCloudMatInfo(shape_type=self.info.shape_type).build(pcd_data)
        return self.unsafe_update_mat(pcd_data)

    def unsafe_update_mat(self, pcd_data):
        if self.shmIO_mode and self.shmIO_writer:
            self.shmIO_writer.write(pcd_data)
        self._pcd_data = pcd_data
        return self

    def data(self):
        if self.shmIO_mode and self.shmIO_reader:
            return self.shmIO_reader.read()
        return self._pcd_data

    def is_ndarray(self):
        return isinstance(self._pcd_data, np.ndarray)

    def is_torch_tensor(self):
        return _is_tensor(self._pcd_data)

    def require_ndarray(self):
        if not self.is_ndarray():
            raise TypeError(f"Expected np.ndarray, got {type(self._pcd_data)}")

    def require_torch_tensor(self):
        if not self.is_torch_tensor():
            raise TypeError(f"Expected torch.Tensor, got {type(self._pcd_data)}")

    def require_shape_type(self, shape_type):
        if self.info.shape_type != shape_type:
            raise TypeError(f"Expected shape type {ShapeType(shape_type).value}, got {self.info.shape_type.value}")


class PointCloudMatProcessor(BaseModel):
    """Plug-in base: subclasses set pydantic fields (their JSON config) and override
    `validate_pcd` and `forward_raw(pcds_data, pcds_info, meta) -> list`."""
    model_config = ConfigDict(arbitrary_types_allowed=True)
    title: str = ''
    uuid: str = ''
    save_results_to_meta: bool = False
    input_mats: List[PointCloudMat] = []
    out_mats: List[PointCloudMat] = []
    meta: dict = {}
    num_devices: List[str] = ['cpu']
    num_gpus: int = 0
    forward_T: List[List[List[float]]] = []
    _enable: bool = True

    def model_post_init(self, context):
        if not self.title:
            self.title = self.__class__.__name__
        if not self.uuid:
            self.uuid = f'{self.__class__.__name__}:{_uuid.uuid4()}'

    def print(self, *args):
        print(f'[{self.uuid}]', *args)

    def is_enable(self):
        return self._enable

    def on(self):
        self._enable = True

    def off(self):
        self._enable = False

    def devices_info(self, gpu=True, multi_gpu=-1):
        """cuda:i list for round-robin placement of mat i (reference :425-434)."""
        from . import _native
        self.num_devices, self.num_gpus = ['cpu'], 0
        if gpu:
            n = _native.device_count()
            if n > 0:
                use = n if multi_gpu <= 0 or multi_gpu > n else multi_gpu
                self.num_gpus = n
                self.num_devices = [f"cuda:{i}" for i in range(use)]
        return self.num_devices

    def validate_pcd(self, idx: int, pcd: PointCloudMat):
        raise NotImplementedError()

    def validate(self, pcds: List[PointCloudMat], meta: Dict = {}, run=True):
        mats = []
        for i, pcd in enumerate(pcds):
            self.validate_pcd(i, pcd)
            if pcd.shmIO_mode == 'writer':
                pcd = pcd.model_copy()
                pcd.to_shmIO_reader()
            elif pcd.shmIO_mode is not False:
                raise ValueError(f"pcd shmIO_mode must be False or writer. Got {pcd.shmIO_mode}")
            mats.append(pcd)
        self.input_mats = mats
        self.forward_T = [np.eye(4).tolist() for _ in mats]
        return self(self.input_mats, meta) if run else self.input_mats

    def build_out_mats(self, validated_pcds, converted_raw_pcds):
        ni, no = len(validated_pcds), len(converted_raw_pcds)
        if ni == no:
            types = [m.info.shape_type for m in validated_pcds]
        elif ni == 1:
            types = [validated_pcds[0].info.shape_type] * no
        elif no == 1:
            kinds = {m.info.shape_type for m in validated_pcds}
            if len(kinds) > 1:
                raise ValueError(f"Shape type mismatch: {kinds}. All PointCloudMats must have the same shape type.")
            types = [kinds.pop()]
        else:
            raise ValueError(f"[{self.uuid}] Length mismatch: {ni} vs {no}")
        self.out_mats = [PointCloudMat(shape_type=t).build(d) for t, d in zip(types, converted_raw_pcds)]
        return self.out_mats

    def forward_raw(self, pcds: List[Any], pcd_infos: List[PointCloudMatInfo] = [], meta={}) -> List[Any]:
        raise NotImplementedError()

    def forward(self, pcds: List[PointCloudMat], meta: Dict) -> Tuple[List[PointCloudMat], Dict]:
        if not (pcds and pcds[0].shmIO_mode is False) and self.input_mats:
            pcds = self.input_mats
        infos = [p.info for p in pcds]
        raw = [p.data() for p in pcds]
        if self._enable:
            raw = self.forward_raw(raw, infos, meta)
        if len(self.out_mats) == len(raw):
            outs = [m.unsafe_update_mat(d) for m, d in zip(self.out_mats, raw)]
        else:
            outs = self.build_out_mats(self.input_mats or pcds, raw)
            if any(p.shmIO_mode for p in pcds):
                for o in outs:
                    o.build_shmIO_writer()
        self.out_mats = outs
        if self.save_results_to_meta:
            meta[self.uuid] = self
        return outs, meta

    def __call__(self, pcds, meta={}):
        return self.forward(pcds, meta)

    def release(self):
        for m in list(self.input_mats) + list(self.out_mats):
            m.release()
