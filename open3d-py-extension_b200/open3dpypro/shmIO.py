"""Named shared-memory frame transport between pipeline processes.

Same role and entry points as the reference's shmIO.py (`NumpyFloat32SharedMemoryStreamIO.reader /
.writer`, reference shmIO.py:281-296; fixed shape, fixed dtype, one segment per stream, name =
stream key with ':' replaced by '_'), re-implemented on a small common core.  Two differences that
matter to the device path: a writer publishes a sequence number after the payload so a reader can
detect a torn frame (the reference has no synchronisation at all, shmIO.py:229-240), and `read`
can copy straight into a caller-provided (e.g. page-locked) buffer.
"""
import json
from multiprocessing import shared_memory
from typing import Any, List, Optional

import numpy as np
from pydantic import BaseModel

_HDR = 64  # bytes: [0:8] sequence number (odd while a write is in flight)


class _ShmArray(BaseModel):
    shm_name: str
    array_shape: List[int]
    dtype_str: str = "float32"
    create: bool = False
    auto_del: bool = True
    pin: bool = False          # page-lock the segment (o3dpp_host_register): uploads from it are direct DMA
    _pinned: Any = None
    _shm: Any = None
    _arr: Any = None
    _seq: Any = None

    @property
    def shm_size(self):
        return int(np.prod(self.array_shape)) * np.dtype(self.dtype_str).itemsize + _HDR

    def build_buffer(self):
        if self._shm is not None:
            return self
        self._shm = shared_memory.SharedMemory(name=self.shm_name, create=self.create, size=self.shm_size)
        self._seq = np.ndarray((1,), np.uint64, buffer=self._shm.buf, offset=0)
        self._arr = np.ndarray(tuple(self.array_shape), np.dtype(self.dtype_str), buffer=self._shm.buf, offset=_HDR)
        if self.create:
            self._seq[0] = 0
        if self.pin and self._pinned is None:
            from . import _native
            whole = np.ndarray((self.shm_size,), np.uint8, buffer=self._shm.buf)
            _native.host_register(whole)
            self._pinned = whole
        return self

    def model_dump_json_dict(self):
        return json.loads(self.model_dump_json())

    def get_steam_info(self):
        return dict(shm_name=self.shm_name, array_shape=self.array_shape, dtype=self.dtype_str)

    def close(self):
        if self._pinned is not None:
            try:
                from . import _native
                _native.host_unregister(self._pinned)
            except Exception:
                pass
            self._pinned = None
        shm, self._arr, self._seq, self._shm = self._shm, None, None, None
        if shm is not None:
            try:
                shm.close()
                if self.create and self.auto_del:
                    shm.unlink()
            except Exception:
                pass

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _Reader(_ShmArray):
    def read(self, copy=True, out: Optional[np.ndarray] = None, retries=8):
        self.build_buffer()
        if not copy and out is None:
            return self._arr
        dst = np.empty_like(self._arr) if out is None else out
        for _ in range(retries):
            s0 = int(self._seq[0])
            np.copyto(dst, self._arr)
            if s0 % 2 == 0 and int(self._seq[0]) == s0:
                break
        return dst

    def write(self, data):
        raise ValueError("reader cannot write")


class _Writer(_ShmArray):
    def write(self, data):
        self.build_buffer()
        if tuple(data.shape) != tuple(self.array_shape):
            raise ValueError(f"shape mismatch: {tuple(data.shape)} vs {tuple(self.array_shape)}")
        self._seq[0] += 1
        np.copyto(self._arr, data, casting="unsafe")
        self._seq[0] += 1

    def read(self):
        raise ValueError("writer cannot read")


class GeneralNumpySharedMemoryIO:
    Reader, Writer = _Reader, _Writer
    DTYPE = "uint8"

    @classmethod
    def reader(cls, stream_key, array_shape, dtype=None, pin=False):
        return _Reader(shm_name=stream_key.replace(":", "_"), array_shape=list(array_shape),
                       dtype_str=np.dtype(dtype or cls.DTYPE).name, create=False, pin=pin).build_buffer()

    @classmethod
    def writer(cls, stream_key, array_shape, dtype=None, pin=False):
        return _Writer(shm_name=stream_key.replace(":", "_"), array_shape=list(array_shape),
                       dtype_str=np.dtype(dtype or cls.DTYPE).name, create=True, pin=pin).build_buffer()


class NumpyUInt8SharedMemoryStreamIO(GeneralNumpySharedMemoryIO):
    DTYPE = "uint8"


class NumpyFloat32SharedMemoryStreamIO(GeneralNumpySharedMemoryIO):
    DTYPE = "float32"
