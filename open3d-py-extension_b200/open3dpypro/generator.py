"""Frame sources for the block pipeline (reference open3dpypro/generator.py:15-141): a generator
owns one output `PointCloudMat` per source and refreshes it on every `next()`."""
import time
import uuid as _uuid
from typing import Any, List, Literal

import numpy as np
from pydantic import BaseModel, ConfigDict

from .PointCloudMat import PointCloudMat, ShapeType


class PointCloudMatGenerator(BaseModel):
    model_config = ConfigDict(arbitrary_types_allowed=True)
    sources: List[Any]
    shape_types: List[ShapeType]
    uuid: str = ''
    shmIO_mode: Literal[False, 'writer', 'reader'] = False
    fps: int = -1
    output_mats: List[PointCloudMat] = []
    _frame_generators: List = []

    def model_post_init(self, context):
        self.uuid = f'{self.__class__.__name__}:{_uuid.uuid4()}'
        if not self.sources:
            raise ValueError("Empty sources.")
        if not self.shape_types:
            raise ValueError("Empty shape_types.")
        self._frame_generators = [self.create_frame_generator(i, s) for i, s in enumerate(self.sources)]
        if not self.output_mats:
            # like the reference (:44) the first frame of every source is consumed to size the mats
            self.output_mats = [PointCloudMat(shape_type=self.shape_types[i]).build(next(g))
                                for i, g in enumerate(self._frame_generators)]
        for mat in self.output_mats:
            if self.shmIO_mode == 'writer' and not mat.shmIO_writer:
                mat.build_shmIO_writer()

    def create_frame_generator(self, idx, source):
        raise NotImplementedError("Subclasses must implement `create_frame_generator`")

    def __iter__(self):
        return self

    def __next__(self):
        t0 = time.time()
        frames = [next(g) for g in self._frame_generators]   # StopIteration propagates
        for frame, mat in zip(frames, self.output_mats):
            mat.unsafe_update_mat(frame)
        if self.fps > 0:
            rest = 1.0 / self.fps - (time.time() - t0)
            if rest > 0:
                time.sleep(rest)
        return self.output_mats

    def reset_generators(self):
        self._frame_generators = [self.create_frame_generator(i, s) for i, s in enumerate(self.sources)]

    def release(self):
        for mat in self.output_mats:
            mat.release()


def _frames(arr, loop):
    k = 0
    while loop or k < len(arr):
        d = arr[k % len(arr)]
        yield np.ascontiguousarray(d[~np.isnan(d.sum(1))])   # drop NaN rows (reference :137-139)
        k += 1


class NumpyRawFrameFileGenerator(PointCloudMatGenerator):
    """sources: paths of (F,N,D) .npy recordings"""
    loop: bool = True

    def create_frame_generator(self, idx, source):
        return _frames(np.load(source), self.loop)


class ArrayFrameGenerator(PointCloudMatGenerator):
    """sources: in-memory (F,N,D) arrays (used by the benchmark's batched-frame config)"""
    loop: bool = False

    def create_frame_generator(self, idx, source):
        return _frames(np.asarray(source), self.loop)
