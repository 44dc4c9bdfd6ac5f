"""ctypes binding of libo3dpp_b200.so (include/o3dpp.h).

This is the only place the package touches native code.  There is no CPU fallback: if the
library is missing it is built with nvcc; if it cannot be loaded, or no CUDA device is usable,
the error is raised to the caller.
"""
import ctypes as C
import os
import threading

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.environ.get("O3DPP_LIB") or os.path.join(_ROOT, "libo3dpp_b200.so")   # O3DPP_LIB: tuning builds

ATTR_POINTS, ATTR_NORMALS, ATTR_COLORS = 0, 1, 2
SEARCH_KNN, SEARCH_RADIUS, SEARCH_HYBRID = 0, 1, 2
SAMPLER_PHILOX, SAMPLER_MT19937, SAMPLER_GIVEN = 0, 1, 2
F32, F64 = 0, 1
MAX_KNN = 128
(OPT_KEEP_NEIGHBORS, OPT_STATS, OPT_KNN_RHO, OPT_KNN_GSPLIT, OPT_KNN_MAXSHRINK, OPT_KNN_SHRINK, OPT_KNN_BITS,
 OPT_KNN_FORCE_LIST, OPT_QUERY_FIRST) = range(1, 10)

# every symbol include/o3dpp.h declares (tests check the library exports all of them)
SYMBOLS = [
    "o3dpp_last_error", "o3dpp_version", "o3dpp_device_count", "o3dpp_create", "o3dpp_destroy",
    "o3dpp_synchronize", "o3dpp_stream", "o3dpp_host_alloc", "o3dpp_host_free", "o3dpp_launch_count",
    "o3dpp_last_timings", "o3dpp_debug_counters", "o3dpp_upload", "o3dpp_upload_dev", "o3dpp_download", "o3dpp_download_dev",
    "o3dpp_size", "o3dpp_has", "o3dpp_drop", "o3dpp_bounds", "o3dpp_transform", "o3dpp_voxel",
    "o3dpp_voxel_fetch", "o3dpp_voxel_select", "o3dpp_estimate_normals", "o3dpp_knn",
    "o3dpp_segment_plane", "o3dpp_fetch_inliers", "o3dpp_fetch_samples", "o3dpp_select",
    "o3dpp_select_inliers", "o3dpp_sample_philox", "o3dpp_planes_from_points", "o3dpp_score_planes",
    "o3dpp_plane_inliers", "o3dpp_inlier_moments", "o3dpp_fetch_kth_bound", "o3dpp_knn_mean_distance", "o3dpp_zsplat", "o3dpp_download_async",
    "o3dpp_set_option", "o3dpp_fetch_neighbors", "o3dpp_voxel_bounded", "o3dpp_download_as", "o3dpp_download_rows6",
    "o3dpp_host_register", "o3dpp_host_unregister", "o3dpp_fp64_peak", "o3dpp_transform_points",
    "o3dpp_bounds_static", "o3dpp_fetch_kth_bound_dev",
]

_lib = None
_lock = threading.Lock()


class O3dppError(RuntimeError):
    """Raised for every non-zero status of the C ABI; `.code` is the O3DPP_E_* value.

    Open3D raises RuntimeError for the same conditions (bad voxel_size, ransac_n < 3, ...), so
    code written against the reference keeps catching it."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def lib():
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    import importlib.util
                    spec = importlib.util.spec_from_file_location("_o3dpp_build", os.path.join(_ROOT, "build.py"))
                    mod = importlib.util.module_from_spec(spec)
                    spec.loader.exec_module(mod)
                    mod.build()
                L = C.CDLL(LIB_PATH)
                L.o3dpp_last_error.restype = C.c_char_p
                L.o3dpp_version.restype = C.c_char_p
                _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        msg = lib().o3dpp_last_error()
        raise O3dppError(rc, (msg or b"").decode("utf-8", "replace") or f"o3dpp error {rc}")


def device_count():
    n = C.c_int(0)
    _check(lib().o3dpp_device_count(C.byref(n)))
    return n.value


BOUNDS_STATIC_MAX_POINTS = 1 << 26      # O3DPP_BOUNDS_STATIC_MAX_POINTS of include/o3dpp.h


def bounds_static(device, dev_rows, n, dev_out6, stream=0):
    """o3dpp_bounds_static: min / max bound of n x 3 float64 device rows into 6 device doubles without any allocation
    (static device scratch, n <= BOUNDS_STATIC_MAX_POINTS); enqueues on `stream` and returns"""
    _check(lib().o3dpp_bounds_static(C.c_int(device), C.c_size_t(stream), C.c_size_t(dev_rows), C.c_int64(n), C.c_size_t(dev_out6)))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PinnedArray:
    """numpy array over cudaHostAlloc memory (page-locked, so H2D/D2H run at full PCIe speed)."""

    def __init__(self, shape, dtype=np.float64):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._ptr = C.c_void_p()
        _check(lib().o3dpp_host_alloc(C.byref(self._ptr), C.c_uint64(max(nbytes, 1))))
        buf = (C.c_char * max(nbytes, 1)).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._ptr is not None and self._ptr.value:
            self.array = None
            try:
                lib().o3dpp_host_free(self._ptr)
            except Exception:
                pass
            self._ptr = None

    def __del__(self):
        self.free()


def host_register(arr):
    """page-lock the memory behind a numpy array the caller owns (e.g. a shared-memory frame segment)"""
    _check(lib().o3dpp_host_register(C.c_void_p(arr.ctypes.data), C.c_uint64(arr.nbytes)))


def host_unregister(arr):
    _check(lib().o3dpp_host_unregister(C.c_void_p(arr.ctypes.data)))


class Context:
    """One device-resident cloud on one GPU (o3dpp_ctx).  Created lazily, per process.

    `generation` counts the changes of the device points (upload, transform, selection): the pipeline blocks use it
    to tell whether the cloud they left on the device is still the one they find (processors._DeviceBlock)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        self.device = device
        self._pid = os.getpid()
        self.generation = 0
        self.n_inliers = 0
        _check(lib().o3dpp_create(C.c_int(device), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value and self._pid == os.getpid():
            lib().o3dpp_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- state
    def upload(self, attr, arr):
        a = np.asarray(arr)
        if a.dtype == np.float32:
            dt = F32
        else:
            dt = F64
            if a.dtype != np.float64:
                a = a.astype(np.float64)
        if a.ndim != 2 or a.shape[1] != 3:
            raise ValueError("expected an (n,3) array")
        a = np.ascontiguousarray(a)
        if attr == ATTR_POINTS:
            self.generation += 1
        _check(lib().o3dpp_upload(self._h, C.c_int(attr), _ptr(a), C.c_int64(a.shape[0]), C.c_int(dt)))

    def upload_dev(self, attr, dev_ptr, n, dtype=F64):
        if attr == ATTR_POINTS:
            self.generation += 1
        _check(lib().o3dpp_upload_dev(self._h, C.c_int(attr), C.c_size_t(dev_ptr), C.c_int64(n), C.c_int(dtype)))

    def download_as(self, attr, out):
        """(n,3) float32 or float64 C-contiguous destination"""
        n = self.size()
        assert out.dtype in (np.float32, np.float64) and out.flags.c_contiguous and out.shape == (n, 3)
        if n:
            _check(lib().o3dpp_download_as(self._h, C.c_int(attr), _ptr(out), C.c_int(F32 if out.dtype == np.float32 else F64)))
        return out

    def download_rows6(self, out):
        """(n,6) rows [x y z nx ny nz], float32 or float64"""
        n = self.size()
        assert out.dtype in (np.float32, np.float64) and out.flags.c_contiguous and out.shape == (n, 6)
        if n:
            _check(lib().o3dpp_download_rows6(self._h, _ptr(out), C.c_int(F32 if out.dtype == np.float32 else F64)))
        return out

    def download(self, attr, out=None):
        n = self.size()
        if out is None:
            out = np.empty((n, 3), np.float64)
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape == (n, 3)
        if n:
            _check(lib().o3dpp_download(self._h, C.c_int(attr), _ptr(out)))
        return out

    def download_async(self, attr, out):
        """starts the copy and returns; `out` is complete after synchronize()"""
        n = self.size()
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape == (n, 3)
        if n:
            _check(lib().o3dpp_download_async(self._h, C.c_int(attr), _ptr(out)))
        return out

    def download_dev(self, attr, dev_ptr, dtype=F64):
        _check(lib().o3dpp_download_dev(self._h, C.c_int(attr), C.c_size_t(dev_ptr), C.c_int(dtype)))

    def size(self):
        n = C.c_int64(0)
        _check(lib().o3dpp_size(self._h, C.byref(n)))
        return n.value

    def has(self, attr):
        f = C.c_int(0)
        _check(lib().o3dpp_has(self._h, C.c_int(attr), C.byref(f)))
        return bool(f.value)

    def drop(self, attr):
        if attr == ATTR_POINTS:
            self.generation += 1
        _check(lib().o3dpp_drop(self._h, C.c_int(attr)))

    def synchronize(self):
        _check(lib().o3dpp_synchronize(self._h))

    def stream(self):
        s = C.c_size_t(0)
        _check(lib().o3dpp_stream(self._h, C.byref(s)))
        return s.value

    def launch_count(self):
        n = C.c_uint64(0)
        _check(lib().o3dpp_launch_count(self._h, C.byref(n)))
        return n.value

    def timings(self):
        t = np.zeros(8, np.float64)
        _check(lib().o3dpp_last_timings(self._h, _ptr(t)))
        return t

    def debug_counters(self):
        t = np.zeros(8, np.uint64)
        _check(lib().o3dpp_debug_counters(self._h, _ptr(t)))
        return t

    def bounds(self):
        mn, mx = np.zeros(3), np.zeros(3)
        _check(lib().o3dpp_bounds(self._h, _ptr(mn), _ptr(mx)))
        return mn, mx

    # ---- operations
    def transform_points(self, T):
        """points only; normals (if any) stay as they are"""
        T = np.ascontiguousarray(np.asarray(T, dtype=np.float64).reshape(16))
        self.generation += 1
        _check(lib().o3dpp_transform_points(self._h, _ptr(T)))

    def transform(self, T):
        T = np.ascontiguousarray(np.asarray(T, dtype=np.float64).reshape(16))
        self.generation += 1
        _check(lib().o3dpp_transform(self._h, _ptr(T)))

    def voxel(self, voxel_size, min_bound=None, max_bound=None):
        """min_bound / max_bound: explicit grid origin / far corner (None: the cloud's own bounds)"""
        V = C.c_int64(0)
        if min_bound is None and max_bound is None:
            _check(lib().o3dpp_voxel(self._h, C.c_double(voxel_size), C.byref(V)))
        else:
            mn = None if min_bound is None else np.ascontiguousarray(np.asarray(min_bound, dtype=np.float64).reshape(3))
            mx = None if max_bound is None else np.ascontiguousarray(np.asarray(max_bound, dtype=np.float64).reshape(3))
            _check(lib().o3dpp_voxel_bounded(self._h, C.c_double(voxel_size), _ptr(mn), _ptr(mx), C.byref(V)))
        return V.value

    def voxel_fetch(self, V, n, fields=("rep_asc",)):
        """fields: any of rep_asc, rep, count, centroid, normal, color, cubic_id, order, seg"""
        spec = {
            "rep_asc": ((V,), np.int32), "rep": ((V,), np.int32), "count": ((V,), np.int32),
            "centroid": ((V, 3), np.float64), "normal": ((V, 3), np.float64), "color": ((V, 3), np.float64),
            "cubic_id": ((V, 8), np.int32), "order": ((n,), np.int32), "seg": ((V + 1,), np.int64),
        }
        out = {k: np.empty(spec[k][0], spec[k][1]) for k in fields}
        args = [_ptr(out.get(k)) for k in ("rep_asc", "rep", "count", "centroid", "normal", "color", "cubic_id", "order", "seg")]
        _check(lib().o3dpp_voxel_fetch(self._h, *args))
        return out

    def voxel_select(self):
        self.generation += 1
        _check(lib().o3dpp_voxel_select(self._h))

    def estimate_normals(self, mode, knn, radius):
        _check(lib().o3dpp_estimate_normals(self._h, C.c_int(mode), C.c_int(knn), C.c_double(radius)))

    def set_option(self, option, value):
        _check(lib().o3dpp_set_option(self._h, C.c_int(option), C.c_double(float(value))))

    def fetch_neighbors(self, knn, queries=None):
        """neighbour sets of the last estimate_normals (OPT_KEEP_NEIGHBORS set before it): (idx (m,knn) int32 ascending
        original indices, -1 padded; cnt (m,) int32; tier (m,) int8 = which kernel of the chain answered)"""
        if queries is None:
            m, q = self.size(), None
        else:
            q = np.ascontiguousarray(np.asarray(queries, dtype=np.int64).reshape(-1))
            m = q.size
        idx = np.empty((m, knn), np.int32)
        cnt = np.empty(m, np.int32)
        tier = np.empty(m, np.int8)
        if m:
            _check(lib().o3dpp_fetch_neighbors(self._h, _ptr(q), C.c_int64(m), _ptr(idx), _ptr(cnt), _ptr(tier)))
        return idx, cnt, tier

    def knn(self, mode, knn, radius=0.0, want_d2=True):
        n = self.size()
        idx = np.empty((n, knn), np.int32)
        d2 = np.empty((n, knn), np.float64) if want_d2 else None
        cnt = np.empty(n, np.int32)
        _check(lib().o3dpp_knn(self._h, C.c_int(mode), C.c_int(knn), C.c_double(radius), _ptr(idx), _ptr(d2), _ptr(cnt)))
        return idx, d2, cnt

    def segment_plane(self, distance_threshold, ransac_n, num_iterations, probability=0.99999999,
                      sampler=SAMPLER_PHILOX, seed=0, samples=None, debug=False, fetch_inliers=True, inliers_out=None):
        """inliers_out: optional caller-owned int32 buffer of >= n entries (e.g. page-locked) that receives the inlier list;
        the returned array is then a view of its first n_inliers entries."""
        H = int(num_iterations)
        plane = np.zeros(4)
        best = np.zeros(4)
        m = C.c_int64(0)
        bi = C.c_int32(-1)
        counts = np.empty(max(H, 1), np.int64) if debug else None
        errs = np.empty(max(H, 1), np.float64) if debug else None
        s = None
        if samples is not None:
            s = np.ascontiguousarray(np.asarray(samples, dtype=np.int64).reshape(-1))
            if s.size != H * int(ransac_n):
                raise ValueError("samples must hold num_iterations*ransac_n indices")
            sampler = SAMPLER_GIVEN
        _check(lib().o3dpp_segment_plane(self._h, C.c_double(distance_threshold), C.c_int(ransac_n), C.c_int(H),
                                         C.c_double(probability), C.c_int(sampler), C.c_uint64(seed), _ptr(s),
                                         _ptr(plane), C.byref(m), _ptr(best), C.byref(bi), _ptr(counts), _ptr(errs)))
        inl = None
        self.n_inliers = m.value                     # for fetch_inliers_into after a call with fetch_inliers=False
        if fetch_inliers:
            if inliers_out is not None:
                if inliers_out.dtype != np.int32 or not inliers_out.flags.c_contiguous or inliers_out.size < m.value:
                    raise ValueError("inliers_out must be a contiguous int32 array with room for every inlier")
                inl = inliers_out.reshape(-1)[:m.value]
            else:
                inl = np.empty(m.value, np.int32)
            if m.value:
                _check(lib().o3dpp_fetch_inliers(self._h, _ptr(inl)))
        if debug:
            return plane, inl, dict(best_plane=best, best_iter=bi.value, counts=counts[:H], errs=errs[:H],
                                    n_inliers=m.value)
        return plane, inl

    def fetch_samples(self, H, ransac_n):
        out = np.empty((H, ransac_n), np.int64)
        if H:
            _check(lib().o3dpp_fetch_samples(self._h, _ptr(out)))
        return out

    def select(self, idx, invert=False):
        idx = np.ascontiguousarray(np.asarray(idx, dtype=np.int32).reshape(-1))
        self.generation += 1
        _check(lib().o3dpp_select(self._h, _ptr(idx), C.c_int64(idx.size), C.c_int(int(invert))))

    def select_inliers(self, invert=False):
        self.generation += 1
        _check(lib().o3dpp_select_inliers(self._h, C.c_int(int(invert))))

    # ---- the steps of segment_plane for slab-sharded scans (sharding.py)
    def sample_philox(self, seed, n, ransac_n, num_iterations):
        out = np.empty((int(num_iterations), int(ransac_n)), np.int64)
        if out.size:
            _check(lib().o3dpp_sample_philox(self._h, C.c_uint64(seed), C.c_int64(n), C.c_int(ransac_n),
                                             C.c_int(num_iterations), _ptr(out)))
        return out

    def planes_from_points(self, pts, ransac_n):
        pts = np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, int(ransac_n), 3))
        planes = np.zeros((len(pts), 4))
        if len(pts):
            _check(lib().o3dpp_planes_from_points(self._h, _ptr(pts), C.c_int(ransac_n), C.c_int(len(pts)), _ptr(planes)))
        return planes

    def score_planes(self, planes, distance_threshold, n_use=0):
        planes = np.ascontiguousarray(np.asarray(planes, dtype=np.float64).reshape(-1, 4))
        counts = np.zeros(len(planes), np.int64)
        errs = np.zeros(len(planes), np.float64)
        if len(planes) and self.size():
            _check(lib().o3dpp_score_planes(self._h, _ptr(planes), C.c_int(len(planes)), C.c_double(distance_threshold),
                                            C.c_int64(n_use), _ptr(counts), _ptr(errs)))
        return counts, errs

    def plane_inliers(self, plane, distance_threshold, n_use=0):
        """inlier list of `plane` (kept on the device: fetch_inlier_list / select_inliers); returns (count, coordinate sums)"""
        plane = np.ascontiguousarray(np.asarray(plane, dtype=np.float64).reshape(4))
        m = C.c_int64(0)
        s3 = np.zeros(3)
        _check(lib().o3dpp_plane_inliers(self._h, _ptr(plane), C.c_double(distance_threshold), C.c_int64(n_use), C.byref(m), _ptr(s3)))
        return m.value, s3

    def fetch_inliers_into(self, out):
        """the inlier list of the last segment_plane (fetch_inliers=False) into a caller-owned int32 buffer (e.g. page-locked);
        returns the view of its first n_inliers entries"""
        m = self.n_inliers
        if out.dtype != np.int32 or not out.flags.c_contiguous or out.size < m:
            raise ValueError("out must be a contiguous int32 array with room for every inlier")
        inl = out.reshape(-1)[:m]
        if m:
            _check(lib().o3dpp_fetch_inliers(self._h, _ptr(inl)))
        return inl

    def fetch_inlier_list(self, m):
        inl = np.empty(int(m), np.int32)
        if m:
            _check(lib().o3dpp_fetch_inliers(self._h, _ptr(inl)))
        return inl

    def inlier_moments(self, centroid):
        cen = np.ascontiguousarray(np.asarray(centroid, dtype=np.float64).reshape(3))
        mom = np.zeros(6)
        _check(lib().o3dpp_inlier_moments(self._h, _ptr(cen), _ptr(mom)))
        return mom

    def knn_mean_distance(self, knn):
        out = np.empty(self.size(), np.float64)
        if out.size:
            _check(lib().o3dpp_knn_mean_distance(self._h, C.c_int(knn), _ptr(out)))
        return out

    def zsplat(self, grid_size, x_min, x_max, y_min, y_max):
        """-> (float32 (grid, grid) z of the last point per cell, NaN where empty; int32 index of that point, -1 where empty)"""
        g = int(grid_size)
        z = np.empty((g, g), np.float32)
        last = np.empty((g, g), np.int32)
        _check(lib().o3dpp_zsplat(self._h, C.c_int(g), C.c_double(x_min), C.c_double(x_max), C.c_double(y_min), C.c_double(y_max),
                                  _ptr(z), _ptr(last)))
        return z, last

    def fp64_peak(self):
        """measured FP64 FMA throughput of the device, TFLOP/s"""
        v = C.c_double(0.0)
        _check(lib().o3dpp_fp64_peak(self._h, C.byref(v)))
        return v.value

    def kth_bound(self):
        """per point: upper bound of the squared distance to its k-th neighbour (last estimate_normals)"""
        out = np.empty(self.size(), np.float32)
        if out.size:
            _check(lib().o3dpp_fetch_kth_bound(self._h, _ptr(out)))
        return out

    def kth_bound_dev(self, dev_ptr):
        """the same bounds into size() float32 of device memory"""
        if self.size():
            _check(lib().o3dpp_fetch_kth_bound_dev(self._h, C.c_size_t(dev_ptr)))


_contexts = {}
_ctx_locks = {}
_ctx_create_lock = threading.RLock()


def context_lock(device=0):
    """re-entrant lock that serialises the calls of one process on the shared context of a device"""
    key = (os.getpid(), int(device))
    lk = _ctx_locks.get(key)
    if lk is None:
        with _lock:
            lk = _ctx_locks.setdefault(key, threading.RLock())
    return lk


def context(device=0, slot=0):
    """Process-wide shared context per device, re-created after fork (CUDA state does not survive it).
    slot > 0: further contexts on the same device (the blocks give every mat of a list its own, so that several mats on one
    GPU stay resident side by side and can be worked on from several threads; one context = one cloud, one stream)."""
    key = (os.getpid(), int(device), int(slot))
    ctx = _contexts.get(key)
    if ctx is None:
        lib()                                   # loads the library under _lock: must not happen inside the lock below
        with _ctx_create_lock:
            ctx = _contexts.get(key)
            if ctx is None:
                ctx = Context(device)
                _contexts[key] = ctx
    return ctx
