"""ctypes front-end of the CPU oracle (oracle/o3d_oracle.c).

TEST INFRASTRUCTURE ONLY — may be imported from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs, never from the product
package.  PARITY UNPINNED: see the header of o3d_oracle.c.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liborc.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.orc_voxel_index.restype = C.c_int64
        _LIB.orc_mt19937_nth.restype = C.c_uint32
    return _LIB


def _pts(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    assert a.ndim == 2 and a.shape[1] == 3
    return a


def _opt(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def transform(pts, T, normals=None):
    p = _pts(pts).copy()
    nrm = None if normals is None else _pts(normals).copy()
    T = np.ascontiguousarray(T, dtype=np.float64).reshape(16)
    lib().orc_transform(p.ctypes.data_as(C.c_void_p), _opt(nrm), C.c_int64(len(p)), T.ctypes.data_as(C.c_void_p))
    return p, nrm


def voxel_trace(pts, voxel_size, normals=None, colors=None, min_bound=None, max_bound=None):
    """dict with rows in ascending voxel key (ix,iy,iz).  min_bound / max_bound: the explicit grid origin / far corner of
    voxel_down_sample_and_trace(voxel_size, min_bound, max_bound) (reference PointCloud.py:339-340); None = own bounds."""
    p = _pts(pts)
    n = len(p)
    vidx = np.empty((n, 3), np.int32)
    octant = np.empty(n, np.int32)
    order = np.empty(n, np.int32)
    seg = np.empty(n + 1, np.int64)
    mnb = None if min_bound is None else np.ascontiguousarray(np.asarray(min_bound, dtype=np.float64).reshape(3))
    mxb = None if max_bound is None else np.ascontiguousarray(np.asarray(max_bound, dtype=np.float64).reshape(3))
    f = lib().orc_voxel_index_bounded
    f.restype = C.c_int64
    V = f(p.ctypes.data_as(C.c_void_p), C.c_int64(n), C.c_double(voxel_size), _opt(mnb), _opt(mxb),
          vidx.ctypes.data_as(C.c_void_p), octant.ctypes.data_as(C.c_void_p),
          order.ctypes.data_as(C.c_void_p), seg.ctypes.data_as(C.c_void_p))
    if V == -1:
        raise RuntimeError("voxel_size <= 0.")
    if V == -2:
        raise RuntimeError("voxel_size is too small.")
    seg = seg[:V + 1].copy()
    nrm = None if normals is None else _pts(normals)
    col = None if colors is None else _pts(colors)
    cubic = np.empty((V, 8), np.int32)
    rep = np.empty(V, np.int32)
    cnt = np.empty(V, np.int32)
    cen = np.empty((V, 3), np.float64)
    on = np.empty((V, 3), np.float64) if nrm is not None else None
    oc = np.empty((V, 3), np.float64) if col is not None else None
    lib().orc_voxel_reduce(p.ctypes.data_as(C.c_void_p), _opt(nrm), _opt(col), C.c_int64(V),
                           order.ctypes.data_as(C.c_void_p), seg.ctypes.data_as(C.c_void_p),
                           octant.ctypes.data_as(C.c_void_p), cubic.ctypes.data_as(C.c_void_p),
                           rep.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p),
                           cen.ctypes.data_as(C.c_void_p), _opt(on), _opt(oc))
    keys = vidx[order[seg[:-1]]] if V else np.empty((0, 3), np.int32)
    return dict(V=int(V), vidx=vidx, octant=octant, order=order, seg=seg, keys=keys, cubic_id=cubic,
                rep=rep, count=cnt, centroid=cen, normal=on, color=oc)


def knn(pts, k, queries=None, mode=0, radius=0.0, brute=False, nthreads=0):
    p = _pts(pts)
    q = p if queries is None else _pts(queries)
    idx = np.empty((len(q), k), np.int32)
    d2 = np.empty((len(q), k), np.float64)
    cnt = np.empty(len(q), np.int32)
    rc = lib().orc_knn(p.ctypes.data_as(C.c_void_p), C.c_int64(len(p)), q.ctypes.data_as(C.c_void_p),
                       C.c_int64(len(q)), C.c_int(mode), C.c_int(k), C.c_double(radius), C.c_int(int(brute)),
                       idx.ctypes.data_as(C.c_void_p), d2.ctypes.data_as(C.c_void_p),
                       cnt.ctypes.data_as(C.c_void_p), C.c_int(nthreads))
    if rc:
        raise RuntimeError(f"orc_knn rc={rc}")
    return idx, d2, cnt


def fast_eigen3x3(C6):
    C6 = np.ascontiguousarray(C6, dtype=np.float64).reshape(-1, 6)
    out = np.empty((len(C6), 3), np.float64)
    f = lib().orc_fast_eigen3x3
    for i in range(len(C6)):
        f(C6[i].ctypes.data_as(C.c_void_p), out[i].ctypes.data_as(C.c_void_p))
    return out


MODES = {"knn": 0, "radius": 1, "hybrid": 2}


def estimate_normals(pts, mode="knn", k=30, radius=0.0, old_normals=None, want_cov=False, nthreads=0):
    p = _pts(pts)
    out = np.empty_like(p)
    cov = np.empty((len(p), 6), np.float64) if want_cov else None
    old = None if old_normals is None else _pts(old_normals)
    rc = lib().orc_estimate_normals(p.ctypes.data_as(C.c_void_p), C.c_int64(len(p)), C.c_int(MODES[mode]),
                                    C.c_int(k), C.c_double(radius), _opt(old), out.ctypes.data_as(C.c_void_p),
                                    _opt(cov), C.c_int(nthreads))
    if rc:
        raise RuntimeError(f"orc_estimate_normals rc={rc}")
    return (out, cov) if want_cov else out


def sample_mt19937(seed, n, ransac_n, H):
    out = np.empty((H, ransac_n), np.int64)
    if lib().orc_sample_mt19937(C.c_uint32(seed), C.c_int64(n), C.c_int(ransac_n), C.c_int(H),
                                out.ctypes.data_as(C.c_void_p)):
        raise RuntimeError("There must be at least 'ransac_n' points.")
    return out


def sample_philox(seed, n, ransac_n, H):
    out = np.empty((H, ransac_n), np.int64)
    if lib().orc_sample_philox(C.c_uint64(seed), C.c_int64(n), C.c_int(ransac_n), C.c_int(H),
                               out.ctypes.data_as(C.c_void_p)):
        raise RuntimeError("There must be at least 'ransac_n' points.")
    return out


def philox(seed, c0, c1, c2, c3):
    out = np.empty(4, np.uint32)
    lib().orc_philox(C.c_uint64(seed), C.c_uint32(c0), C.c_uint32(c1), C.c_uint32(c2), C.c_uint32(c3),
                     out.ctypes.data_as(C.c_void_p))
    return out


def mt19937_nth(seed, nth):
    return int(lib().orc_mt19937_nth(C.c_uint32(seed), C.c_int(nth)))


def plane_model(pts, sample):
    p = _pts(pts)
    s = np.ascontiguousarray(sample, dtype=np.int64)
    out = np.empty(4, np.float64)
    lib().orc_plane_model(p.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p), C.c_int(len(s)),
                          out.ctypes.data_as(C.c_void_p))
    return out


def score_plane(pts, plane, thr):
    p = _pts(pts)
    pl = np.ascontiguousarray(plane, dtype=np.float64)
    c = C.c_int64(0)
    e = C.c_double(0)
    lib().orc_score_plane(p.ctypes.data_as(C.c_void_p), C.c_int64(len(p)), pl.ctypes.data_as(C.c_void_p),
                          C.c_double(thr), C.byref(c), C.byref(e))
    return c.value, e.value


def segment_plane(pts, thr, ransac_n, H, samples, prob=0.99999999, nthreads=1, debug=False):
    p = _pts(pts)
    s = np.ascontiguousarray(samples, dtype=np.int64).reshape(H, -1) if H else np.empty((0, max(ransac_n, 1)), np.int64)
    plane = np.zeros(4, np.float64)
    best = np.zeros(4, np.float64)
    inl = np.empty(max(len(p), 1), np.int32)
    m = C.c_int64(0)
    bi = C.c_int32(-1)
    counts = np.empty(max(H, 1), np.int64)
    errs = np.empty(max(H, 1), np.float64)
    rc = lib().orc_segment_plane(p.ctypes.data_as(C.c_void_p), C.c_int64(len(p)), C.c_double(thr), C.c_int(ransac_n),
                                 C.c_int(H), C.c_double(prob), s.ctypes.data_as(C.c_void_p),
                                 plane.ctypes.data_as(C.c_void_p), inl.ctypes.data_as(C.c_void_p), C.byref(m),
                                 best.ctypes.data_as(C.c_void_p), C.byref(bi), counts.ctypes.data_as(C.c_void_p),
                                 errs.ctypes.data_as(C.c_void_p), C.c_int(nthreads))
    if rc == -3:
        raise RuntimeError("ransac_n should be set to higher than or equal to 3.")
    if rc == -4:
        raise RuntimeError("There must be at least 'ransac_n' points.")
    if rc:
        raise RuntimeError(f"orc_segment_plane rc={rc}")
    inl = inl[:m.value].copy()
    if debug:
        return plane, inl, dict(best_plane=best, best_iter=bi.value, counts=counts[:H], errs=errs[:H])
    return plane, inl


def num_threads():
    return int(lib().orc_num_threads())
