"""Where a 131072-point frame spends its time in the block pipeline (config 4): per-call wall clock and the library's own stage timings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
from open3dpypro.processors import Processors, PointCloudMatProcessors
N = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
frame = lidar_like(N, 2000, np.float32)
c = nat.Context(0)
c.set_option(nat.OPT_STATS, 1)
def t(f, reps=20):
    f(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); ts.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(ts))
print("upload f32            %.3f ms" % t(lambda: c.upload(nat.ATTR_POINTS, frame)))
def normals():
    c.drop(nat.ATTR_NORMALS); c.estimate_normals(nat.SEARCH_KNN, 30, 0.0)
print("estimate_normals      %.3f ms" % t(normals, 5), "stages", [round(v, 3) for v in c.timings()])
c.set_option(nat.OPT_STATS, 0)
print("estimate_normals      %.3f ms (no stats)" % t(normals))
out6 = np.empty((N, 6)); print("download_rows6 f64    %.3f ms" % t(lambda: c.download_rows6(out6)))
print("segment_plane         %.3f ms" % t(lambda: c.segment_plane(0.05, 3, 1000, seed=7, fetch_inliers=False)), "stages", [round(v, 3) for v in c.timings()])
T = np.eye(4); T[:3, 3] = [0.1, 0.2, 0.3]
print("transform             %.3f ms" % t(lambda: c.transform(T)))
o32 = np.empty((N, 3), np.float32); print("download_as f32       %.3f ms" % t(lambda: c.download_as(nat.ATTR_POINTS, o32)))
os.environ["O3DPP_DEVICE"] = "0"
det = Processors.PlaneDetection(distance_threshold=0.05, num_iterations=1000, alpha=1.0, seed=7)
blocks = [Processors.CPUNormals(), det, Processors.PlaneNormalize(detection_uuid=det.uuid)]
mat = PointCloudMat(shape_type=ShapeType.XYZ).build(frame)
PointCloudMatProcessors.run_once([mat], {}, blocks, validate=True)
def run():
    mat.unsafe_update_mat(frame); PointCloudMatProcessors.run_once([mat], {}, blocks)
print("run_once (3 blocks)   %.3f ms" % t(run))
meta = {}
m1, meta = blocks[0]([mat], meta)
print("  CPUNormals          %.3f ms" % t(lambda: blocks[0]([mat], {})))
print("  PlaneDetection      %.3f ms" % t(lambda: blocks[1](m1, {})))
m2, meta = blocks[1](m1, meta)
print("  PlaneNormalize      %.3f ms" % t(lambda: blocks[2](m2, dict(meta))))
