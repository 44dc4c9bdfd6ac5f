"""Config 4: a batch of LiDAR frames through the block pipeline, frame f -> rank f mod G.

    python tools/bench_frames.py [--frames 256] [--points 131072]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P tools/bench_frames.py

Every rank runs CPUNormals (estimate_normals k=30) -> PlaneDetection (segment_plane 0.05/3/1000) -> PlaneNormalize on
its frames through PointCloudMatProcessors.run_once, host float32 frames in, host arrays out (the reference's block
contract).  Prints one JSON line: frames/s and Mpoints/s of the whole job (max time over ranks)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=256)
ap.add_argument("--points", type=int, default=131072)
ap.add_argument("--distinct", type=int, default=8)
a = ap.parse_args()
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
os.environ["O3DPP_DEVICE"] = str(local)
dist = None
if world > 1:
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from open3dpypro import sharding as sh
from open3dpypro.PointCloudMat import PointCloudMat, ShapeType
from open3dpypro.processors import Processors, PointCloudMatProcessors
from open3dpypro.synthetic import lidar_like

mine = sh.frames_of_rank(a.frames, rank, world)
frames = {f: lidar_like(a.points, 2000 + f) for f in range(min(a.distinct, a.frames))}      # float32 (N,3)
det = Processors.PlaneDetection(distance_threshold=0.05, num_iterations=1000, alpha=1.0, seed=7)
pipes = [Processors.CPUNormals(pinned_outputs=True), det, Processors.PlaneNormalize(detection_uuid=det.uuid, pinned_outputs=True)]      # O3DPP_DEVICE pins this rank's GPU
mat = PointCloudMat(shape_type=ShapeType.XYZ).build(frames[0])
PointCloudMatProcessors.run_once([mat], {}, pipes, validate=True)

def barrier():
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()

planes = []
for rep in range(2):                      # first repetition = warm-up
    planes.clear()
    barrier(); t0 = time.perf_counter()
    for f in mine:
        mat.unsafe_update_mat(frames[f % len(frames)])
        out, meta = PointCloudMatProcessors.run_once([mat], {}, pipes)
        planes.append(meta[det.uuid][0])
    barrier(); t1 = time.perf_counter()
t = torch.tensor([t1 - t0], dtype=torch.float64, device=f"cuda:{local}")
if dist is not None:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
allp = sh.gather_frame_results(np.asarray(planes).reshape(-1, 4), a.frames, rank, world)
if rank == 0:
    sec = float(t[0])
    print(json.dumps({"workload": f"C4: {a.frames} frames x {a.points} points, CPUNormals -> PlaneDetection -> PlaneNormalize blocks, frame f -> rank f mod G",
                      "n_gpus": world, "seconds": sec, "frames_per_s": a.frames / sec, "mpoints_per_s": a.frames * a.points / sec / 1e6,
                      "ms_per_frame_per_gpu": 1e3 * sec / max(len(mine), 1), "planes_gathered": int(np.isfinite(allp).all(1).sum()),
                      "plane_frame0": [float(v) for v in allp[0]]}))
if dist is not None:
    dist.destroy_process_group()
