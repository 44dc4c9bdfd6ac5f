"""Config 5: one large scan sharded across ranks by spatial slab, NCCL over NVLink.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P \
        tools/bench_slab.py --points 20000000 [--mode hybrid|knn] [--steps 2]

Every rank generates 1/G of the scan, then: slab_cuts (2 all-reduces) -> redistribute (all-to-all) ->
estimate_normals with halo exchange -> segment_plane with all-reduced scores.  Prints one JSON line with the
per-stage times (max over ranks; host numpy + collectives are inside the times) and Mpoints/s of the scan."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
import numpy as np
import torch
import torch.distributed as dist
from open3dpypro import _native as nat
from open3dpypro import sharding as sh
from open3dpypro.synthetic import lidar_like

ap = argparse.ArgumentParser()
ap.add_argument("--points", type=int, default=20_000_000)
ap.add_argument("--mode", default="hybrid")
ap.add_argument("--steps", type=int, default=2)
a = ap.parse_args()
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = nat.Context(local)
part = lidar_like(a.points // world, 1236 + rank).astype(np.float64)
mode = nat.SEARCH_HYBRID if a.mode == "hybrid" else nat.SEARCH_KNN

def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()

rows = []
for step in range(a.steps + 1):
    sync(); t0 = time.perf_counter()
    axis, cuts = sh.slab_cuts(part, world, 0.1)
    owned = sh.redistribute(part, axis, cuts, rank, world)
    sync(); t1 = time.perf_counter()
    nrm, info = sh.slab_estimate_normals(ctx, owned, axis, cuts, rank, world, mode, 30, 0.5)
    sync(); t2 = time.perf_counter()
    plane, inl, off, pinfo = sh.slab_segment_plane(ctx, owned, rank, world, 0.05, 3, 1000, seed=1, n_use=len(owned))
    sync(); t3 = time.perf_counter()
    if step:
        rows.append([t1 - t0, t2 - t1, t3 - t2, t3 - t0])
t = torch.tensor(rows, dtype=torch.float64, device=f"cuda:{local}").mean(0)
cnt = torch.tensor([len(owned), info["halo_points"]], dtype=torch.float64, device=f"cuda:{local}")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    mx = cnt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
else:
    mx = cnt
if rank == 0:
    t = t.tolist()
    print(json.dumps({"workload": f"C5-style slab-sharded scan, {a.points} points, estimate_normals({a.mode}, k=30, r=0.5) + segment_plane(0.05,3,1000)",
                      "n_gpus": world, "steps": a.steps, "cut_redistribute_s": t[0], "normals_s": t[1], "segment_s": t[2], "total_s": t[3],
                      "mpoints_per_s_normals_segment": a.points / (t[1] + t[2]) / 1e6, "max_owned_points": int(mx[0]), "max_halo_points": int(mx[1]),
                      "halo_width": info["halo_width"], "plane": [float(v) for v in plane], "n_inliers": pinfo["n_inliers"],
                      "note": "host numpy masks/concatenates and H2D/D2H of every stage are inside these times"}))
ctx.close()
if world > 1:
    dist.destroy_process_group()
