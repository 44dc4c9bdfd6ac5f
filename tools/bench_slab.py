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
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"      # NCCL's version banner goes to stdout: keep the JSON line alone there
import torch
import torch.distributed as dist
from open3dpypro import _native as nat
from open3dpypro import sharding as sh
from open3dpypro.synthetic import lidar_like

ap = argparse.ArgumentParser()
ap.add_argument("--points", type=int, default=20_000_000)
ap.add_argument("--mode", default="hybrid")
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--device", action="store_true", help="device-resident bookkeeping (torch tensors + NCCL all_to_all_single)")
ap.add_argument("--check", action="store_true", help="with --device: compare against the host-array path on the same input")
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
part_t = torch.from_numpy(part).cuda() if a.device else None
for step in range(a.steps + 1):
    sync(); t0 = time.perf_counter()
    if a.device:
        axis, cuts = sh.dev_slab_cuts(part_t, world, 0.1)
        owned = sh.dev_redistribute(part_t, axis, cuts, rank, world)
        sync(); t1 = time.perf_counter()
        nrm, info = sh.dev_slab_estimate_normals(ctx, owned, axis, cuts, rank, world, mode, 30, 0.5)
        sync(); t2 = time.perf_counter()
        plane, m_loc, off, pinfo = sh.dev_slab_segment_plane(ctx, owned, rank, world, 0.05, 3, 1000, seed=1)
    else:
        axis, cuts = sh.slab_cuts(part, world, 0.1)
        owned = sh.redistribute(part, axis, cuts, rank, world)
        sync(); t1 = time.perf_counter()
        nrm, info = sh.slab_estimate_normals(ctx, owned, axis, cuts, rank, world, mode, 30, 0.5)
        sync(); t2 = time.perf_counter()
        plane, inl, off, pinfo = sh.slab_segment_plane(ctx, owned, rank, world, 0.05, 3, 1000, seed=1, n_use=len(owned))
    sync(); t3 = time.perf_counter()
    if step:
        rows.append([t1 - t0, t2 - t1, t3 - t2, t3 - t0])
checked = None
if a.device and a.check:
    inl_dev = ctx.fetch_inlier_list(m_loc)
    axis2, cuts2 = sh.slab_cuts(part, world, 0.1)
    owned2 = sh.redistribute(part, axis2, cuts2, rank, world)
    nrm2, info2 = sh.slab_estimate_normals(ctx, owned2, axis2, cuts2, rank, world, mode, 30, 0.5)
    plane2, inl2, off2, pinfo2 = sh.slab_segment_plane(ctx, owned2, rank, world, 0.05, 3, 1000, seed=1, n_use=len(owned2))
    ok = (axis == axis2 and np.array_equal(cuts, cuts2) and np.array_equal(owned.cpu().numpy(), owned2)
          and np.array_equal(nrm.cpu().numpy(), nrm2) and np.array_equal(plane, plane2) and np.array_equal(inl_dev, inl2)
          and info["halo_points"] == info2["halo_points"])
    okt = torch.tensor([1.0 if ok else 0.0], device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    checked = bool(okt.item() == 1.0)
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
                      "bookkeeping": "device tensors + NCCL all_to_all_single" if a.device else "host numpy + all_to_all of host blocks",
                      "equal_to_host_path": checked}))
ctx.close()
if world > 1:
    dist.destroy_process_group()
