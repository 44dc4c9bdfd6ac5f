"""Multi-GPU sharding of the hot path: one process per GPU (torch.distributed; NCCL over NVLink on the
box, gloo in the CPU tests).

  frames (config 4)  frame f -> rank f mod G, the reference's own placement rule
                     (`cuda:{i % num_gpus}`, reference processors.py:206, :339, :433).  No data-path
                     collective; the only exchange is the optional all-gather of the 4-double plane
                     models at the end of a batch.
  slabs (config 5)   one very large scan is cut along its longest axis at voxel-index planes (so no
                     voxel straddles two ranks) into slabs of near-equal point count; every rank also
                     receives the HALO: the points of its neighbours within `halo` of the cut, which
                     is what estimate_normals(hybrid radius r) needs with halo = r, and what
                     KNN(k) needs after one re-exchange for the few queries whose k-th distance
                     exceeds their distance to the cut.  segment_plane shards by all-reducing the H
                     per-hypothesis (count, error) pairs.
"""
from typing import List, Optional, Sequence, Tuple

import numpy as np


def _dev(group=None):
    import torch.distributed as dist
    return "cuda" if dist.get_backend(group) == "nccl" else "cpu"


# ---------------------------------------------------------------------------------- frames
def frames_of_rank(n_frames: int, rank: int, world: int) -> List[int]:
    """frame f is processed by rank f mod world (reference processors.py:206)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, int(n_frames), world))


def gather_frame_results(local: np.ndarray, n_frames: int, rank: int, world: int, group=None) -> np.ndarray:
    """all-gather per-frame result rows (e.g. the (4,) plane models) into frame order.
    `local` holds this rank's rows in the order of frames_of_rank()."""
    local = np.ascontiguousarray(local, dtype=np.float64)
    width = local.shape[1] if local.ndim == 2 else 1
    local = local.reshape(-1, width)
    if world == 1:
        return local.copy()
    import torch
    import torch.distributed as dist
    per = (n_frames + world - 1) // world
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    buf = torch.full((per, width), float("nan"), dtype=torch.float64, device=dev)
    buf[: len(local)] = torch.from_numpy(local).to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = np.empty((n_frames, width))
    for r in range(world):
        fr = frames_of_rank(n_frames, r, world)
        out[fr] = parts[r][: len(fr)].cpu().numpy()
    return out


# ---------------------------------------------------------------------------------- slabs
def slab_cuts(points: np.ndarray, world: int, voxel_size: float, min_bound: Optional[np.ndarray] = None,
              group=None) -> Tuple[int, np.ndarray]:
    """Cut planes of a slab decomposition.  Every rank passes the points it currently holds (any
    split of the scan); returns (axis, cuts) with `world + 1` cut coordinates, identical on all
    ranks.  Cuts lie on voxel planes min_bound[axis] + i * voxel_size so that voxel_down_sample needs
    no halo.  Collectives: one all-reduce(min/max), one all-reduce(sum) of the axis histogram."""
    import torch
    import torch.distributed as dist
    distributed = world > 1 and dist.is_initialized()   # uninitialised: the caller holds the whole scan
    p = np.asarray(points, dtype=np.float64)
    mn = p.min(0) if len(p) else np.full(3, np.inf)
    mx = p.max(0) if len(p) else np.full(3, -np.inf)
    if distributed:
        t = torch.from_numpy(np.concatenate([mn, -mx])).to(_dev(group))
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        t = t.cpu()
        mn, mx = t[:3].numpy().copy(), -t[3:].numpy()
    if min_bound is None:
        min_bound = mn
    axis = int(np.argmax(mx - mn))
    nvox = int(np.floor((mx[axis] - min_bound[axis]) / voxel_size)) + 1
    stride = max(1, nvox // 65536)                      # histogram bins are whole numbers of voxels
    nb = (nvox + stride - 1) // stride
    idx = np.floor((p[:, axis] - min_bound[axis]) / voxel_size).astype(np.int64) // stride
    hist = np.bincount(np.clip(idx, 0, nb - 1), minlength=nb).astype(np.int64)
    if distributed:
        t = torch.from_numpy(hist).to(_dev(group))
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        hist = t.cpu().numpy()
    cum = np.cumsum(hist)
    total = int(cum[-1])
    cuts = [min_bound[axis]]
    for r in range(1, world):
        b = int(np.searchsorted(cum, total * r / world, side="left")) + 1     # bins [0, b) go left
        cuts.append(min_bound[axis] + min(b, nb) * stride * voxel_size)
    cuts.append(np.inf)
    cuts[0] = -np.inf
    return axis, np.maximum.accumulate(np.asarray(cuts))


def slab_of(points: np.ndarray, axis: int, cuts: Sequence[float]) -> np.ndarray:
    """owner rank of every point: cuts[r] <= x < cuts[r+1]"""
    return np.clip(np.searchsorted(np.asarray(cuts)[1:-1], points[:, axis], side="right"), 0, len(cuts) - 2)


def halo_mask(points: np.ndarray, axis: int, cuts: Sequence[float], rank: int, halo: float) -> np.ndarray:
    """points owned by OTHER ranks that rank `rank` needs: within `halo` of its slab along the axis"""
    x = points[:, axis]
    lo, hi = cuts[rank], cuts[rank + 1]
    return ((x >= lo - halo) & (x < lo)) | ((x >= hi) & (x < hi + halo))


def exchange_halo(owned: np.ndarray, axis: int, cuts: Sequence[float], rank: int, world: int, halo: float,
                  group=None) -> np.ndarray:
    """Send the boundary points to ranks +-1 (and further if `halo` spans more than one slab) and return
    the halo points received.  One all-to-all of (n,3) float64 blocks."""
    if world == 1:
        return np.empty((0, 3))
    import torch
    import torch.distributed as dist
    send = [owned[halo_mask(owned, axis, cuts, r, halo)] if r != rank else np.empty((0, 3)) for r in range(world)]
    counts = torch.tensor([len(s) for s in send], dtype=torch.int64, device=_dev(group))
    allc = [torch.empty_like(counts) for _ in range(world)]
    dist.all_gather(allc, counts, group=group)
    recv_counts = [int(allc[r][rank]) for r in range(world)]
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    outs = [torch.from_numpy(np.ascontiguousarray(s, dtype=np.float64)).reshape(-1, 3).to(dev) for s in send]
    ins = [torch.empty((c, 3), dtype=torch.float64, device=dev) for c in recv_counts]
    if dist.get_backend(group) == "nccl":
        dist.all_to_all(ins, outs, group=group)
    else:   # gloo has no all_to_all for uneven lists: pairwise isend/irecv
        reqs = []
        for r in range(world):
            if r == rank:
                continue
            if len(outs[r]):
                reqs.append(dist.isend(outs[r], r, group=group))
            if recv_counts[r]:
                reqs.append(dist.irecv(ins[r], r, group=group))
        for q in reqs:
            q.wait()
    got = [t.cpu().numpy() for r, t in enumerate(ins) if r != rank and recv_counts[r]]
    return np.concatenate(got, 0) if got else np.empty((0, 3))


def allreduce_scores(counts: np.ndarray, errs: np.ndarray, group=None):
    """segment_plane across slabs: sum the per-hypothesis inlier counts and distance sums"""
    import torch
    import torch.distributed as dist
    c = torch.from_numpy(np.ascontiguousarray(counts, dtype=np.int64)).to(_dev(group))
    e = torch.from_numpy(np.ascontiguousarray(errs, dtype=np.float64)).to(_dev(group))
    dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(e, op=dist.ReduceOp.SUM, group=group)
    return c.cpu().numpy(), e.cpu().numpy()
