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


def cut_voxel_indices(cuts: Sequence[float], axis: int, min_bound, voxel_size: float) -> np.ndarray:
    """voxel index along `axis` of every inner cut plane (slab_cuts puts them at min_bound + k * voxel_size)"""
    inner = np.asarray(cuts, dtype=np.float64)[1:-1]
    return np.rint((inner - float(np.asarray(min_bound)[axis])) / float(voxel_size)).astype(np.int64)


def slab_of(points: np.ndarray, axis: int, cuts: Sequence[float], grid=None) -> np.ndarray:
    """owner rank of every point: cuts[r] <= x < cuts[r+1].  grid = (min_bound, voxel_size): ownership by the voxel
    index floor((x - min_bound) / voxel_size), the very expression voxel_down_sample keys with, so that no voxel can
    straddle two ranks even when x sits within a rounding error of a cut plane."""
    if grid is not None:
        mn, vs = grid
        ia = np.floor((points[:, axis] - float(np.asarray(mn)[axis])) / float(vs)).astype(np.int64)
        return np.clip(np.searchsorted(cut_voxel_indices(cuts, axis, mn, vs), ia, side="right"), 0, len(cuts) - 2)
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
    send = [owned[halo_mask(owned, axis, cuts, r, halo)] if r != rank else np.empty((0, 3)) for r in range(world)]
    got = all_to_all_rows(send, rank, world, group)
    got = [g for r, g in enumerate(got) if r != rank and len(g)]
    return np.concatenate(got, 0) if got else np.empty((0, 3))


def all_to_all_rows(send: Sequence[np.ndarray], rank: int, world: int, group=None) -> List[np.ndarray]:
    """uneven all-to-all of (n_r, 3) float64 blocks: send[r] goes to rank r; returns the block received from
    every rank (its own block is passed through).  NCCL: one all_to_all over NVLink; gloo: pairwise isend/irecv."""
    import torch
    import torch.distributed as dist
    counts = torch.tensor([len(x) for x in send], dtype=torch.int64, device=_dev(group))
    allc = [torch.empty_like(counts) for _ in range(world)]
    dist.all_gather(allc, counts, group=group)
    recv_counts = [int(allc[r][rank]) for r in range(world)]
    dev = _dev(group)
    outs = [torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).reshape(-1, 3).to(dev) for x in send]
    ins = [torch.empty((c, 3), dtype=torch.float64, device=dev) for c in recv_counts]
    if dist.get_backend(group) == "nccl":
        dist.all_to_all(ins, outs, group=group)
    else:   # gloo has no all_to_all for uneven lists
        ins[rank].copy_(outs[rank])
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
    return [t.cpu().numpy() for t in ins]


def redistribute(part: np.ndarray, axis: int, cuts: Sequence[float], rank: int, world: int, group=None, grid=None) -> np.ndarray:
    """every rank holds an arbitrary part of the scan; afterwards it holds the points of its own slab
    (rank-major, original order inside every source block).  One all-to-all.  grid: see slab_of."""
    part = np.asarray(part, dtype=np.float64).reshape(-1, 3)
    if world == 1:
        return part
    owner = slab_of(part, axis, cuts, grid)
    got = all_to_all_rows([part[owner == r] for r in range(world)], rank, world, group)
    return np.concatenate(got, 0)


def allreduce_scores(counts: np.ndarray, errs: np.ndarray, group=None):
    """segment_plane across slabs: sum the per-hypothesis inlier counts and distance sums"""
    import torch
    import torch.distributed as dist
    c = torch.from_numpy(np.ascontiguousarray(counts, dtype=np.int64)).to(_dev(group))
    e = torch.from_numpy(np.ascontiguousarray(errs, dtype=np.float64)).to(_dev(group))
    dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(e, op=dist.ReduceOp.SUM, group=group)
    return c.cpu().numpy(), e.cpu().numpy()


# ---------------------------------------------------------------------------------- slab-sharded operators
def _allreduce(arr: np.ndarray, op: str, group=None) -> np.ndarray:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return np.array(arr, copy=True)
    t = torch.from_numpy(np.ascontiguousarray(arr)).to(_dev(group))
    dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op], group=group)
    return t.cpu().numpy()


def slab_estimate_normals(ctx, owned: np.ndarray, axis: int, cuts: Sequence[float], rank: int, world: int,
                          mode: int, knn: int, radius: float = 0.0, group=None):
    """estimate_normals of the points this rank owns, equal to the single-GPU result on the whole scan.

    hybrid / radius search: one halo exchange of width `radius` (exact).  Pure KNN: a first pass over the
    owned points alone answers every query whose k-th distance bound (o3dpp_fetch_kth_bound) stays short of
    the cut; the halo width is the largest bound among the others (all-reduce MAX), one exchange, and a
    second pass over just the strips along the cuts (own points within two halo widths + the halo) for those
    queries.  Exact ties in d2 at the k-th neighbour are broken by LOCAL index (owned points first, then halo), which can pick
    another member of the tie than the whole-scan call does.  Returns (normals of the owned points, info dict).  The context is left holding the owned points
    first (hybrid / radius: followed by the halo)."""
    from . import _native as nat
    owned = np.ascontiguousarray(owned, dtype=np.float64).reshape(-1, 3)
    n_own = len(owned)
    info = {"halo_width": 0.0, "halo_points": 0, "passes": 1}
    if world == 1:
        ctx.upload(nat.ATTR_POINTS, owned)
        ctx.drop(nat.ATTR_NORMALS)
        ctx.estimate_normals(mode, knn, radius)
        return _download_normals(ctx, n_own), info
    if mode == nat.SEARCH_KNN and knn > 32:
        raise ValueError("slab-sharded pure KNN needs knn <= 32 (the k-th distance bounds come from the certified selection path); "
                         "use KDTreeSearchParamHybrid(radius, max_nn) for larger neighbourhoods")
    if mode == nat.SEARCH_KNN:
        # pass 1: the owned points alone.  A query whose k-th distance bound r stays short of the cut already has its global
        # neighbours (safe); the others need the points of the neighbour slabs within r of the cut
        ctx.upload(nat.ATTR_POINTS, owned)
        ctx.drop(nat.ATTR_NORMALS)
        h_need = 0.0
        unsafe = np.zeros(n_own, bool)
        nrm = np.empty((0, 3))
        if n_own:
            ctx.estimate_normals(mode, knn, radius)
            nrm = _download_normals(ctx, n_own)
            r = np.sqrt(ctx.kth_bound().astype(np.float64)) * (1.0 + 1e-6)
            x = owned[:, axis]
            to_cut = np.minimum(x - cuts[rank], cuts[rank + 1] - x)        # +-inf at the outer faces
            unsafe = r > to_cut
            if unsafe.any():
                h_need = float(r[unsafe].max())
        halo_w = float(_allreduce(np.array([h_need]), "max", group)[0])
        if np.isfinite(halo_w):
            # pass 2, only for the unsafe queries: they lie within halo_w of a cut and their balls reach at most 2 halo_w into
            # the own slab and halo_w into the neighbours', so a cloud of just those strips holds all their global neighbours
            halo = exchange_halo(owned, axis, cuts, rank, world, halo_w, group)
            info.update(halo_width=halo_w, halo_points=int(len(halo)), passes=2)
            if unsafe.any():
                # the unsafe queries first: only they are queried (O3DPP_OPT_QUERY_FIRST), the rest of the strips are candidates
                near = to_cut <= 2.0 * halo_w
                m = int(unsafe.sum())
                ctx.upload(nat.ATTR_POINTS, np.concatenate([owned[unsafe], owned[near & ~unsafe], halo], 0))
                ctx.drop(nat.ATTR_NORMALS)
                ctx.set_option(nat.OPT_QUERY_FIRST, m)
                try:
                    ctx.estimate_normals(mode, knn, radius)
                finally:
                    ctx.set_option(nat.OPT_QUERY_FIRST, 0)
                nrm[unsafe] = _download_normals(ctx, m)
                info["second_pass_points"] = int(near.sum()) + int(len(halo))
                info["second_pass_queries"] = m
            ctx.upload(nat.ATTR_POINTS, owned)          # the context is left holding the owned points (slab_segment_plane)
            ctx.drop(nat.ATTR_NORMALS)
            return nrm, info
        # a slab with fewer than k points: no bound at all; one pass over everything with a halo as wide as the whole scan
        # (global extent along the axis: empty gaps between the slabs count too)
        lo = float(_allreduce(np.array([owned[:, axis].min() if n_own else np.inf]), "min", group)[0])
        hi = float(_allreduce(np.array([owned[:, axis].max() if n_own else -np.inf]), "max", group)[0])
        halo_w = (hi - lo) + 1.0
        info["passes"] = 2
    else:
        halo_w = float(radius)
    halo = exchange_halo(owned, axis, cuts, rank, world, halo_w, group)
    info.update(halo_width=halo_w, halo_points=int(len(halo)))
    loc = np.concatenate([owned, halo], 0) if len(halo) else owned
    ctx.upload(nat.ATTR_POINTS, loc)
    ctx.drop(nat.ATTR_NORMALS)
    if len(loc):
        ctx.estimate_normals(mode, knn, radius)
    return _download_normals(ctx, n_own), info


def _download_normals(ctx, n_own):
    from . import _native as nat
    n = ctx.size()
    if n == 0:
        return np.empty((0, 3))
    out = np.empty((n, 3), np.float64)
    ctx.download(nat.ATTR_NORMALS, out)
    return out[:n_own]


class RansacReplay:
    """Open3D's serial best-so-far rule with the early-exit bound (the loop of csrc/ransac.cuh
    segment_plane_impl), fed batch by batch so that scoring can stop as early as the
    single-GPU path does.  feed() returns True once the iteration budget is exhausted."""

    def __init__(self, n_total, ransac_n, num_iterations, probability):
        self.n_total, self.rn, self.H, self.prob = float(n_total), int(ransac_n), int(num_iterations), float(probability)
        self.best_fit, self.best_rmse, self.best, self.best_pl = 0.0, 0.0, -1, np.zeros(4)
        self.break_iteration, self.it, self.h = 1.7976931348623157e308, 0, 0

    def done(self):
        return self.it > self.break_iteration or self.h >= self.H

    def feed(self, planes, counts, errs):
        import math
        for pl, c, e in zip(planes, counts, errs):
            if self.it > self.break_iteration:
                break
            h = self.h
            self.h += 1
            if not pl.any():
                continue
            c = int(c)
            fit = c / self.n_total if c else 0.0
            rmse = float(e) / math.sqrt(c) if c else 0.0
            if fit > self.best_fit or (fit == self.best_fit and rmse < self.best_rmse):
                self.best_fit, self.best_rmse, self.best, self.best_pl = fit, rmse, h, np.array(pl, dtype=np.float64)
                if fit < 1.0:
                    den = math.log(1.0 - fit ** self.rn) if fit > 0.0 else 0.0
                    num = math.log(1.0 - self.prob) if self.prob < 1.0 else -math.inf
                    bi = num / den if den != 0.0 else (-math.inf if num < 0 else math.nan)   # C: negative / +0.0
                    self.break_iteration = 1.7976931348623157e308 if not (bi >= 0.0) else math.floor(min(bi, float(self.H)))
                else:
                    self.break_iteration = 0.0
            self.it += 1
        return self.done()


def ransac_select(planes: np.ndarray, counts: np.ndarray, errs: np.ndarray, n_total: int, ransac_n: int,
                  num_iterations: int, probability: float):
    """one-shot form of RansacReplay over hypotheses scored in full.  Returns (best index, best plane)."""
    r = RansacReplay(n_total, ransac_n, num_iterations, probability)
    r.feed(planes[:int(num_iterations)], counts, errs)
    return r.best, r.best_pl


def _score_and_select(ctx, planes, distance_threshold, n_own, n_total, rn, H, probability, group, batch=64):
    """score the hypotheses in batches of `batch` on this rank's points, all-reduce every batch, replay the
    selection, and stop as soon as the early-exit bound says so (exactly where the single-GPU path stops)"""
    replay = RansacReplay(n_total, rn, H, probability)
    counts_all = np.full(H, -1, np.int64)
    for h0 in range(0, H, batch):
        pl = planes[h0:h0 + batch]
        c, e = ctx.score_planes(pl, distance_threshold, n_own) if n_own else (np.zeros(len(pl), np.int64), np.zeros(len(pl)))
        c = _allreduce(c, "sum", group)
        e = _allreduce(e, "sum", group)
        counts_all[h0:h0 + len(pl)] = c
        if replay.feed(pl, c, e):
            break
    return replay.best, replay.best_pl, counts_all


def refit_plane(centroid: np.ndarray, mom: np.ndarray) -> np.ndarray:
    """plane through `centroid` from the six centred moments (xx, xy, xz, yy, yz, zz): Open3D GetPlaneFromPoints"""
    import math
    xx, xy, xz, yy, yz, zz = [float(v) for v in mom]
    det_x, det_y, det_z = yy * zz - yz * yz, xx * zz - xz * xz, xx * yy - xy * xy
    if det_x > det_y and det_x > det_z:
        abc = [det_x, xz * yz - xy * zz, xy * yz - xz * yy]
    elif det_y > det_z:
        abc = [xz * yz - xy * zz, det_y, xy * xz - yz * xx]
    else:
        abc = [xy * yz - xz * yy, xy * xz - yz * xx, det_z]
    nn = math.sqrt((abc[0] * abc[0] + abc[1] * abc[1]) + abc[2] * abc[2])
    if nn == 0.0:
        return np.zeros(4)
    a, b, c = abc[0] / nn, abc[1] / nn, abc[2] / nn
    return np.array([a, b, c, -((a * centroid[0] + b * centroid[1]) + c * centroid[2])])


def slab_segment_plane(ctx, owned: np.ndarray, rank: int, world: int, distance_threshold: float, ransac_n: int,
                       num_iterations: int, probability: float = 0.99999999, seed: int = 0, group=None,
                       n_use: Optional[int] = None):
    """segment_plane of a scan whose points are spread over the ranks.  The context must hold this rank's
    owned points FIRST (e.g. as slab_estimate_normals left it; n_use = number of owned points).  Global point
    index = position in the rank-major concatenation of the owned arrays; the hypothesis sequence is the
    Philox sequence of the single-GPU call on that concatenation, so counts, winner and inlier sets are
    identical to it (error sums differ in summation order: they only break exact count ties).
    Returns (plane [a,b,c,d], local inlier indices into `owned` ascending, global offset of this rank, info)."""
    owned = np.asarray(owned, dtype=np.float64).reshape(-1, 3)
    n_own = len(owned) if n_use is None else int(n_use)
    H, rn = int(num_iterations), int(ransac_n)
    counts_all = _allreduce(np.eye(world, dtype=np.int64)[rank] * n_own, "sum", group) if world > 1 else np.array([n_own])
    offset, n_total = int(counts_all[:rank].sum()), int(counts_all.sum())
    if rn < 3:
        raise RuntimeError("ransac_n should be set to higher than or equal to 3.")
    if n_total < rn:
        raise RuntimeError("There must be at least 'ransac_n' points.")
    samples = ctx.sample_philox(seed, n_total, rn, H)                       # identical on every rank
    pts = np.zeros((H * rn, 3))
    flat = samples.reshape(-1)
    mine = (flat >= offset) & (flat < offset + n_own)
    pts[mine] = owned[flat[mine] - offset]
    pts = _allreduce(pts, "sum", group)                                       # owners publish the sampled points
    planes = ctx.planes_from_points(pts, rn)
    best, best_pl, counts = _score_and_select(ctx, planes, distance_threshold, n_own, n_total, rn, H, probability, group)
    m, s3 = ctx.plane_inliers(best_pl, distance_threshold, n_own) if n_own else (0, np.zeros(3))
    tot = _allreduce(np.concatenate([[float(m)], s3]), "sum", group)
    m_total = int(round(tot[0]))
    plane = np.zeros(4)
    if m_total > 0:
        cen = tot[1:] / float(m_total)
        mom = ctx.inlier_moments(cen) if n_own else np.zeros(6)
        mom = _allreduce(mom, "sum", group)
        plane = refit_plane(cen, mom)
    inl = ctx.fetch_inlier_list(m) if n_own else np.empty(0, np.int32)
    return plane, inl, offset, {"best_iter": best, "best_plane": best_pl, "n_inliers": m_total, "n_total": n_total,
                                "counts": counts, "samples": samples}


# ---------------------------------------------------------------------------------- slab-sharded voxel_down_sample
def global_bounds(points, world: int, group=None) -> Tuple[np.ndarray, np.ndarray]:
    """(min_bound, max_bound) of the whole scan: one all-reduce(min) of 6 doubles.
    `points`: this rank's part, a host array or a CUDA tensor."""
    import torch
    import torch.distributed as dist
    if hasattr(points, "is_cuda"):
        big = torch.full((3,), float("inf"), dtype=torch.float64, device=points.device)
        mn = points.min(0).values.double() if len(points) else big
        mx = points.max(0).values.double() if len(points) else -big
        t = torch.cat([mn, -mx])
    else:
        p = np.asarray(points, dtype=np.float64)
        mn = p.min(0) if len(p) else np.full(3, np.inf)
        mx = p.max(0) if len(p) else np.full(3, -np.inf)
        t = torch.from_numpy(np.concatenate([mn, -mx]))
    if world > 1 and dist.is_initialized():
        t = t.to(_dev(group))
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    t = t.cpu().numpy()
    return t[:3].copy(), -t[3:]


def slab_voxel_down_sample(ctx, owned, rank: int, world: int, voxel_size: float, axis: int, cuts, min_bound, max_bound,
                           group=None, fields=("rep", "count", "centroid"), want_keys=True):
    """voxel_down_sample_and_trace of a scan sharded by slab (reference call: PointCloud.py:338-341, which passes the
    scan's min/max bound explicitly): every rank runs o3dpp_voxel_bounded on its own points with the GLOBAL bounds, so its
    voxel keys are the whole scan's.  The cuts lie on voxel planes and ownership is by voxel index (redistribute(grid=)),
    so no voxel straddles two ranks: no halo, no merge.  Collectives: one all-gather of the row counts.
    -> dict: V (local rows), V_total, row_offset (rows of the lower ranks), keys (V,3) int64 voxel indices, the fetched
    `fields` (rep indexes this rank's `owned`), globally_ordered (rank-major concatenation is the scan's ascending
    (ix,iy,iz) order iff the slabs cut the most significant axis, x; gather_voxel_rows sorts otherwise).
    `owned`: host (n,3) array or CUDA tensor."""
    import torch
    import torch.distributed as dist
    from . import _native as nat
    mn = np.asarray(min_bound, dtype=np.float64).reshape(3)
    mx = np.asarray(max_bound, dtype=np.float64).reshape(3)
    n = len(owned)
    if hasattr(owned, "is_cuda"):
        t = owned.contiguous()
        _sync_torch()
        ctx.upload_dev(nat.ATTR_POINTS, t.data_ptr(), n, nat.F32 if t.dtype == torch.float32 else nat.F64)
    else:
        ctx.upload(nat.ATTR_POINTS, owned)
    ctx.drop(nat.ATTR_NORMALS)
    V = ctx.voxel(float(voxel_size), mn, mx) if n else 0
    want = tuple(dict.fromkeys(tuple(fields) + ("rep",)))
    res = ctx.voxel_fetch(V, n, list(want)) if n else {k: np.empty((0,) + ((3,) if k == "centroid" else ()), np.float64 if k == "centroid" else np.int32) for k in want}
    rep = res["rep"]
    keys = None
    if want_keys:
        rp = owned[torch.as_tensor(rep, device=owned.device).long()].double().cpu().numpy() if hasattr(owned, "is_cuda") else np.asarray(owned, dtype=np.float64)[rep]
        keys = np.floor((rp - mn) / float(voxel_size)).astype(np.int64)           # FP64 subtract, divide, floor: the device's expression
    if V and want_keys:
        k = cut_voxel_indices(cuts, axis, mn, voxel_size)
        lo = k[rank - 1] if rank > 0 else np.iinfo(np.int64).min
        hi = k[rank] if rank < world - 1 else np.iinfo(np.int64).max
        if keys[:, axis].min() < lo or keys[:, axis].max() >= hi:
            raise ValueError("a voxel straddles a slab cut: redistribute the scan with grid=(min_bound, voxel_size)")
    counts = [V]
    if world > 1 and dist.is_initialized():
        tv = torch.tensor([V], dtype=torch.int64, device=_dev(group))
        parts = [torch.empty_like(tv) for _ in range(world)]
        dist.all_gather(parts, tv, group=group)
        counts = [int(p[0]) for p in parts]
    out = dict(res)
    out.update(V=V, V_total=int(sum(counts)), row_offset=int(sum(counts[:rank])), keys=keys, globally_ordered=(axis == 0 or world == 1))
    return out


def gather_voxel_rows(res: dict, rank: int, world: int, group=None) -> dict:
    """all ranks' voxel rows (keys, count, centroid) on every rank, in the scan's ascending (ix,iy,iz) order"""
    import torch
    import torch.distributed as dist
    cols = np.concatenate([res["keys"].astype(np.float64), np.asarray(res["count"], dtype=np.float64).reshape(-1, 1),
                           np.asarray(res["centroid"], dtype=np.float64).reshape(-1, 3)], 1) if res["V"] else np.empty((0, 7))
    if world > 1 and dist.is_initialized():
        dev = _dev(group)
        nloc = torch.tensor([len(cols)], dtype=torch.int64, device=dev)
        ns = [torch.empty_like(nloc) for _ in range(world)]
        dist.all_gather(ns, nloc, group=group)
        ns = [int(v[0]) for v in ns]
        buf = torch.zeros((max(ns + [1]), 7), dtype=torch.float64, device=dev)
        buf[: len(cols)] = torch.from_numpy(cols).to(dev)
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf, group=group)
        cols = np.concatenate([parts[r][: ns[r]].cpu().numpy() for r in range(world)], 0)
    keys = cols[:, :3].astype(np.int64)
    order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))
    return dict(keys=keys[order], count=cols[order, 3].astype(np.int32), centroid=cols[order, 4:7].copy(), V_total=len(cols))


# ---------------------------------------------------------------------------------- the same, device resident
# torch CUDA tensors in, torch CUDA tensors out: ownership, halo selection and the exchanges run on the GPU
# (NCCL all_to_all_single over NVLink) and the library reads / writes the tensors in place
# (o3dpp_upload_dev / o3dpp_download_dev), so no (n,3) array crosses PCIe.  Same results as the numpy versions.
def _sync_torch():
    import torch
    torch.cuda.current_stream().synchronize()


def dev_slab_cuts(part, world: int, voxel_size: float, min_bound=None, group=None):
    import torch
    import torch.distributed as dist
    distributed = world > 1 and dist.is_initialized()
    big = torch.full((3,), float("inf"), dtype=torch.float64, device=part.device)
    mn = part.min(0).values if len(part) else big
    mx = part.max(0).values if len(part) else -big
    t = torch.cat([mn, -mx])
    if distributed:
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    t = t.cpu().numpy()
    mn, mx = t[:3].copy(), -t[3:]
    if min_bound is None:
        min_bound = mn
    axis = int(np.argmax(mx - mn))
    nvox = int(np.floor((mx[axis] - min_bound[axis]) / voxel_size)) + 1
    stride = max(1, nvox // 65536)
    nb = (nvox + stride - 1) // stride
    idx = torch.div(torch.floor((part[:, axis] - float(min_bound[axis])) / voxel_size).long(), stride, rounding_mode="floor")
    hist = torch.bincount(idx.clamp_(0, nb - 1), minlength=nb)
    if distributed:
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    cum = np.cumsum(hist.cpu().numpy())
    total = int(cum[-1])
    cuts = [min_bound[axis]]
    for r in range(1, world):
        b = int(np.searchsorted(cum, total * r / world, side="left")) + 1
        cuts.append(min_bound[axis] + min(b, nb) * stride * voxel_size)
    cuts.append(np.inf)
    cuts[0] = -np.inf
    return axis, np.maximum.accumulate(np.asarray(cuts))


def _dev_all_to_all_rows(send, in_splits, world: int, group=None):
    """send: (sum(in_splits), 3) float64 rows grouped by destination rank -> rows received, grouped by source rank"""
    import torch
    import torch.distributed as dist
    ins = torch.tensor(in_splits, dtype=torch.int64, device=send.device)
    outs = torch.empty_like(ins)
    dist.all_to_all_single(outs, ins, group=group)
    out_splits = [int(v) for v in outs.tolist()]
    recv = torch.empty((sum(out_splits), 3), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv, send.contiguous(), out_splits, [int(v) for v in in_splits], group=group)
    return recv, out_splits


def dev_redistribute(part, axis: int, cuts, rank: int, world: int, group=None, grid=None):
    import torch
    if world == 1:
        return part
    if grid is not None:       # ownership by voxel index (see slab_of)
        mn, vs = grid
        ia = torch.floor((part[:, axis] - float(np.asarray(mn)[axis])) / float(vs)).long()
        inner = torch.as_tensor(cut_voxel_indices(cuts, axis, mn, vs), dtype=torch.int64, device=part.device)
        owner = torch.bucketize(ia, inner, right=True).clamp_(0, world - 1)
    else:
        inner = torch.as_tensor(np.asarray(cuts)[1:-1], dtype=part.dtype, device=part.device)
        owner = torch.bucketize(part[:, axis].contiguous(), inner, right=True).clamp_(0, world - 1)
    order = torch.argsort(owner, stable=True)
    counts = torch.bincount(owner, minlength=world).tolist()
    recv, _ = _dev_all_to_all_rows(part[order], counts, world, group)
    return recv


def dev_exchange_halo(owned, axis: int, cuts, rank: int, world: int, halo: float, group=None):
    import torch
    if world == 1:
        return owned[:0]
    x = owned[:, axis]
    parts, counts = [], []
    for r in range(world):
        if r == rank:
            counts.append(0)
            continue
        lo, hi = float(cuts[r]), float(cuts[r + 1])
        m = ((x >= lo - halo) & (x < lo)) | ((x >= hi) & (x < hi + halo))
        rows = owned[m]
        parts.append(rows)
        counts.append(len(rows))
    send = torch.cat(parts, 0) if parts else owned[:0]
    recv, _ = _dev_all_to_all_rows(send, counts, world, group)
    return recv


def dev_slab_estimate_normals(ctx, owned, axis: int, cuts, rank: int, world: int, mode: int, knn: int, radius: float = 0.0,
                              group=None):
    """slab_estimate_normals on device tensors: -> (normals of the owned points as a CUDA tensor, info).
    The context is left holding the owned points first (hybrid / radius: followed by the halo)."""
    import torch
    from . import _native as nat
    n_own = len(owned)
    info = {"halo_width": 0.0, "halo_points": 0, "passes": 1}

    def run(loc):
        loc = loc.contiguous()
        _sync_torch()
        ctx.upload_dev(nat.ATTR_POINTS, loc.data_ptr(), len(loc), nat.F64)
        ctx.drop(nat.ATTR_NORMALS)
        if len(loc):
            ctx.estimate_normals(mode, knn, radius)
        return loc

    import time

    def fetch(n_keep):
        out = torch.empty((ctx.size(), 3), dtype=torch.float64, device=owned.device)
        if len(out):
            ctx.download_dev(nat.ATTR_NORMALS, out.data_ptr(), nat.F64)
            ctx.synchronize()
        return out[:n_keep]

    def exchange(width):
        _sync_torch()
        t0 = time.perf_counter()
        halo = dev_exchange_halo(owned, axis, cuts, rank, world, width, group)
        _sync_torch()
        info.update(halo_width=width, halo_points=int(len(halo)), halo_bytes=int(len(halo)) * 24,
                    exchange_ms=(time.perf_counter() - t0) * 1e3)
        return halo

    if world > 1 and mode == nat.SEARCH_KNN and knn > 32:
        raise ValueError("slab-sharded pure KNN needs knn <= 32 (the k-th distance bounds come from the certified selection path); "
                         "use KDTreeSearchParamHybrid(radius, max_nn) for larger neighbourhoods")
    if world > 1 and mode == nat.SEARCH_KNN:
        # the same two passes as slab_estimate_normals: all owned points alone, then only the strips along the cuts for the
        # queries whose k-th distance bound reaches across a cut
        h_need = 0.0
        unsafe = None
        nrm = owned[:0]
        if n_own:
            run(owned)
            nrm = fetch(n_own)
            kb = torch.empty(n_own, dtype=torch.float32, device=owned.device)
            ctx.kth_bound_dev(kb.data_ptr())
            ctx.synchronize()
            r = kb.double().sqrt_() * (1.0 + 1e-6)
            del kb
            x = owned[:, axis]
            to_cut = torch.minimum(x - float(cuts[rank]), float(cuts[rank + 1]) - x)
            unsafe = r > to_cut
            if bool(unsafe.any()):
                h_need = float(r[unsafe].max())
        halo_w = float(_allreduce(np.array([h_need]), "max", group)[0])
        if np.isfinite(halo_w):
            halo = exchange(halo_w)
            info["passes"] = 2
            if unsafe is not None and bool(unsafe.any()):
                # the unsafe queries first: only they are queried (O3DPP_OPT_QUERY_FIRST), the rest of the strips are candidates
                near = to_cut <= 2.0 * halo_w
                first = owned[unsafe]
                m = int(len(first))
                rest = owned[near & ~unsafe]
                ctx.set_option(nat.OPT_QUERY_FIRST, m)
                try:
                    run(torch.cat([first, rest, halo], 0))
                finally:
                    ctx.set_option(nat.OPT_QUERY_FIRST, 0)
                nrm[unsafe] = fetch(m)
                info["second_pass_points"] = m + int(len(rest)) + int(len(halo))
                info["second_pass_queries"] = m
                del first, rest
            if n_own:
                _sync_torch()
                ctx.upload_dev(nat.ATTR_POINTS, owned.contiguous().data_ptr(), n_own, nat.F64)   # left holding the owned points
                ctx.drop(nat.ATTR_NORMALS)
            return nrm, info
        lo = float(_allreduce(np.array([float(owned[:, axis].min()) if n_own else np.inf]), "min", group)[0])
        hi = float(_allreduce(np.array([float(owned[:, axis].max()) if n_own else -np.inf]), "max", group)[0])
        halo_w = (hi - lo) + 1.0
        info["passes"] = 2
    else:
        halo_w = float(radius) if world > 1 else 0.0
    halo = exchange(halo_w)
    loc = run(torch.cat([owned, halo], 0) if len(halo) else owned)
    return fetch(n_own), info


def dev_slab_segment_plane(ctx, owned, rank: int, world: int, distance_threshold: float, ransac_n: int,
                           num_iterations: int, probability: float = 0.99999999, seed: int = 0, group=None):
    """slab_segment_plane with the owned points given as a CUDA tensor (only the 3H sampled points are read back)."""
    import torch
    n_own = len(owned)
    H, rn = int(num_iterations), int(ransac_n)
    counts_all = _allreduce(np.eye(world, dtype=np.int64)[rank] * n_own, "sum", group) if world > 1 else np.array([n_own])
    offset, n_total = int(counts_all[:rank].sum()), int(counts_all.sum())
    if rn < 3:
        raise RuntimeError("ransac_n should be set to higher than or equal to 3.")
    if n_total < rn:
        raise RuntimeError("There must be at least 'ransac_n' points.")
    samples = ctx.sample_philox(seed, n_total, rn, H)
    flat = samples.reshape(-1)
    mine = (flat >= offset) & (flat < offset + n_own)
    pts = np.zeros((H * rn, 3))
    if mine.any():
        sel = torch.from_numpy(flat[mine] - offset).to(owned.device)
        pts[mine] = owned[sel].cpu().numpy()
    pts = _allreduce(pts, "sum", group)
    planes = ctx.planes_from_points(pts, rn)
    best, best_pl, counts = _score_and_select(ctx, planes, distance_threshold, n_own, n_total, rn, H, probability, group)
    m, s3 = ctx.plane_inliers(best_pl, distance_threshold, n_own) if n_own else (0, np.zeros(3))
    tot = _allreduce(np.concatenate([[float(m)], s3]), "sum", group)
    m_total = int(round(tot[0]))
    plane = np.zeros(4)
    if m_total > 0:
        cen = tot[1:] / float(m_total)
        mom = _allreduce(ctx.inlier_moments(cen) if n_own else np.zeros(6), "sum", group)
        plane = refit_plane(cen, mom)
    return plane, m, offset, {"best_iter": best, "best_plane": best_pl, "n_inliers": m_total, "n_total": n_total,
                              "counts": counts}
