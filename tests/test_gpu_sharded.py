"""Slab-sharded scan on the GPU: two ranks (gloo rendezvous, both on cuda:0 so the
test runs on a one-GPU box; the NCCL path is the same code with device tensors) must reproduce the
single-context results: normals of every owned point, RANSAC winner, inlier set, refit plane.
Host-side pieces (selection replay, refit) are also checked against the oracle on the CPU."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_ransac_select_and_refit_match_oracle(orc):
    """host replay used by the sharded segment_plane == the oracle's SegmentPlane on the same scores"""
    from open3dpypro import sharding as sh
    from open3dpypro.synthetic import lidar_like
    pts = lidar_like(20_000, 5)
    H = 60
    samples = orc.sample_philox(11, len(pts), 3, H)
    planes = np.array([orc.plane_model(pts, samples[h]) for h in range(H)])
    sc = [orc.score_plane(pts, planes[h], 0.05) for h in range(H)]
    counts = np.array([c for c, _ in sc], np.int64)
    errs = np.array([e for _, e in sc])
    best, best_pl = sh.ransac_select(planes, counts, errs, len(pts), 3, H, 0.99999999)
    eplane, einl = orc.segment_plane(pts, 0.05, 3, H, samples)
    d = np.abs(pts @ best_pl[:3] + best_pl[3])
    inl = np.nonzero(d < 0.05)[0]
    assert np.array_equal(inl, einl)
    cen = pts[inl].mean(0)
    r = pts[inl] - cen
    mom = np.array([(r[:, 0] * r[:, 0]).sum(), (r[:, 0] * r[:, 1]).sum(), (r[:, 0] * r[:, 2]).sum(),
                    (r[:, 1] * r[:, 1]).sum(), (r[:, 1] * r[:, 2]).sum(), (r[:, 2] * r[:, 2]).sum()])
    assert np.abs(sh.refit_plane(cen, mom) - eplane).max() < 1e-9


@pytest.mark.gpu
def test_parts_equal_segment_plane_single_rank(ctx):
    """world = 1: the split entry points chained by sharding.slab_segment_plane == o3dpp_segment_plane"""
    from open3dpypro import _native as nat
    from open3dpypro import sharding as sh
    from open3dpypro.synthetic import lidar_like
    pts = lidar_like(150_000, 77)
    ctx.upload(nat.ATTR_POINTS, pts)
    plane, inl, dbg = ctx.segment_plane(0.05, 3, 300, seed=5, debug=True)
    assert np.array_equal(ctx.sample_philox(5, len(pts), 3, 300), ctx.fetch_samples(300, 3))
    p2, inl2, off, info = sh.slab_segment_plane(ctx, pts, 0, 1, 0.05, 3, 300, seed=5)
    assert off == 0 and info["best_iter"] == dbg["best_iter"]
    assert np.array_equal(info["best_plane"], dbg["best_plane"])
    assert np.array_equal(inl2, inl)
    assert np.abs(p2 - plane).max() < 1e-9
    scored = dbg["counts"] >= 0
    assert np.array_equal(info["counts"][scored], dbg["counts"][scored])


@pytest.mark.gpu
def test_device_resident_path_equals_host_path_single_rank(ctx):
    """dev_slab_* (torch CUDA tensors, in-place library I/O) == slab_* (host arrays) at world = 1; the world-2 NCCL
    comparison is `tools/bench_slab.py --device --check` (profiles/r01f_slab2_dev_hybrid.json: equal_to_host_path)"""
    import torch
    from open3dpypro import _native as nat
    from open3dpypro import sharding as sh
    from open3dpypro.synthetic import lidar_like
    pts = lidar_like(100_000, 12).astype(np.float64)
    t = torch.from_numpy(pts).cuda()
    axis, cuts = sh.slab_cuts(pts, 1, 0.1)
    axis_d, cuts_d = sh.dev_slab_cuts(t, 1, 0.1)
    assert axis == axis_d and np.array_equal(cuts, cuts_d)
    for mode, radius in ((nat.SEARCH_KNN, 0.0), (nat.SEARCH_HYBRID, 0.5)):
        n_h, _ = sh.slab_estimate_normals(ctx, pts, axis, cuts, 0, 1, mode, 30, radius)
        n_d, _ = sh.dev_slab_estimate_normals(ctx, t, axis, cuts, 0, 1, mode, 30, radius)
        assert np.array_equal(n_d.cpu().numpy(), n_h)
    p_h, inl_h, _, info_h = sh.slab_segment_plane(ctx, pts, 0, 1, 0.05, 3, 300, seed=4)
    p_d, m_d, _, info_d = sh.dev_slab_segment_plane(ctx, t, 0, 1, 0.05, 3, 300, seed=4)
    assert np.array_equal(p_h, p_d) and m_d == len(inl_h) and info_h["best_iter"] == info_d["best_iter"]
    assert np.array_equal(ctx.fetch_inlier_list(m_d), inl_h)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))
    import torch.distributed as dist
    from open3dpypro import _native as nat
    from open3dpypro import sharding as sh
    from open3dpypro.synthetic import lidar_like
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        ctx = nat.Context(0)
        pts = lidar_like(120_000, 31)
        part = pts[rank::world]                                  # an arbitrary half of the scan
        axis, cuts = sh.slab_cuts(part, world, 0.1)
        owned = sh.redistribute(part, axis, cuts, rank, world)
        assert np.all(sh.slab_of(owned, axis, cuts) == rank)
        out = {"rank": rank, "owned": owned}
        for mode, knn, radius in ((nat.SEARCH_KNN, 30, 0.0), (nat.SEARCH_HYBRID, 30, 0.5)):
            nrm, info = sh.slab_estimate_normals(ctx, owned, axis, cuts, rank, world, mode, knn, radius)
            out[f"normals{mode}"] = nrm
            out[f"info{mode}"] = info
        plane, inl, off, info = sh.slab_segment_plane(ctx, owned, rank, world, 0.05, 3, 400, seed=3, n_use=len(owned))
        out.update(plane=plane, inl=inl, off=off, best_iter=info["best_iter"], n_inliers=info["n_inliers"])
        # voxel_down_sample across the slabs: global bounds as the grid origin, ownership by voxel index
        gmn, gmx = sh.global_bounds(part, world)
        vaxis, vcuts = sh.slab_cuts(part, world, 0.1, min_bound=gmn)
        mine_v = sh.redistribute(part, vaxis, vcuts, rank, world, grid=(gmn, 0.1))
        vres = sh.slab_voxel_down_sample(ctx, mine_v, rank, world, 0.1, vaxis, vcuts, gmn, gmx)
        assert np.array_equal(mine_v[vres["rep"]].shape, (vres["V"], 3))
        out["vox"] = sh.gather_voxel_rows(vres, rank, world)
        out["vox_local"] = (vres["V"], vres["row_offset"], vres["V_total"], vaxis)
        # transform: every rank maps its own rows, no exchange
        T = np.eye(4); T[:3, :3] = np.linalg.qr(np.random.default_rng(9).normal(size=(3, 3)))[0]; T[:3, 3] = (1.5, -2.0, 0.25)
        ctx.upload(nat.ATTR_POINTS, owned)
        ctx.transform(T)
        moved = np.empty_like(owned)
        ctx.download(nat.ATTR_POINTS, moved)
        out["transformed"] = moved
        out["T"] = T
        ctx.close()
        q.put((rank, "ok", out))
    except Exception:   # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_world2_slabs_equal_single_context(ctx):
    import torch.multiprocessing as mp
    from open3dpypro import _native as nat
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=600) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), [r[1] for r in res]
    o0, o1 = res[0][2], res[1][2]
    whole = np.concatenate([o0["owned"], o1["owned"]], 0)          # the global (rank-major) order
    n0 = len(o0["owned"])
    assert o0["off"] == 0 and o1["off"] == n0 and len(whole) == 120_000
    ctx.upload(nat.ATTR_POINTS, whole)
    for mode, knn, radius in ((nat.SEARCH_KNN, 30, 0.0), (nat.SEARCH_HYBRID, 30, 0.5)):
        ctx.drop(nat.ATTR_NORMALS)
        ctx.estimate_normals(mode, knn, radius)
        ref = np.empty((len(whole), 3))
        ctx.download(nat.ATTR_NORMALS, ref)
        got = np.concatenate([o0[f"normals{mode}"], o1[f"normals{mode}"]], 0)
        cosang = np.abs(np.einsum("ij,ij->i", got, ref))
        # same neighbour sets; the cumulants are summed in another order, so compare as directions (1e-4 rad bar)
        assert cosang.min() > 1 - 5e-9, (mode, float(cosang.min()), int((cosang < 1 - 5e-9).sum()))
        assert o0[f"info{mode}"]["halo_points"] > 0
    plane, inl, dbg = ctx.segment_plane(0.05, 3, 400, seed=3, debug=True)
    assert o0["best_iter"] == o1["best_iter"] == dbg["best_iter"]
    ginl = np.concatenate([o0["inl"].astype(np.int64), o1["inl"].astype(np.int64) + n0])
    assert np.array_equal(ginl, inl)
    assert o0["n_inliers"] == len(inl)
    assert np.abs(o0["plane"] - plane).max() < 1e-9 and np.array_equal(o0["plane"], o1["plane"])
    # sharded voxel_down_sample == the single-context call == the oracle, rows in ascending (ix,iy,iz)
    from oracle import orc
    from open3dpypro.synthetic import lidar_like
    pts = lidar_like(120_000, 31)
    ctx.upload(nat.ATTR_POINTS, pts)
    V = ctx.voxel(0.1)
    one = ctx.voxel_fetch(V, len(pts), ["count", "centroid"])
    ref = orc.voxel_trace(pts, 0.1)
    for o in (o0, o1):
        vx = o["vox"]
        assert vx["V_total"] == V == ref["V"]
        assert np.array_equal(vx["keys"], ref["keys"]) and np.array_equal(vx["count"], ref["count"]) and np.array_equal(vx["count"], one["count"])
        assert np.abs(vx["centroid"] - ref["centroid"]).max() <= 1e-5 * np.abs(ref["centroid"]).max()      # north_star bar; sums differ in order only
        assert np.abs(vx["centroid"] - one["centroid"]).max() <= 1e-12 * np.abs(ref["centroid"]).max()
    assert o0["vox_local"][0] + o1["vox_local"][0] == V and o0["vox_local"][1] == 0 and o1["vox_local"][1] == o0["vox_local"][0]
    # sharded transform: the concatenation of the ranks' rows is bit-equal to the oracle's transform of the whole scan
    assert np.array_equal(o0["T"], o1["T"])
    assert np.array_equal(np.concatenate([o0["transformed"], o1["transformed"]], 0), orc.transform(whole, o0["T"])[0])
