"""o3dpp_bounds_static: the allocation-free variant of o3dpp_bounds (static __device__ scratch, no counterpart in the
reference; include/o3dpp.h).  Called through the C ABI on the caller's device buffers, twice in a row with different
inputs (the second smaller, so stale partial bounds of the first call would show), and compared bit for bit with the
existing entry point o3dpp_bounds (min / max are order independent: the tolerance is zero) and numpy's min / max.
Collected last on purpose."""
import numpy as np
import pytest

from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like

pytestmark = pytest.mark.gpu


def _existing(ctx, pts):
    ctx.upload(nat.ATTR_POINTS, pts)
    mn, mx = ctx.bounds()
    return np.concatenate([mn, mx])


def test_bounds_static_equals_bounds_twice_in_a_row(ctx):
    import torch
    dev = torch.device("cuda", 0)
    out = torch.full((6,), float("nan"), dtype=torch.float64, device=dev)
    clouds = [lidar_like(300_001, 41), lidar_like(4097, 42) * 0.25 + 3.0, np.array([[1.5, -2.0, 0.25]]),
              np.random.default_rng(7).normal(size=(4096, 3)), np.random.default_rng(8).normal(size=(12_289, 3)) * 1e6]
    for pts in clouds:
        pts = np.ascontiguousarray(pts, np.float64)
        rows = torch.from_numpy(pts).to(dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        for _ in range(2):                              # the same input again: the scratch keeps the last call's contents
            out.fill_(float("nan"))
            nat.bounds_static(0, rows.data_ptr(), len(pts), out.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
            torch.cuda.synchronize()
            got = out.cpu().numpy()
            assert np.array_equal(got, _existing(ctx, pts)), (len(pts), got)
            assert np.array_equal(got, np.concatenate([pts.min(0), pts.max(0)]))
        assert torch.cuda.memory_allocated(dev) == before


def test_bounds_static_empty_and_limit(ctx):
    import torch
    dev = torch.device("cuda", 0)
    out = torch.full((6,), 7.0, dtype=torch.float64, device=dev)
    rows = torch.zeros((1, 3), dtype=torch.float64, device=dev)
    nat.bounds_static(0, rows.data_ptr(), 0, out.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.zeros(6))
    with pytest.raises(nat.O3dppError, match="exceed the static scratch") as e:      # refused before anything is launched
        nat.bounds_static(0, rows.data_ptr(), nat.BOUNDS_STATIC_MAX_POINTS + 1, out.data_ptr())
    assert e.value.code != 0
    with pytest.raises(nat.O3dppError):
        nat.bounds_static(0, 0, 10, out.data_ptr())
    with pytest.raises(nat.O3dppError):
        nat.bounds_static(0, rows.data_ptr(), -1, out.data_ptr())


def test_bounds_static_at_the_limit_tile_count(ctx):
    """a cloud large enough for several thousand tiles (the final kernel's strided loop), then a one-tile call"""
    import torch
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(3)
    rows = torch.randn((10_000_003, 3), dtype=torch.float64, device=dev, generator=g)
    out = torch.empty(6, dtype=torch.float64, device=dev)
    nat.bounds_static(0, rows.data_ptr(), len(rows), out.data_ptr())
    torch.cuda.synchronize()
    exp = torch.cat([rows.min(0).values, rows.max(0).values])
    assert torch.equal(out, exp)
    nat.bounds_static(0, rows.data_ptr(), 100, out.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(out, torch.cat([rows[:100].min(0).values, rows[:100].max(0).values]))
