"""Diagnostic: per-call wall-clock timeline of L scans in flight (L host threads, one context each): argv[1] = L, argv[2] = steps per lane."""
import sys, threading, time
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.')
import numpy as np
from open3dpypro import _native as nat
from open3dpypro.synthetic import lidar_like
N = 10_000_000
L = int(sys.argv[1]) if len(sys.argv) > 1 else 3
S = int(sys.argv[2]) if len(sys.argv) > 2 else 3
pts = lidar_like(N, 1235).astype(np.float64)
lanes = []
for _ in range(L):
    c = nat.Context(0); hi = nat.PinnedArray((N, 3), np.float64); hi.array[...] = pts; ho = nat.PinnedArray((N, 3), np.float64)
    lanes.append((c, hi, ho))
log = []
TOKEN = threading.Lock() if len(sys.argv) > 3 and sys.argv[3] == 'token' else None
def step(j, ln, rec):
    c, hi, ho = ln
    ops = [("up", lambda: c.upload(nat.ATTR_POINTS, hi.array)), ("sync", c.synchronize), ("nrm", lambda: (c.drop(nat.ATTR_NORMALS), c.estimate_normals(0, 30, 0.0))),
           ("seg", lambda: c.segment_plane(0.05, 3, 1000, seed=1)), ("dasy", lambda: c.download_async(nat.ATTR_NORMALS, ho.array)), ("wait", c.synchronize)]
    for name, f in ops:
        if TOKEN and name == "nrm":
            t0 = time.perf_counter(); TOKEN.acquire(); t1 = time.perf_counter()
            if rec: log.append((j, "lock", t0, t1))
        t0 = time.perf_counter(); f(); t1 = time.perf_counter()
        if rec: log.append((j, name, t0, t1))
        if TOKEN and name == "dasy": TOKEN.release()
for ln in lanes:
    step(0, ln, False); step(0, ln, False)
T0 = time.perf_counter()
def worker(j):
    for _ in range(S): step(j, lanes[j], True)
th = [threading.Thread(target=worker, args=(j,)) for j in range(L)]
[t.start() for t in th]; [t.join() for t in th]
print("total ms", (time.perf_counter() - T0) * 1e3, "for", L * S, "steps:", (time.perf_counter() - T0) * 1e3 / (L * S), "ms per step")
for j, name, t0, t1 in sorted(log, key=lambda r: r[2]):
    print(f"lane {j} {name:5s} {1e3*(t0-T0):8.2f} -> {1e3*(t1-T0):8.2f}  ({1e3*(t1-t0):6.2f} ms)")
