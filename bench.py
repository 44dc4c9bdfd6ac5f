#!/usr/bin/env python
"""Benchmark of the point-cloud hot path (BASELINE.json metric):
    Mpoints/s of estimate_normals + segment_plane on a 10 M-point synthetic LiDAR-like scan.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--points P]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one pass of the hot path over one scan: voxel_down_sample(0.1) (timed, reported under
`stages_ms`, not part of the metric), then estimate_normals(KNN k=30) and segment_plane(0.05, 3,
1000) on the full scan.  `value` = points / (t_normals + t_segment), inputs resident in HBM.
`e2e` = the same two calls through the C ABI with page-locked HOST buffers: H2D of the float64
(N,3) points, D2H of the float64 normals, the inlier list and the plane inside the timed region.
At N > 1 every rank owns its own scan (frame/scan sharding, no data-path collective): weak scaling.
`--impl reference` times the CPU oracle (a restatement of Open3D 0.19's algorithms, the reference's
CPU path; the Open3D wheel itself is not installable here) on the host cores.
"""
import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "open3d-py-extension_b200"))

import numpy as np  # noqa: E402

METRIC = "Mpoints/s estimate_normals+segment_plane @10M pts"
VOXEL, KNN, THR, RANSAC_N, ITERS = 0.1, 30, 0.05, 3, 1000
ALGO_BYTES_PER_POINT_KNN = 384.0   # N*(12 + 12k + 12) at k = 30


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons of one GPU sampled DURING the timed region: NVML from a thread every 5 ms
    (the timed region of the device leg is ~0.1 s), `nvidia-smi -lms` as the fallback."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index, uuid=None):
        self.rows, self.proc, self.nv, self.h, self.mx = [], None, None, None, None
        self.alive = True
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid:
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)) if not str(uuid).startswith("GPU-") else str(uuid))
                except Exception:
                    h = None
            if h is None:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                phys = index
                if vis:
                    ent = vis.split(",")[index].strip()
                    phys = int(ent) if ent.isdigit() else index
                h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.nv, self.h = pynvml, h
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        reasons_of = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while self.alive:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(reasons_of(self.h))
                self.rows.append((time.time(), mhz, mask))
            except Exception:
                pass
            time.sleep(0.005)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.nv is not None:
            time.sleep(0.02)
            self.alive = False
            sm = [r[1] for r in self.rows if t0 - 0.01 <= r[0] <= t1 + 0.01]
            mask = 0
            for r in self.rows:
                if t0 - 0.01 <= r[0] <= t1 + 0.01:
                    mask |= r[2]
            return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(k for k, b in self.BITS.items() if mask & b), "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [v.strip() for v in line.split(",")]
            if len(f) < 6:
                continue
            try:
                mx = float(f[1])
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[0]))
                    for nme, v in zip(names, f[2:6]):
                        if v.lower().startswith("active"):
                            reasons.add(nme)
            except ValueError:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvidia-smi"}


def bind_to_gpu_numa_node(index):
    """Several ranks on one host: run this process (and place its page-locked buffers, first touch) on the CPUs of the
    NUMA node its GPU hangs off, so that the 0.5 GB of copies per step do not cross the socket interconnect.
    Returns the node number, or None when the topology is not visible (nothing is changed then)."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(index).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(index), "pci_domain_id", 0)
        dev = getattr(torch.cuda.get_device_properties(index), "pci_device_id", 0)
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if len(allowed) < 2:
            return None
        os.sched_setaffinity(0, allowed)
        return node
    except Exception:
        return None


def oracle_pass(pts, threads):
    """one CPU pass of the metric's two calls on `pts`; returns seconds"""
    from oracle import orc
    t = time.perf_counter()
    orc.estimate_normals(pts, "knn", KNN, nthreads=threads)
    samples = orc.sample_philox(1, len(pts), RANSAC_N, ITERS)
    orc.segment_plane(pts, THR, RANSAC_N, ITERS, samples, nthreads=threads)
    return time.perf_counter() - t


def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import orc
    from open3dpypro.synthetic import lidar_like, SEED_C3
    threads = os.cpu_count() or 1
    n = min(args.points, args.ref_points)
    pts = lidar_like(n, SEED_C3)
    for _ in range(min(args.warmup, 1)):
        oracle_pass(pts[: max(n // 10, 1000)], threads)
    ts = [oracle_pass(pts, threads) for _ in range(args.steps)]
    ms = 1e3 * sum(ts) / len(ts)
    val = n / (ms * 1e-3) / 1e6
    sample = f"{n}-point scan from the same generator (seed {SEED_C3}), KNN k={KNN} normals + segment_plane({THR},{RANSAC_N},{ITERS}), OpenMP on all host threads"
    emit(({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Mpoints/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C3: synthetic 10M-point scan, estimate_normals(k=30)+segment_plane(0.05,3,1000)",
                   "points_per_step": n, "note": "CPU oracle = C restatement of Open3D 0.19 (wheel unavailable on this box)"},
        "cpu_baseline": {"value": val, "unit": "Mpoints/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "Mpoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


_JSON_FD = None


def claim_stdout():
    """The driver reads ONE JSON line from stdout, and NCCL prints its version banner there (NCCL_DEBUG=VERSION or
    higher, set by the box).  Keep the real stdout for the result line and point fd 1 at stderr for everything else."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--points", type=int, default=10_000_000)
    ap.add_argument("--ref-points", type=int, default=10_000_000, help="bounded CPU sample size")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-token", dest="token", action="store_false",
                    help="end-to-end lanes: let the kernels of several scans share the GPU instead of taking turns")
    ap.add_argument("--inflight", type=int, default=2, help="scans in flight in the end-to-end measurement (1 = strictly serial)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    claim_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    from open3dpypro import _native as nat
    from open3dpypro.synthetic import lidar_like, SEED_C3
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    numa = bind_to_gpu_numa_node(local) if world > 1 else None   # before the page-locked buffers are allocated
    N = args.points
    pts = lidar_like(N, SEED_C3 + rank)          # every rank owns its own scan
    ctx = nat.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local)
    hin = nat.PinnedArray((N, 3), np.float64)
    hin.array[...] = pts
    hout = nat.PinnedArray((N, 3), np.float64)
    hinl = nat.PinnedArray((N,), np.int32)       # page-locked home of the inlier list
    ctx.upload(nat.ATTR_POINTS, hin.array)

    def device_step(stages):
        ctx.drop(nat.ATTR_NORMALS)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record(stream)
        V = ctx.voxel(VOXEL)
        e[1].record(stream)
        ctx.estimate_normals(nat.SEARCH_KNN, KNN, 0.0)
        tk = ctx.timings()
        e[2].record(stream)
        plane, _ = ctx.segment_plane(THR, RANSAC_N, ITERS, seed=1, fetch_inliers=False)
        ts = ctx.timings()
        e[3].record(stream)
        e[3].synchronize()
        stages.append({"voxel": e[0].elapsed_time(e[1]), "normals": e[1].elapsed_time(e[2]),
                       "segment": e[2].elapsed_time(e[3]), "grid_build": tk[0], "knn_kernel": tk[1],
                       "knn_tier1": tk[2], "knn_tier2": tk[3], "knn_tier3": tk[6], "tier2_queries": tk[4], "tier3_queries": tk[5],
                       "ransac_score": ts[1], "ransac_scored": ts[7], "V": V})
        return plane

    # With several scans in flight the kernels of one scan at a time own the GPU (a host-side token), so that the
    # lanes fall into a staggered schedule: one computes while the others copy.  Without it the lanes run in lock
    # step, share the SMs, finish together and then all queue on PCIe with the GPU idle.
    token = threading.Lock() if args.token else contextlib.nullcontext()

    def e2e_step(lane=None):
        c, hi, ho, hx = lane or (ctx, hin, hout, hinl)
        c.upload(nat.ATTR_POINTS, hi.array)            # H2D 24 B/point from page-locked memory
        if lane is None:                               # one scan at a time: the normals leave while segment_plane runs
            c.drop(nat.ATTR_NORMALS)
            c.estimate_normals(nat.SEARCH_KNN, KNN, 0.0)
            c.download_async(nat.ATTR_NORMALS, ho.array)   # D2H 24 B/point on the copy stream
            plane, inl = c.segment_plane(THR, RANSAC_N, ITERS, seed=1, inliers_out=hx.array)   # D2H inliers + plane
        else:                                          # several in flight: the normals leave while the NEXT scan computes
            c.synchronize()
            with token:
                c.drop(nat.ATTR_NORMALS)
                c.estimate_normals(nat.SEARCH_KNN, KNN, 0.0)
                plane, inl = c.segment_plane(THR, RANSAC_N, ITERS, seed=1, inliers_out=hx.array)
                c.download_async(nat.ATTR_NORMALS, ho.array)
        c.synchronize()                                # normals complete on the host
        return plane, inl

    scratch = []
    for _ in range(args.warmup):
        device_step(scratch)
    barrier()
    sampler = ClockSampler(local, getattr(torch.cuda.get_device_properties(local), "uuid", None))
    l0 = ctx.launch_count()
    stages = []
    wall0 = time.time()
    for _ in range(args.steps):
        device_step(stages)
    barrier()
    wall1 = time.time()
    launches = ctx.launch_count() - l0
    clocks = sampler.stop(wall0, wall1)
    t_metric = sum(s["normals"] + s["segment"] for s in stages) / len(stages)      # ms
    t_step = sum(s["voxel"] + s["normals"] + s["segment"] for s in stages) / len(stages)

    for _ in range(max(1, min(args.warmup, 2))):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        plane, inl = e2e_step()
    barrier()
    t_e2e_serial = (time.perf_counter() - t0) * 1e3 / args.steps
    # the same steps with two scans in flight: two host threads, each with its own library context (the C ABI is
    # thread-safe per handle) and page-locked buffers, so that the PCIe copies of one scan overlap the kernels of
    # the other.  Every step still uploads its own input and downloads its own normals, plane and inlier list.
    t_e2e = t_e2e_serial
    if args.inflight > 1:
        lanes = [(ctx, hin, hout, hinl)]
        for _ in range(args.inflight - 1):
            c2 = nat.Context(local)
            hi2 = nat.PinnedArray((N, 3), np.float64)
            hi2.array[...] = pts
            lanes.append((c2, hi2, nat.PinnedArray((N, 3), np.float64), nat.PinnedArray((N,), np.int32)))
        for ln in lanes[1:]:
            e2e_step(ln)
        todo = list(range(args.steps))
        lock = threading.Lock()
        last = {}

        def worker(ln):
            while True:
                with lock:
                    if not todo:
                        return
                    todo.pop()
                last[id(ln)] = e2e_step(ln)

        barrier()
        t0 = time.perf_counter()
        th = [threading.Thread(target=worker, args=(ln,)) for ln in lanes]
        for t in th:
            t.start()
        for t in th:
            t.join()
        barrier()
        t_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
        for c2, hi2, ho2, hx2 in lanes[1:]:
            assert np.array_equal(ho2.array, hout.array), "pipelined normals differ from the serial result"
            hi2.free(); ho2.free(); hx2.free(); c2.close()
    n_inl = int(len(inl))

    tm = torch.tensor([t_metric, t_step, t_e2e, t_e2e_serial], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    t_metric, t_step, t_e2e, t_e2e_serial = [float(v) for v in tm.tolist()]

    if rank == 0:
        peak, peak_src = peaks()
        knn_ms = sum(s["knn_kernel"] for s in stages) / len(stages)
        achieved = ALGO_BYTES_PER_POINT_KNN * N / (knn_ms * 1e-3) / 1e9
        med = {k: statistics.median(s[k] for s in stages) for k in stages[0]}
        out = {
            "metric": METRIC, "value": world * N / (t_metric * 1e-3) / 1e6, "unit": "Mpoints/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_metric, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3: synthetic 10M-point scan, estimate_normals(k=30)+segment_plane(0.05,3,1000)"
                                   " (voxel_down_sample(0.1) timed beside it)",
                       "points_per_gpu": N, "sharding": "one scan per GPU, no data-path collective",
                       "l2": "inputs (240 MB FP64 planes per scan) exceed the 126 MB L2",
                       "timing": "CUDA events on the library stream; max over ranks",
                       "numa_node_of_rank0": numa},
            "stages_ms": med, "chain_ms_per_step": t_step,
            "e2e": {"value": world * N / (t_e2e * 1e-3) / 1e6, "unit": "Mpoints/s", "ms_per_step": t_e2e,
                    "scans_in_flight": args.inflight, "serial_value": world * N / (t_e2e_serial * 1e-3) / 1e6,
                    "serial_ms_per_step": t_e2e_serial,
                    "h2d_bytes_per_step": 24 * N, "d2h_bytes_per_step": 24 * N + 4 * n_inl + 32},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "knn_lane_kernel (tier 1) + knn_range_kernel + knn_exact_kernel: the whole neighbour search of one "
                                                        "estimate_normals call, tiers 2 and 3 included in kernel_ms", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_POINT_KNN * N, "kernel_ms": knn_ms},
            "clocks": clocks,
        }
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            n = min(N, args.ref_points)
            sec = oracle_pass(lidar_like(n, SEED_C3), threads)
            out["cpu_baseline"] = {"value": n / sec / 1e6, "unit": "Mpoints/s", "cores": threads, "kind": "port",
                                   "sample": f"one pass over a {n}-point scan of the same generator ({sec:.1f} s), "
                                             f"C oracle (Open3D 0.19 restatement), OpenMP on all host threads"}
        traffic = os.path.join(ROOT, "profiles", "knn_traffic.json")
        if os.path.exists(traffic):
            try:
                out["roofline"]["traffic"] = json.load(open(traffic)).get("dram_bytes_per_launch")
            except Exception:
                pass
        emit(out)
    hin.free()
    hout.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
