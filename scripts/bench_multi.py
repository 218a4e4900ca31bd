#!/usr/bin/env python
"""The in-library multi-GPU path (pb2_multi_*: what pine::CudaPathIntegrator drives) on every visible GPU:
ONE process, one host BVH build, sharded staging + ncclAllGather, one host thread per GPU, one ncclReduce.
  python scripts/bench_multi.py [--workload c5] [--steps 3] [--mode samples|rows]
Prints one JSON line: Msamples/s of the blocking render call (host wall clock, film delivered to the host),
the cold call (upload + first render) and the stage / reduce times."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import pine_b200 as pb

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c5")
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--mode", default="samples")
ap.add_argument("--devices", type=int, default=0, help="use the first N devices (0 = all)")
a = ap.parse_args()
S = pb.Scene()
cfg = bench.build_scene(S, a.workload)
w, h = S.size
spp, depth = cfg["spp"], cfg["depth"]
M = pb.MultiBackend(list(range(a.devices)) if a.devices else None)
M.upload_tables(spp)
film = np.empty((h, w, 4), np.float32)
t0 = time.perf_counter()
M.upload_scene(S)
t_up = time.perf_counter() - t0
M.render(spp, depth, mode=a.mode, out=film)
t_cold = time.perf_counter() - t0
c = M.counters()
M.render(spp, depth, mode=a.mode, out=film)
times = []
for _ in range(a.steps):
    t = time.perf_counter()
    M.render(spp, depth, mode=a.mode, out=film)
    times.append(time.perf_counter() - t)
c2 = M.counters()
t = time.perf_counter(); M.restage(); t_restage = time.perf_counter() - t
t = time.perf_counter(); M.restage(); t_restage2 = time.perf_counter() - t
secs = float(np.mean(times))
print(json.dumps({"path": "pb2_multi (in-library, one process)", "n_gpus": M.n, "mode": a.mode, "workload": a.workload,
                  "value": w * h * spp / secs / 1e6, "unit": "Msamples/s (blocking call, film on the host)",
                  "seconds_per_render": secs, "device_ms_slowest_gpu": c2["last_render_ms"], "reduce_ms": c2["reduce_ms"],
                  "cold_seconds": t_cold, "upload_seconds": t_up, "host_bvh_build_ms": c["bvh_build_ms"],
                  "first_stage_ms": c["stage_ms"], "restage_first_s": t_restage, "restage_pinned_s": t_restage2,
                  "film_sum_rgb": float(film[..., :3].astype(np.float64).sum())}))
M.close()
