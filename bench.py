#!/usr/bin/env python
"""bench.py -- throughput of the PathIntegrator hot path on B200, and of the reference on host cores.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c4|c3|c1] [--impl reference]
  torchrun --nproc-per-node N bench.py --gpus N ...          (N > 1: one rank per GPU, NCCL)

One "step" = one complete render of the workload: every (pixel, sample) of the film traced to
`max_path_length` by the wavefront kernels (pine_b200/csrc/render.cu). The default workload is C5 of
BASELINE.md: 2000 icosphere(4) meshes flattened to 10.24 M triangles, 3840x2160, BlueSampler(64),
depth 6 -- the largest configuration of BASELINE.json and the one it shards over 1/2/4/8 GPUs.
With N GPUs the scene is replicated and the SAMPLE range is split across ranks (fixed total work
=> "strong" scaling); the float films are summed with one NCCL reduce inside the timed region.

Printed JSON (rank 0, one line): value = Msamples/s of the whole job with the scene resident in HBM
(CUDA events, max over ranks); e2e = the same through the host-buffer C ABI: scene + sampler tables
staged host->device from pinned memory and the film copied back, every step; roofline = the
closest-hit traversal kernel against the measured HBM peak using SURVEY 8(d)'s algorithmic bytes
per ray (profiles/b_ray.json); cpu_baseline = the compiled reference (oracle/_ref) on a bounded
sample of the same workload on this box's host cores.
`--impl reference` times only the reference (rank 0; other ranks exit 0).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene function, kwargs, description)
    "c5": ("mesh_field", dict(n_instances=2000, subdiv=4, seed=42, size=(3840, 2160), spp=64, depth=6,
                              with_shapes=False),
           "C5: 2000 x icosphere(4) flattened = 10.24M triangles, 3840x2160, BlueSampler(64), depth 6"),
    "c5s": ("mesh_field", dict(n_instances=2000, subdiv=4, seed=42, size=(960, 540), spp=16, depth=6,
                               with_shapes=False),
            "C5 geometry (10.24M triangles) at 960x540, BlueSampler(16), depth 6 -- profiling size"),
    "c4": ("mesh_field", dict(n_instances=400, subdiv=4, seed=1234, size=(1920, 1080), spp=64, depth=6,
                              with_shapes=True),
           "C4 stand-in (room.glb is missing): 400 x icosphere(4) = 2.05M triangles + sphere/disk, "
           "1920x1080, BlueSampler(64), depth 6"),
    "c3": ("cbox_indirect", dict(size=(1024, 1024)),
           "C3: scenes/cbox-indirect.pine geometry, 1024x1024, BlueSampler(256), depth 8"),
    "c1": ("cbox_readme", dict(size=(640, 640)),
           "C1: README Cornell box (cones + 2 boxes), 640x640, BlueSampler(16), depth 4"),
    "tiny": ("mesh_field", dict(n_instances=27, subdiv=2, seed=42, size=(320, 180), spp=4, depth=4,
                                with_shapes=True), "tiny smoke workload"),
}


def build_scene(target, workload):
    from pine_b200 import scenes
    fn, kw, _ = WORKLOADS[workload]
    return scenes.SCENES[fn](target, **kw)


class ClockSampler:
    """nvidia-smi clocks/throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


class _DevicePtr:
    """A context-owned device buffer as a __cuda_array_interface__ object (for torch.as_tensor)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (ptr, False),
                                         "version": 3, "strides": None}


def run_reference(workload, steps, warmup, step_seconds):
    """The reference's own PathIntegrator (oracle/_ref: the unmodified C++ compiled in place) on
    all host threads. Each step renders a bounded sample of the workload's film: every `row_step`-th
    row, all spp, BVH built once before the steps (oracle/ref_harness.cpp BoundedPath)."""
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/libpine_ref.so missing (built by __graft_entry__.build() "
                           "where /root/reference exists; it travels with the repo snapshot)")
    R = ref.RefScene()
    cfg = build_scene(R, workload)
    w, h = R.size
    spp, depth = cfg["spp"], cfg["depth"]
    prep_s = R.bounded_prepare(spp, depth)
    cores = ref.lib().ref_num_threads()
    # calibrate on a few spread-out rows, then size a step to ~step_seconds
    cal_rows = max(1, min(h, 4))
    cal_step = max(1, h // cal_rows)
    t_cal, _ = R.bounded_render_rows(0, h, cal_step)
    n_cal = len(range(0, h, cal_step))
    per_row = t_cal / n_cal
    rows = int(max(1, min(h, step_seconds / max(per_row, 1e-9))))
    row_step = max(1, h // rows)
    n_rows = len(range(0, h, row_step))
    times = []
    for i in range(warmup + steps):
        t, _ = R.bounded_render_rows(0, h, row_step)
        if i >= warmup:
            times.append(t)
    secs = float(np.mean(times)) if times else float("nan")
    samples = n_rows * w * spp
    return {"value": samples / secs / 1e6, "unit": "Msamples/s", "cores": int(cores), "kind": "reference",
            "sample": f"{n_rows} of {h} rows (every {row_step}th) x {w} px x {spp} spp per step = "
                      f"{samples} samples; pine::BVH built once before the steps ({prep_s:.2f} s, "
                      f"excluded); mean of {len(times)} steps of {secs:.2f} s",
            "seconds_per_step": secs, "bvh_build_s": prep_s}


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import pine_b200 as pb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torchrun")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the pine_b200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION/INFO on some
        # boxes) goes to stdout too, so keep NCCL at WARN unless told otherwise
        if not os.environ.get("PB2_KEEP_NCCL_DEBUG"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- scene: every rank builds and uploads its own replica ----
    S = pb.Scene()
    cfg = build_scene(S, args.workload)
    spp, depth = cfg["spp"], cfg["depth"]
    w, h = S.size
    B = pb.Backend(local_rank)
    t0 = time.perf_counter()
    B.upload_scene(S)
    upload_s = time.perf_counter() - t0
    c0 = B.counters()
    from pine_b200.multigpu import sample_shards
    samples = sample_shards(spp, world)[rank]  # contiguous sample range of this rank
    stream = torch.cuda.current_stream().cuda_stream
    TIME = 2  # PB2_RENDER_TIME_KERNELS

    film_t = None

    def step(flags=0):
        nonlocal film_t
        ptr = B.render_device(spp, depth, samples=samples, stream=stream, flags=flags)
        if world > 1:
            if film_t is None or film_t.data_ptr() != ptr:
                film_t = torch.as_tensor(_DevicePtr(ptr, (h, w, 4)), device=dev)
            dist.reduce(film_t, dst=0, op=dist.ReduceOp.SUM)
        return ptr

    for _ in range(args.warmup):
        step()
    # ---- timed region: K steps, scene resident ----
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if clocks:
        clocks.start()
    launches0 = B.counters()["kernel_launches"]
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    clock_info = clocks.stop() if clocks else None
    ms = ev0.elapsed_time(ev1)
    launches = B.counters()["kernel_launches"] - launches0
    ms_t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms = float(ms_t.item())
    ms_per_step = ms / args.steps
    total_samples = w * h * spp
    value = total_samples / (ms_per_step * 1e-3) / 1e6

    # ---- per-kernel pass (separate, event pairs around every launch): roofline of the trace kernel
    kt = {"trace_ms": 0.0, "shadow_ms": 0.0, "shade_ms": 0.0, "other_ms": 0.0, "closest": 0, "shadow": 0,
          "trace_launches": 0, "render_ms": 0.0}
    n_k = max(1, min(args.steps, 3))
    for _ in range(n_k):
        step(TIME)
        torch.cuda.synchronize()
        c = B.counters()
        kt["trace_ms"] += c["last_trace_ms"]; kt["shadow_ms"] += c["last_shadow_ms"]
        kt["shade_ms"] += c["last_shade_ms"]; kt["other_ms"] += c["last_other_ms"]
        kt["closest"] += c["last_closest_rays"]; kt["shadow"] += c["last_shadow_rays"]
        kt["trace_launches"] += c["last_trace_launches"]; kt["render_ms"] += c["last_render_ms"]
    rays = torch.tensor([kt["closest"] / n_k, kt["shadow"] / n_k], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    closest_per_step, shadow_per_step = float(rays[0].item()), float(rays[1].item())

    # ---- end to end: host buffers in, host film out, every step ----
    host_film = torch.empty((h, w, 4), dtype=torch.float32).pin_memory() if rank == 0 else None
    scene_bytes = c0["scene_bytes"] + 65536 + 2 * 131072

    from pine_b200.multigpu import restage_sharded

    def e2e_step():
        # N = 1: the whole scene from pinned host memory. N > 1: each rank pulls 1/N of the large arrays
        # through PCIe and the replicas are completed by in-place all-gathers over NVLink
        restage_sharded(B, rank, world, stream)
        ptr = step()
        if rank == 0:
            src = film_t if world > 1 else torch.as_tensor(_DevicePtr(ptr, (h, w, 4)), device=dev)
            host_film.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    e2e_t = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_ms_per_step = float(e2e_t.item()) / args.steps
    e2e_value = total_samples / (e2e_ms_per_step * 1e-3) / 1e6

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    film_sum = float(host_film[..., :3].double().sum().item())

    # ---- roofline of the dominant kernel (closest-hit traversal) ----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    b_ray, traffic = None, None
    bpath = os.path.join(ROOT, "profiles", "b_ray.json")
    if os.path.exists(bpath):
        entry = json.load(open(bpath)).get(args.workload)
        if entry:
            b_ray = float(entry["b_ray_closest"])
            traffic = entry.get("ncu_dram_bytes_per_launch")
    roof = {"bound": "hbm", "kernel": "k_trace (closest hit, pine_b200/csrc/render.cu)",
            "peak": peak, "peak_source": peak_src, "unit": "GB/s", "traffic": traffic}
    trace_s = kt["trace_ms"] / n_k * 1e-3
    rank0_closest = kt["closest"] / n_k
    if b_ray and trace_s > 0:
        roof["achieved"] = b_ray * rank0_closest / trace_s / 1e9
        roof["frac"] = roof["achieved"] / peak
        roof["bytes_per_ray"] = b_ray
        # the algorithmic bytes one launch stands for, to read `traffic` (ncu DRAM bytes per launch) against
        roof["algorithmic_bytes_per_launch"] = b_ray * rank0_closest / max(kt["trace_launches"] / n_k, 1)
        roof["traffic_unit"] = "bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum, profiles/b_ray.json ncu_source)"
    else:
        roof["achieved"], roof["frac"] = None, None
    roof["launches_per_step"] = kt["trace_launches"] / n_k
    roof["avg_launch_ms"] = kt["trace_ms"] / max(kt["trace_launches"], 1)
    roof["mrays_per_s_in_kernel"] = rank0_closest / trace_s / 1e6 if trace_s > 0 else None
    shares = {k: kt[k + "_ms"] / max(kt["render_ms"], 1e-9) for k in ("trace", "shadow", "shade", "other")}

    out = {
        "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mrays_per_s": (closest_per_step + shadow_per_step) / (ms_per_step * 1e-3) / 1e6,
        "rays_per_step": {"closest": closest_per_step, "shadow": shadow_per_step},
        "config": {"workload": WORKLOADS[args.workload][2], "name": args.workload, "width": w, "height": h,
                   "spp": spp, "max_path_length": depth, "integrator": "PathIntegrator(BlueSampler(spp), depth)",
                   "triangles_and_shapes": c0["n_primitives"], "bvh2_nodes": c0["bvh2_nodes"],
                   "wide_nodes": c0["wide_nodes"], "sharding": f"samples {spp}//{world} per GPU, scene replicated"
                   + (", one NCCL reduce(sum) of the float film per step" if world > 1 else ""),
                   "l2": f"no flush: scene {c0['scene_bytes'] / 1e6:.0f} MB + wavefront state of a "
                         f"{min(w * h, 1 << 24)}-slot wave exceed the 126 MB L2",
                   "host_bvh_build_ms": c0["bvh_build_ms"], "host_build_threads": c0["build_threads"],
                   "upload_scene_s": upload_s, "film_sum_rgb": film_sum},
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "ms_per_step": e2e_ms_per_step,
                "h2d_bytes_per_step": int(scene_bytes), "d2h_bytes_per_step": int(w * h * 16),
                "what": ("pb2_restage_scene" if world == 1 else f"pb2_restage_scene_part (1/{world} of the replica per rank through PCIe, all-gathered over NVLink)") + " (scene, BVH and sampler tables from pinned host memory) + "
                        "render + film to pinned host memory, every step; host BVH build not repeated"},
        "gpu_launches": int(launches),
        "clocks": clock_info,
        "roofline": roof,
        "kernel_time_shares": shares,
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            out["cpu_baseline"] = run_reference(args.workload, 1, 0, args.cpu_seconds)
        except Exception as e:  # the checker is missing: say so, do not fail the GPU number
            out["cpu_baseline"] = {"value": None, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
                                   "sample": f"unavailable: {e}"}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("PB2_WORKLOAD", "c5"), choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="size of the cpu_baseline sample")
    ap.add_argument("--ref-step-seconds", type=float, default=6.0, help="--impl reference: size of one step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "graft":
        args.warmup = max(args.warmup, 0)  # honoured as given; the contract asks the driver for W >= 3
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        r = run_reference(args.workload, args.steps, args.warmup, args.ref_step_seconds)
        fn, kw, desc = WORKLOADS[args.workload]
        out = {"impl": "reference", "metric": "Msamples/s", "value": r["value"], "unit": "Msamples/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": desc, "name": args.workload,
                          "integrator": "PathIntegrator(BVH(), BlueSobolSampler(spp), UniformLightSampler(), depth)"},
               "cpu_baseline": r,
               "e2e": {"value": r["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(out), flush=True)
        return
    run_gpu(args)


if __name__ == "__main__":
    main()
