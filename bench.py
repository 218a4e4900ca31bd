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

Printed JSON (rank 0, ONE line on stdout; everything else -- NCCL's own log included -- goes to stderr):
  value        Msamples/s of the whole job with the scene resident in HBM (CUDA events, max over ranks)
  e2e          the same through the host-buffer C ABI: pb2_restage_scene (scene + BVH + sampler tables from
               pinned host memory) + pb2_render into a pinned HOST film, every step
  e2e_cold     the wall of ONE drop-in call as `Integrator::render(Scene&)` makes it: flatten, pb2_upload_scene
               (host SAH build, layout, page-lock, H2D), render, film to the host
  roofline     the closest-hit traversal kernel against the measured HBM peak on SURVEY 8(d)'s algorithmic
               bytes per ray (profiles/b_ray.json), with the ncu-measured DRAM fraction beside it
  scenes       (N = 1) the other named configurations C1, C3, C4 measured the same way in the same run, C1/C3
               against the L2 read bandwidth measured live (their scenes are cache resident)
  cpu_baseline the compiled reference (oracle/_ref) on a bounded sample of the same workload, host cores
`--impl reference` times only the reference (rank 0; other ranks exit 0).
"""
import argparse
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
    "c4": ("room_proc", dict(size=(1920, 1080), spp=64, depth=6),
           "C4 room-proc (scenes/room.glb is missing): closed 8x3x6 m room, 60 icosphere(5) + 64 tessellated boxes = "
           "2.02M triangles, Diffuse/Metal/Glossy, one 1x1 m emitter, 1920x1080, BlueSampler(64), depth 6"),
    "c4_field": ("mesh_field", dict(n_instances=400, subdiv=4, seed=1234, size=(1920, 1080), spp=64, depth=6,
                                    with_shapes=True),
                 "open field of 400 x icosphere(4) = 2.05M triangles + sphere/disk (round 1's C4 stand-in), "
                 "1920x1080, BlueSampler(64), depth 6"),
    "c3": ("cbox_indirect", dict(size=(1024, 1024)),
           "C3: scenes/cbox-indirect.pine geometry, 1024x1024, BlueSampler(256), depth 8"),
    "c1": ("cbox_readme", dict(size=(640, 640)),
           "C1: README Cornell box (cones + 2 boxes), 640x640, BlueSampler(16), depth 4"),
    "tiny": ("mesh_field", dict(n_instances=27, subdiv=2, seed=42, size=(320, 180), spp=4, depth=4,
                                with_shapes=True), "tiny smoke workload"),
}
L2_RESIDENT = {"c1", "c3"}  # scenes + BVH below 1 MB: traversal is L1/L2 bound, HBM carries queues and film only


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


def run_reference(workload, steps, warmup, step_seconds, accel="bvh"):
    """The reference's own PathIntegrator (oracle/_ref: the unmodified C++ compiled in place) on
    all host threads. Each step renders a bounded sample of the workload's film: every `row_step`-th
    row, all spp, BVH built once before the steps (oracle/ref_harness.cpp BoundedPath).
    accel "bvh": pine::BVH (the bit-exact oracle); "embree": EmbreeAccel over the vendored Embree 4.3.1, the
    accel PRL's PathIntegrator(sampler, depth) really constructs (program_context.cpp:79-81) -- needs a process
    started with PINE_REF_EMBREE=1 (see run_reference_embree)."""
    from oracle import ref
    if accel == "embree":
        ref.use_embree_accel(True)
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
            "accel": "EmbreeAccel (vendored Embree 4.3.1, ISAs <= AVX2, tasking INTERNAL)" if accel == "embree" else "pine::BVH",
            "film": [w, h], "spp": spp, "max_path_length": depth,
            "sample": f"{n_rows} of {h} rows (every {row_step}th) x {w} px x {spp} spp per step = "
                      f"{samples} samples; pine::BVH built once before the steps ({prep_s:.2f} s, "
                      f"excluded); mean of {len(times)} steps of {secs:.2f} s",
            "seconds_per_step": secs, "bvh_build_s": prep_s}


def run_reference_embree(workload, step_seconds):
    """The Embree-backed reference in a child process (its library exports the same symbols as the oracle's)."""
    from oracle import ref
    if not ref.embree_available():
        return {"value": None, "sample": "oracle/_ref/libpine_ref_embree.so not built (`make -C oracle embree`)"}
    env = dict(os.environ, PINE_REF_EMBREE="1")
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--accel", "embree",
                        "--workload", workload, "--steps", "1", "--warmup", "0", "--ref-step-seconds", str(step_seconds)],
                       env=env, capture_output=True, text=True, timeout=900)
    try:
        return json.loads(r.stdout.strip().splitlines()[-1])["cpu_baseline"]
    except Exception:
        return {"value": None, "sample": "embree reference run failed: " + (r.stderr or r.stdout)[-300:]}


def _config(workload, w, h, spp, depth):
    return {"workload": WORKLOADS[workload][2], "name": workload, "width": w, "height": h, "spp": spp,
            "max_path_length": depth, "integrator": "PathIntegrator(BlueSampler(spp), max_path_length)",
            # timing rule: no L2 flush between steps because a step's inputs are larger than the 126 MB L2 (the
            # wavefront state of w*h*spp samples alone is GBs; C4 / C5 add a 0.3-1.4 GB scene); sizes in details.l2
            "l2": "no flush: per-step inputs (wavefront state of width*height*spp samples + scene) exceed the 126 MB L2"}


def _peaks():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _b_ray(workload):
    bpath = os.path.join(ROOT, "profiles", "b_ray.json")
    if os.path.exists(bpath):
        return json.load(open(bpath)).get(workload)
    return None


def measure(args, workload, steps, warmup, world, rank, dev, ts, l2_gbs=None, main=True):
    """One workload on this rank's GPU -> result dict on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist
    import pine_b200 as pb
    from pine_b200.multigpu import restage_sharded, sample_shards

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the drop-in call, cold: flatten + upload (host build, layout, H2D) + render + film to the host ----
    B = pb.Backend(dev.index)
    host_film = torch.empty((1,), dtype=torch.float32)
    t0 = time.perf_counter()
    S = pb.Scene()
    cfg = build_scene(S, workload)
    t_scene = time.perf_counter() - t0
    spp, depth = cfg["spp"], cfg["depth"]
    w, h = S.size
    host_film = torch.empty((h, w, 4), dtype=torch.float32).pin_memory()
    film_np = host_film.numpy()
    samples = sample_shards(spp, world)[rank]  # contiguous sample range of this rank
    B.upload_tables(spp)
    barrier()
    t0 = time.perf_counter()
    if world == 1:
        B.upload_scene(S)
    else:
        # ONE host build per node: rank 0 builds and exports its built scene to /dev/shm, the other ranks map it
        # (pb2_export_host_scene / pb2_import_host_scene) instead of building the same tree N times on one host
        shared = f"/dev/shm/pb2_scene_{os.environ.get('MASTER_PORT', '0')}_{workload}.bin"
        if rank == 0:
            B.upload_scene(S)
            B.export_scene(shared)
        dist.barrier()
        if rank != 0:
            B.import_scene(shared, S.size)
        dist.barrier()
        if rank == 0:
            os.unlink(shared)
    upload_s = time.perf_counter() - t0
    film_t = None

    def step(flags=0):
        """resident scene -> film on the device (rank 0 holds the sum); asynchronous on ts"""
        nonlocal film_t
        with torch.cuda.stream(ts):
            ptr = B.render_device(spp, depth, samples=samples, stream=ts.cuda_stream, flags=flags)
            if world > 1:
                if film_t is None or film_t.data_ptr() != ptr:
                    film_t = torch.as_tensor(_DevicePtr(ptr, (h, w, 4)), device=dev)
                dist.reduce(film_t, dst=0, op=dist.ReduceOp.SUM)
        return ptr

    def to_host(ptr):
        with torch.cuda.stream(ts):
            if rank == 0:
                src = film_t if world > 1 else torch.as_tensor(_DevicePtr(ptr, (h, w, 4)), device=dev)
                host_film.copy_(src, non_blocking=True)
        ts.synchronize()

    if world == 1:
        B.render(spp, depth, out=film_np)          # pb2_render: host film buffer
    else:
        to_host(step())
    barrier()
    cold_s = time.perf_counter() - t0
    cold_t = torch.tensor([cold_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(cold_t, op=dist.ReduceOp.MAX)
    cold_s = float(cold_t.item())
    c0 = B.counters()

    for _ in range(warmup):
        step()
    # ---- timed region: K steps, scene resident ----
    clocks = ClockSampler(dev.index) if (rank == 0 and main) else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if clocks:
        clocks.start()
    launches0 = B.counters()["kernel_launches"]
    ev0.record(ts)
    for _ in range(steps):
        step()
    ev1.record(ts)
    ts.synchronize()
    barrier()
    clock_info = clocks.stop() if clocks else None
    ms = ev0.elapsed_time(ev1)
    launches = B.counters()["kernel_launches"] - launches0
    ms_t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms_t.item()) / steps
    total_samples = w * h * spp
    value = total_samples / (ms_per_step * 1e-3) / 1e6

    # ---- per-kernel pass (separate; the lanes run one after the other with an event after every launch):
    # duration of the traversal launches alone, for the roofline
    TIME = 2  # PB2_RENDER_TIME_KERNELS
    kt = {"trace_ms": 0.0, "shadow_ms": 0.0, "shade_ms": 0.0, "other_ms": 0.0, "closest": 0, "shadow": 0,
          "trace_launches": 0, "render_ms": 0.0, "fb_closest": 0}
    n_k = max(1, min(steps, 3))
    for _ in range(n_k):
        step(TIME)
        ts.synchronize()
        c = B.counters()
        kt["trace_ms"] += c["last_trace_ms"]; kt["shadow_ms"] += c["last_shadow_ms"]
        kt["shade_ms"] += c["last_shade_ms"]; kt["other_ms"] += c["last_other_ms"]
        kt["closest"] += c["last_closest_rays"]; kt["shadow"] += c["last_shadow_rays"]
        kt["trace_launches"] += c["last_trace_launches"]; kt["render_ms"] += c["last_render_ms"]
        kt["fb_closest"] += c["last_closest_fallbacks"]
    rays = torch.tensor([kt["closest"] / n_k, kt["shadow"] / n_k], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    closest_per_step, shadow_per_step = float(rays[0].item()), float(rays[1].item())

    # ---- end to end: host buffers in, host film out, every step ----
    scene_bytes = c0["scene_bytes"] + 65536 + 2 * 131072

    def e2e_step():
        if world == 1:
            # the C ABI's host-buffer calls on the context's own stream: pb2_restage_scene (scene, BVH and
            # tables from pinned host memory) then pb2_render (film into pinned host memory)
            B.restage()
            B.render(spp, depth, out=film_np)
        else:
            # each rank pulls 1/N of the large arrays through PCIe, the replicas are completed by in-place
            # all-gathers over NVLink; render, reduce and the film copy follow on the same stream
            restage_sharded(B, rank, world, ts)
            to_host(step())

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    e2e_t = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_ms_per_step = float(e2e_t.item()) / steps
    e2e_value = total_samples / (e2e_ms_per_step * 1e-3) / 1e6
    if rank != 0:
        B.close()
        return None
    film_sum = float(host_film[..., :3].double().sum().item())

    # ---- roofline of the dominant kernel (closest-hit traversal) ----
    peak, peak_src = _peaks()
    entry = _b_ray(workload)
    b_ray = float(entry["b_ray_closest"]) if entry else None
    traffic = entry.get("ncu_dram_bytes_per_launch") if entry else None
    roof = {"bound": "hbm", "kernel": "k_trace (closest hit, pine_b200/csrc/render.cu)",
            "peak": peak, "peak_source": peak_src, "unit": "GB/s", "traffic": traffic}
    trace_s = kt["trace_ms"] / n_k * 1e-3
    rank0_closest = kt["closest"] / n_k
    avg_launch_s = kt["trace_ms"] / max(kt["trace_launches"], 1) * 1e-3
    if b_ray and trace_s > 0:
        roof["achieved"] = b_ray * rank0_closest / trace_s / 1e9
        roof["frac"] = roof["achieved"] / peak
        roof["frac_basis"] = ("ALGORITHMIC bytes (B_ray of the reference's traversal, SURVEY 8(d)) / launch time / "
                              "peak; the compressed wide tree moves fewer real bytes, see dram_frac")
        roof["bytes_per_ray"] = b_ray
        roof["b_ray_source"] = entry.get("ray_source", "sampled camera + one-bounce rays (round 1)")
        roof["algorithmic_bytes_per_launch"] = b_ray * rank0_closest / max(kt["trace_launches"] / n_k, 1)
        if traffic:
            roof["dram_frac"] = float(traffic) / avg_launch_s / 1e9 / peak
            roof["traffic_unit"] = ("bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum, "
                                    + str(entry.get("ncu_source", "profiles/")) + ")")
        if workload in L2_RESIDENT and l2_gbs:
            roof.update({"bound": "l2", "peak": l2_gbs, "frac": roof["achieved"] / l2_gbs,
                         "peak_source": "pb2_measure_read_gbs: 32 MiB resident buffer, ld.global.cg, measured in this run",
                         "frac_vs_hbm": roof["achieved"] / peak})
    else:
        roof["achieved"], roof["frac"] = None, None
    roof["launches_per_step"] = kt["trace_launches"] / n_k
    roof["avg_launch_ms"] = avg_launch_s * 1e3
    roof["timing"] = "per-kernel pass: lanes serialised, CUDA events around every launch on the launching stream"
    roof["mrays_per_s_in_kernel"] = rank0_closest / trace_s / 1e6 if trace_s > 0 else None
    shares = {k: kt[k + "_ms"] / max(kt["render_ms"], 1e-9) for k in ("trace", "shadow", "shade", "other")}

    per_rank_h2d = scene_bytes if world == 1 else int(scene_bytes / world)
    out = {
        "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mrays_per_s": (closest_per_step + shadow_per_step) / (ms_per_step * 1e-3) / 1e6,
        "rays_per_step": {"closest": closest_per_step, "shadow": shadow_per_step},
        "rays_per_sample": (closest_per_step + shadow_per_step) / total_samples,
        "closest_fallback_share": kt["fb_closest"] / max(kt["closest"], 1),
        # `config` has the same keys and values in both arms (this one and --impl reference)
        "config": _config(workload, w, h, spp, depth),
        "details": {"triangles_and_shapes": c0["n_primitives"], "bvh2_nodes": c0["bvh2_nodes"],
                   "wide_nodes": c0["wide_nodes"], "sharding": f"samples {spp}//{world} per GPU, scene replicated"
                   + (", one NCCL reduce(sum) of the float film per step" if world > 1 else ""),
                   "waves_in_flight": int(os.environ.get("PB2_LANES", "4")),
                   "l2": f"no flush: scene {c0['scene_bytes'] / 1e6:.0f} MB + wavefront state of up to 4 x "
                         f"{min(w * h * spp, 1 << 24)} slots exceed the 126 MB L2",
                   "host_bvh_build_ms": c0["bvh_build_ms"], "host_build_threads": c0["build_threads"],
                   "upload_scene_s": upload_s, "film_sum_rgb": film_sum},
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "ms_per_step": e2e_ms_per_step,
                "h2d_bytes_per_step": int(per_rank_h2d), "d2h_bytes_per_step": int(w * h * 16),
                "h2d_note": ("whole scene through PCIe" if world == 1 else
                             f"per rank: 1/{world} of the {scene_bytes} B replica through PCIe, the other "
                             f"{world - 1}/{world} arrive by in-place all-gather over NVLink"),
                "what": ("pb2_restage_scene + pb2_render with a pinned HOST film buffer" if world == 1 else
                         "pb2_restage_scene_part + all-gather, pb2_render_device, reduce, film to pinned host memory")
                        + ", every step; host BVH build not repeated (see e2e_cold)"},
        "e2e_cold": {"value": total_samples / cold_s / 1e6, "unit": "Msamples/s", "seconds": cold_s,
                     "scene_description_s": t_scene,
                     "what": "ONE drop-in call as Integrator::render(Scene&) makes it: pb2_upload_scene (host SAH "
                             "build + wide layout + H2D) + render + film to the host, first call on a fresh context "
                             "(allocations included); max over ranks"
                             + ("" if world == 1 else "; rank 0 builds once and shares the built scene with the other "
                                "ranks through /dev/shm (pb2_export_host_scene / pb2_import_host_scene)")},
        "gpu_launches": int(launches),
        "clocks": clock_info,
        "roofline": roof,
        "kernel_time_shares": shares,
    }
    B.close()
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torchrun")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the pine_b200 path has no CPU fallback")
    # stdout carries exactly ONE JSON line. NCCL writes its log (NCCL_DEBUG=INFO / VERSION) to the process's
    # stdout, so file descriptor 1 is pointed at stderr for everything else and the JSON line goes to a
    # private copy of the original stdout. NCCL_DEBUG itself is left as the caller set it.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    # render, staging copies and collectives share ONE non-default stream (the context's own stream is
    # non-blocking and NCCL orders against torch's current stream only)
    ts = torch.cuda.Stream(device=dev)

    out = measure(args, args.workload, args.steps, args.warmup, world, rank, dev, ts)
    if rank == 0 and world == 1 and not args.no_scenes:
        import pine_b200 as pb
        Bm = pb.Backend(local_rank)
        l2_gbs = Bm.measure_read_gbs(32 << 20, 200)
        hbm_gbs = Bm.measure_read_gbs(4 << 30, 4)
        Bm.close()
        out["measured_read_gbs"] = {"l2_32MiB_resident": l2_gbs, "hbm_4GiB_stream": hbm_gbs,
                                    "how": "pb2_measure_read_gbs: full grid, 128-bit ld.global.cg, CUDA events"}
        extras = {}
        for wl in ("c1", "c3", "c4"):
            if wl == args.workload:
                continue
            r = measure(args, wl, max(2, min(args.steps, 3)), 3, 1, 0, dev, ts, l2_gbs=l2_gbs, main=False)
            keep = {k: r[k] for k in ("value", "unit", "ms_per_step", "mrays_per_s", "rays_per_sample",
                                      "closest_fallback_share", "gpu_launches", "kernel_time_shares")}
            keep["e2e"] = r["e2e"]["value"]
            keep["e2e_cold_s"] = r["e2e_cold"]["seconds"]
            keep["roofline"] = r["roofline"]
            keep["config"] = r["config"]
            if not args.no_cpu_baseline:
                try:
                    cb = run_reference(wl, 1, 0, max(2.0, args.cpu_seconds / 4))
                    keep["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
                except Exception as e:
                    keep["cpu_baseline"] = {"value": None, "sample": f"unavailable: {e}"}
            extras[wl] = keep
        out["scenes"] = extras
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            try:
                out["cpu_baseline"] = run_reference(args.workload, 1, 0, args.cpu_seconds)
                # beside it: the same reference with the accel PRL really uses (faster than pine::BVH; not bit-exact)
                out["cpu_baseline_embree"] = run_reference_embree(args.workload, max(3.0, args.cpu_seconds / 2))
            except Exception as e:  # the checker is missing: say so, do not fail the GPU number
                out["cpu_baseline"] = {"value": None, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
                                       "sample": f"unavailable: {e}"}
        real_stdout.write(json.dumps(out) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.barrier()
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
    ap.add_argument("--no-scenes", action="store_true", help="skip the C1/C3/C4 lines of the `scenes` key")
    ap.add_argument("--accel", default="bvh", choices=["bvh", "embree"], help="--impl reference: which Accel")
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        r = run_reference(args.workload, args.steps, args.warmup, args.ref_step_seconds, args.accel)
        out = {"impl": "reference", "metric": "Msamples/s", "value": r["value"], "unit": "Msamples/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": _config(args.workload, r["film"][0], r["film"][1], r["spp"], r["max_path_length"]),
               "details": {"constructed_as": "PathIntegrator(BVH(), BlueSobolSampler(spp), UniformLightSampler(), depth): "
                                             "pine::BVH, the bit-exact oracle accel (the PRL default EmbreeAccel is "
                                             "not buildable in this image, BASELINE.md)"},
               "cpu_baseline": r,
               "e2e": {"value": r["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(out), flush=True)
        return
    run_gpu(args)


if __name__ == "__main__":
    main()
