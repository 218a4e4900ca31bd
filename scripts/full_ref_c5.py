#!/usr/bin/env python
"""ONE full-frame STOCK reference render of a bench workload -- PathIntegrator(BVH(), BlueSobolSampler(spp),
UniformLightSampler(), depth).render(scene), every row, BVH build included (oracle/ref.py RefScene.render =
ref_render) -- beside (a) the row-sampled reference arm bench.py times and (b) the CUDA path's film of the same
workload. Validates the sampled arm's Msamples/s and gives a film parity number at the benchmark's own size.
  python scripts/full_ref_c5.py [workload] > profiles/r2/full_ref_render_<workload>.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from oracle import ref

wl = sys.argv[1] if len(sys.argv) > 1 else "c5"
R = ref.RefScene()
cfg = bench.build_scene(R, wl)
w, h = R.size
spp, depth = cfg["spp"], cfg["depth"]
t_build = R.accel_build()                      # BVH().build(&scene) alone
film_ref, secs = R.render(spp, depth)          # stock render(): prologue (build again) + pixel loop
out = {"workload": bench.WORKLOADS[wl][2], "cores": int(ref.lib().ref_num_threads()),
       "stock_render_seconds": secs, "bvh_build_seconds_alone": t_build,
       "stock_msamples_per_s_incl_build": w * h * spp / secs / 1e6,
       "stock_msamples_per_s_excl_build": w * h * spp / max(secs - t_build, 1e-9) / 1e6}
sampled = bench.run_reference(wl, 1, 0, 6.0)
out["row_sampled_arm_msamples_per_s"] = sampled["value"]
out["row_sampled_arm_sample"] = sampled["sample"]
out["sampled_over_stock_excl_build"] = sampled["value"] / out["stock_msamples_per_s_excl_build"]
try:
    import pine_b200 as pb
    S = pb.Scene()
    bench.build_scene(S, wl)
    B = pb.Backend(0)
    B.upload_scene(S)
    t0 = time.perf_counter()
    film = B.render(spp, depth)
    out["gpu_render_seconds"] = time.perf_counter() - t0
    a, b = film[..., :3].astype(np.float64), film_ref[..., :3].astype(np.float64)
    out["film_relmse_gpu_vs_stock_reference"] = float(((a - b) ** 2 / (b ** 2 + 1e-2)).mean())
    out["film_sum_gpu"], out["film_sum_reference"] = float(a.sum()), float(b.sum())
    out["pixels_within_1e-3_rel"] = float((np.abs(a - b) <= 1e-3 * np.abs(b) + 1e-5).all(-1).mean())
    B.close()
except Exception as e:
    out["gpu"] = f"unavailable: {e}"
print(json.dumps(out))
