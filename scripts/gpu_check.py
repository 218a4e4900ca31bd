"""First-light GPU check: traversal parity (exact + wide) and film parity on C1 against the reference
harness, with timings. Writes a summary to gpurun_out/gpu_check.json."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import ref
import pine_b200 as pb
from pine_b200 import scenes

out = {}
name = sys.argv[1] if len(sys.argv) > 1 else "cbox_readme"
kw = {}
if name == "mesh_field":
    kw = dict(n_instances=int(os.environ.get("NI", 64)), subdiv=int(os.environ.get("SUB", 3)))
R = ref.RefScene(); cfg = scenes.SCENES[name](R, **kw)
S = pb.Scene(); scenes.SCENES[name](S, **kw)
spp = int(os.environ.get("SPP", cfg["spp"])); depth = int(os.environ.get("DEPTH", cfg["depth"]))
B = pb.Backend(0); B.upload_scene(S)
c = B.counters(); out["bvh"] = {k: c[k] for k in ("bvh2_nodes", "wide_nodes", "n_primitives", "bvh_build_ms")}

# T2: sampler / RNG
r_ref, d_ref = ref.sampler_stream(spp if spp in (1,2,4,8,16,32,64,128,256) else 16, 37, 411, 0, 16, "21221", 20)
r_gpu, d_gpu = B.test_sampler(spp, 37, 411, 0, 16, "21221", 20)
out["rng_equal"] = bool(np.array_equal(r_ref, r_gpu)); out["sampler_equal"] = bool(np.array_equal(d_ref, d_gpu))
# T3: primary rays
pr_ref = R.primary_rays(0); pr_gpu = B.primary_rays(0)
out["primary_rays_bitexact"] = bool(np.array_equal(pr_ref.view(np.uint32), pr_gpu.view(np.uint32)))
out["primary_rays_mismatch"] = int((pr_ref.view(np.uint32) != pr_gpu.view(np.uint32)).any(axis=1).sum())
# T1: closest hit
t0 = time.time(); hr = R.trace_closest(pr_ref, surf=True); out["ref_trace_s"] = time.time() - t0
for accel, tag in ((1, "exact"), (0, "wide")):
    hg = B.trace_closest(pr_ref, accel=accel, surf=True)
    out[f"closest_{tag}_geom_mismatch"] = int((hr["geom"] != hg["geom"]).sum())
    out[f"closest_{tag}_prim_mismatch"] = int((hr["prim"] != hg["prim"]).sum())
    out[f"closest_{tag}_t_mismatch"] = int((hr["t"].view(np.uint32) != hg["t"].view(np.uint32)).sum())
    out[f"closest_{tag}_surf_mismatch"] = int((hr["surf"].view(np.uint32) != hg["surf"].view(np.uint32)).any(axis=1).sum())
# secondary rays: spawn from hits in random directions
rng = np.random.RandomState(1)
hit = hr["geom"] >= 0
n = int(hit.sum())
d = rng.normal(size=(n, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
inp = np.concatenate([hr["surf"][hit][:, :6], d, np.full((n, 1), 3.0e38, np.float32)], axis=1).astype(np.float32)
sec = ref.spawn_rays(inp)
hr2 = R.trace_closest(sec)
for accel, tag in ((1, "exact"), (0, "wide")):
    hg2 = B.trace_closest(sec, accel=accel)
    out[f"secondary_{tag}_mismatch"] = int(((hr2["geom"] != hg2["geom"]) | (hr2["prim"] != hg2["prim"]) |
                                            (hr2["t"].view(np.uint32) != hg2["t"].view(np.uint32))).sum())
# shadow rays: toward a point near the first light
sh = sec.copy(); sh[:, 7] = rng.uniform(0.2, 3.0, size=n).astype(np.float32)
occ_ref = R.trace_any(sh)
for accel, tag in ((1, "exact"), (0, "wide")):
    out[f"any_{tag}_mismatch"] = int((occ_ref != B.trace_any(sh, accel=accel)).sum())
out["n_rays"] = [int(len(pr_ref)), n]

# film
film_ref, secs = R.render(spp, depth); out["ref_render_s"] = secs
for accel, tag in ((1, "exact"), (0, "wide")):
    t0 = time.time(); film = B.render(spp, depth, accel=accel); wall = time.time() - t0
    film = B.render(spp, depth, accel=accel)
    c = B.counters()
    a, b = film[..., :3].astype(np.float64), film_ref[..., :3].astype(np.float64)
    rel = np.abs(a - b) / (np.abs(b) + 1e-3)
    out[f"film_{tag}"] = {"first_wall_s": wall, "render_ms": c["last_render_ms"], "sum_gpu": float(a.sum()), "sum_ref": float(b.sum()),
                          "relmse": float(((a - b) ** 2 / (b ** 2 + 1e-2)).mean()),
                          "frac_pixels_rel>1e-3": float((rel.max(axis=2) > 1e-3).mean()),
                          "frac_pixels_rel>1e-5": float((rel.max(axis=2) > 1e-5).mean()),
                          "alpha_ok": bool((film[..., 3] == 1).all()),
                          "closest_rays": c["closest_rays"], "shadow_rays": c["shadow_rays"]}
    np.save(f"gpurun_out/film_{name}_{tag}.npy", film)
np.save(f"gpurun_out/film_{name}_ref.npy", film_ref)
print(json.dumps(out, indent=1))
json.dump(out, open(f"gpurun_out/gpu_check_{name}.json", "w"), indent=1)
