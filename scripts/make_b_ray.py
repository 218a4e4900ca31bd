"""Algorithmic bytes per traced ray, B_ray, for the bench workloads (DESIGN.md "Roofline").

    B_ray = 32 (ray) + 16 (hit record; 4 for a shadow ray) + n_node * 56 + sum of primitive bytes

with n_node = binary BVH nodes visited and the primitive tests made by the REFERENCE traversal
(BVHImpl::Intersect / ::hit, bvh.cpp:321-451), counted by the oracle port (oracle/oracle_port.cpp,
pinned against the compiled reference by tests/test_oracle.py) on a sample of the workload's own
rays. With a CUDA device (the GPU box: `python scripts/make_b_ray.py --recorded [workload ...]`) the rays are
the ones a production render of the workload actually traces -- every closest-hit and shadow ray of all depths
of a reduced-film, 1-spp recorded render (pb2_set_ray_record), so the depth mix is the render's own. Without
one: camera rays of sample 0 on a pixel grid, plus one random bounce ray and one shadow-length ray from each
hit point (round 1's method). Writes profiles/b_ray.json, which bench.py reads.
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import common  # noqa: E402
import pine_b200 as pb  # noqa: E402
from oracle import port  # noqa: E402

OUT = os.path.join(ROOT, "profiles", "b_ray.json")
GRID = 160  # sample GRID x GRID camera rays


def totals():
    t = np.zeros(3, np.int64)
    port.lib().port_last_totals(t.ctypes.data_as(C.c_void_p))
    return t


def one(workload):
    S = pb.Scene()
    bench.build_scene(S, workload)
    w, h = S.size
    t0 = time.time()
    P = port.PortScene(S)
    build_s = time.time() - t0
    rays = P.primary_rays(0).reshape(h, w, 8)
    ys = np.linspace(0, h - 1, min(GRID, h)).astype(int)
    xs = np.linspace(0, w - 1, min(GRID, w)).astype(int)
    prim = np.ascontiguousarray(rays[np.ix_(ys, xs)].reshape(-1, 8))
    hp = P.trace_closest(prim)
    tp = totals()
    # bounce rays from the hit points (normals are not needed for counting: offset along -d)
    hit = hp["geom"] >= 0
    o = prim[hit, :3] + prim[hit, 3:6] * hp["t"][hit, None] * np.float32(1 - 1e-4)
    rng = np.random.RandomState(7)
    d = rng.normal(size=(len(o), 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.concatenate([o, d, np.zeros((len(o), 1), np.float32), np.full((len(o), 1), 3.0e38, np.float32)],
                         axis=1).astype(np.float32)
    P.trace_closest(sec)
    ts = totals()
    sh = sec.copy()
    sh[:, 7] = rng.uniform(0.5, 0.5 * float(np.linalg.norm(prim[:, :3].max(0) - o.min(0)) + 1), len(sh))
    P.trace_any(sh)
    ta = totals()
    n_c = len(prim) + len(sec)
    nodes_c, tests_c, bytes_c = (tp + ts)
    e = {
        "rays_sampled": {"camera": len(prim), "bounce": len(sec), "shadow": len(sh)},
        "closest": {"nodes_per_ray": nodes_c / n_c, "prim_tests_per_ray": tests_c / n_c,
                    "prim_bytes_per_ray": bytes_c / n_c},
        "shadow": {"nodes_per_ray": ta[0] / len(sh), "prim_tests_per_ray": ta[1] / len(sh),
                   "prim_bytes_per_ray": ta[2] / len(sh)},
        "b_ray_closest": 32 + 16 + 56 * nodes_c / n_c + bytes_c / n_c,
        "b_ray_shadow": 32 + 4 + 56 * ta[0] / len(sh) + ta[2] / len(sh),
        "oracle_build_s": build_s,
    }
    return {k: (float(v) if isinstance(v, (np.floating, np.integer)) else v) for k, v in e.items()}


def recorded(workload, max_rays=1_500_000):
    """B_ray from the rays of a recorded production render at a reduced film size (same camera, same depth)."""
    S = pb.Scene()
    cfg = bench.build_scene(S, workload)
    w, h = S.size
    depth = cfg["depth"]
    scale = max(1, int(np.ceil(np.sqrt(w * h * (2 * depth - 1) / max_rays))))
    fn, kw, _ = bench.WORKLOADS[workload]
    kw = dict(kw, size=(max(16, w // scale), max(16, h // scale)))
    from pine_b200 import scenes
    S = pb.Scene()
    scenes.SCENES[fn](S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    _, rec = B.render_recorded(1, depth, accel=0)
    B.close()
    t0 = time.time()
    P = port.PortScene(S)
    build_s = time.time() - t0
    P.trace_closest(rec["closest_rays"])
    tc = totals()
    P.trace_any(rec["shadow_rays"])
    ta = totals()
    n_c, n_s = len(rec["closest_rays"]), max(len(rec["shadow_rays"]), 1)
    return {
        "ray_source": f"every ray of a recorded 1-spp production render at {kw['size'][0]}x{kw['size'][1]}, depth {depth} "
                      f"({n_c} closest-hit + {n_s} shadow rays, all depths)",
        "rays_sampled": {"closest": n_c, "shadow": n_s, "camera_share_of_closest": kw["size"][0] * kw["size"][1] / n_c},
        "closest": {"nodes_per_ray": tc[0] / n_c, "prim_tests_per_ray": tc[1] / n_c, "prim_bytes_per_ray": tc[2] / n_c},
        "shadow": {"nodes_per_ray": ta[0] / n_s, "prim_tests_per_ray": ta[1] / n_s, "prim_bytes_per_ray": ta[2] / n_s},
        "b_ray_closest": float(32 + 16 + 56 * tc[0] / n_c + tc[2] / n_c),
        "b_ray_shadow": float(32 + 4 + 56 * ta[0] / n_s + ta[2] / n_s),
        "oracle_build_s": build_s,
    }


def main():
    argv = [a for a in sys.argv[1:] if a != "--recorded"]
    use_recorded = "--recorded" in sys.argv[1:]
    global OUT
    if os.environ.get("PB2_B_RAY_OUT"):
        OUT = os.environ["PB2_B_RAY_OUT"]
    names = argv or ["c1", "c3", "c4", "c4_field", "c5"]
    out = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for n in names:
        e = recorded(n) if use_recorded else one(n)
        e = json.loads(json.dumps(e, default=float))
        old = out.get(n, {})
        for keep in ("ncu_dram_bytes_per_launch", "ncu_source"):
            if keep in old:
                e[keep] = old[keep]
        out[n] = e
        if n == "c5":
            out["c5s"] = dict(e)  # same geometry, smaller film
        print(n, json.dumps(e))
        json.dump(out, open(OUT, "w"), indent=1, default=float)


if __name__ == "__main__":
    main()
