"""Small renders that touch every kernel family, for compute-sanitizer (memcheck / racecheck):
  compute-sanitizer --tool memcheck python scripts/sanitize_small.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pine_b200 as pb
from pine_b200 import scenes
os.environ["PB2_MIN_LANE_SLOTS"] = "1"
for fn, kw, spp, depth in (("cbox_readme", dict(size=(64, 48)), 4, 4), ("materials_room", dict(size=(48, 32)), 2, 5),
                           ("lights_yard", dict(size=(48, 32)), 2, 4),
                           ("mesh_field", dict(n_instances=27, subdiv=2, seed=1, size=(64, 36), with_shapes=True), 2, 4)):
    S = pb.Scene()
    scenes.SCENES[fn](S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    for lanes in ("1", "3"):
        os.environ["PB2_LANES"] = lanes
        f = B.render(spp, depth)
        assert np.isfinite(f).all()
    B.render(spp, depth, accel=1)
    B.render(spp, depth, bands=(8, 3, 1))
    B.restage()
    B.render(spp, depth, samples=(1, spp))
    B.finalize_device(want_mapped=True)
    rays = B.primary_rays(0)
    B.trace_closest(rays, surf=True); B.trace_any(rays)
    B.close()
    print(fn, "ok", flush=True)
