#!/usr/bin/env python
"""Wide-walk work per ray (needs a -DPB2_COUNT build selected with PINE_B200_LIB):
  PINE_B200_LIB=pine_b200/lib_cnt/libpine_b200.so python scripts/walk_stats.py c5s"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pine_b200 as pb
from bench import WORKLOADS, build_scene
wl = sys.argv[1] if len(sys.argv) > 1 else "c5s"
S = pb.Scene(); cfg = build_scene(S, wl)
B = pb.Backend(0); B.upload_scene(S)
out = (C.c_uint64 * 10)()
B.L.pb2_debug_walk_stats(B.h, out)
B.render(cfg["spp"], cfg["depth"])
B.L.pb2_debug_walk_stats(B.h, out)
v = [int(x) for x in out]
c = B.counters()
print("fallback reasons: box-touch %d, tie/stack in walk %d, tie at end %d, gate fail %d" % tuple(v[6:10]))
print("fallbacks closest %d shadow %d" % (c["last_closest_fallbacks"], c["last_shadow_fallbacks"]))
print(wl, "closest: nodes/ray %.2f prims/ray %.2f (rays %d) | any: nodes/ray %.2f prims/ray %.2f (rays %d) | render ms %.2f" % (
    v[0] / max(v[4], 1), v[1] / max(v[4], 1), v[4], v[2] / max(v[5], 1), v[3] / max(v[5], 1), v[5], c["last_render_ms"]))
