"""GPU debugging: where do wide and reference-order renders of a scene part ways?"""
import sys
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import common
import pine_b200 as pb

name = sys.argv[1] if len(sys.argv) > 1 else "cones_48"
S = pb.Scene(); common.build(S, name)
B = pb.Backend(0); B.upload_scene(S)
_, _, spp, depth = common.SMALL_SCENES[name]
f0, r0 = B.render_recorded(spp, depth, accel=0)
f1, r1 = B.render_recorded(spp, depth, accel=1)
print("film equal:", np.array_equal(f0.view(np.uint32), f1.view(np.uint32)))
print("n closest", len(r0["closest_t"]), len(r1["closest_t"]), "n shadow", len(r0["shadow_occluded"]), len(r1["shadow_occluded"]))
def key(r):  # rays are compacted in warp order -> sort by ray bits to compare as sets
    return np.lexsort(r.view(np.uint32).T[::-1])
out = {}
for r in (r0, r1):
    for kind, fields in (("closest", ("closest_t", "closest_geom", "closest_prim")), ("shadow", ("shadow_occluded",))):
        k = key(r[kind + "_rays"])
        r[kind + "_rays"] = r[kind + "_rays"][k]
        for f in fields:
            r[f] = r[f][k]
for kind in ("closest", "shadow"):
    a, b = r0[kind + "_rays"], r1[kind + "_rays"]
    n = min(len(a), len(b))
    # the queues of both renders are identical until the first divergence; compare depth by depth is
    # not needed: walk in order and report the first mismatches
    same_rays = np.array_equal(a[:n].view(np.uint32), b[:n].view(np.uint32))
    print(kind, "ray queues identical:", same_rays and len(a) == len(b))
    if kind == "closest":
        bad = np.nonzero((r0["closest_t"][:n].view(np.uint32) != r1["closest_t"][:n].view(np.uint32)) |
                         (r0["closest_geom"][:n] != r1["closest_geom"][:n]) |
                         (a[:n].view(np.uint32) != b[:n].view(np.uint32)).any(axis=1))[0]
    else:
        bad = np.nonzero((r0["shadow_occluded"][:n] != r1["shadow_occluded"][:n]) |
                         (a[:n].view(np.uint32) != b[:n].view(np.uint32)).any(axis=1))[0]
    print(kind, "first mismatches:", bad[:10], "of", len(bad))
    for i in bad[:5]:
        print("  ray", a[i].tolist(), "hex", [hex(x) for x in a[i].view(np.uint32)])
        if kind == "closest":
            print("   wide:", r0["closest_geom"][i], r0["closest_prim"][i], r0["closest_t"][i],
                  " ref:", r1["closest_geom"][i], r1["closest_prim"][i], r1["closest_t"][i], " same ray:", np.array_equal(a[i].view(np.uint32), b[i].view(np.uint32)))
            h0 = B.trace_closest(a[i:i+1], accel=0); h1 = B.trace_closest(a[i:i+1], accel=1)
            print("   hook wide:", h0["geom"], h0["t"], " hook ref:", h1["geom"], h1["t"])
        else:
            print("   wide:", r0["shadow_occluded"][i], " ref:", r1["shadow_occluded"][i])
            print("   hook wide:", B.trace_any(a[i:i+1], accel=0), " hook ref:", B.trace_any(a[i:i+1], accel=1))
    out[kind + "_bad_rays"] = a[bad[:64]]
for k, v in r0.items(): out["w_" + k] = v
for k, v in r1.items(): out["r_" + k] = v
np.savez("gpurun_out/dbg_%s.npz" % name, **out)
