"""Shared helpers of the test-suite: the small parity scenes and deterministic ray sets."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

# name -> (scene function, kwargs, spp, depth)
SMALL_SCENES = {
    "c1_64": ("cbox_readme", dict(size=(64, 64)), 4, 4),
    "c3_48": ("cbox_indirect", dict(size=(48, 48)), 4, 5),
    "mesh_48": ("mesh_field", dict(n_instances=8, subdiv=2, size=(48, 27), spp=4, depth=5, with_shapes=True), 4, 5),
    "cones_48": ("cbox_readme_noboxes", dict(size=(48, 48)), 4, 4),
}


def build(target, name):
    from pine_b200 import scenes
    fn, kw, _, _ = SMALL_SCENES[name]
    return scenes.SCENES[fn](target, **kw)


def u32(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def secondary_inputs(surf, geom, seed):
    """spawn_ray inputs {p, n, wo, distance} from the hit points of a closest-hit result."""
    rng = np.random.RandomState(seed)
    hit = geom >= 0
    n = int(hit.sum())
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = np.full((n, 1), 3.0e38, np.float32)
    return np.concatenate([surf[hit][:, :6], d, dist], axis=1).astype(np.float32)


def shadow_from(sec, seed):
    rng = np.random.RandomState(seed)
    sh = sec.copy()
    sh[:, 7] = rng.uniform(0.2, 3.0, size=len(sec)).astype(np.float32)
    return sh


def ray_fan(origin, n, seed):
    """n rays {o, d, tmin = 0, tmax} in uniformly random directions from one point."""
    rng = np.random.RandomState(seed)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3], rays[:, 3:6], rays[:, 7] = origin, d, 3.0e38
    return rays


def relmse(a, b, eps=1e-2):
    a, b = a[..., :3].astype(np.float64), b[..., :3].astype(np.float64)
    return float(((a - b) ** 2 / (b ** 2 + eps)).mean())
