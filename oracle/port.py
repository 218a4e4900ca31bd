"""oracle/port.py -- TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle_port.so (oracle_port.cpp): the plain-C++ restatement of the
bit-exact part of the path (RNG, sampler, camera, spawn_ray, shapes, binned-SAH builder, the
reference-order closest/any walks with per-ray node and primitive-test counters). Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle_port.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.port_scene_new.restype = C.c_void_p
        L.port_scene_free.argtypes = [C.c_void_p]
        L.port_add_shape.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.port_add_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.port_build.argtypes = [C.c_void_p]
        L.port_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 5
        L.port_trace_any.argtypes = [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 3
        L.port_bvh_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.port_bvh_nodes.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.port_hash_pixel.restype = C.c_uint64
        L.port_rng_u64.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        L.port_bluenoise.restype = C.c_float
        L.port_bluenoise.argtypes = [C.c_void_p] * 3 + [C.c_int] * 4
        L.port_sampler_stream.argtypes = [C.c_void_p] * 3 + [C.c_int] * 4 + [C.c_void_p, C.c_char_p, C.c_int,
                                                                             C.c_void_p]
        L.port_gen_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.port_primary_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.port_spawn_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        _lib = L
    return _lib


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class PortScene:
    """Built from the flattened records of a `pine_b200.Scene` (the same numbers the C ABI gets)."""

    def __init__(self, scene):
        self.L = lib()
        self.h = C.c_void_p(self.L.port_scene_new())
        inv = {v: k for k, v in __import__("pine_b200")._abi.SHAPE_KINDS.items()}
        for kind, index, _mat in scene.geometries:
            name = inv[kind]
            if name == "mesh":
                v, i, _n, _t = scene.mesh_data[index]
                self.L.port_add_mesh(self.h, v.ctypes.data, len(v), i.ctypes.data, len(i))
            else:
                rec = scene.shapes[name][index]
                self.L.port_add_shape(self.h, kind, C.byref(rec))
        cam = scene._cam
        self.size = scene.size
        self.cam = _f(list(cam.position) + list(cam.c2w) + list(cam.fov2d) +
                      [cam.len_radius, cam.focus_distance]) if cam is not None else None
        self.L.port_build(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.port_scene_free(self.h)
            self.h = None

    def trace_closest(self, rays):
        rays = _f(rays).reshape(-1, 8)
        n = len(rays)
        geom, prim = np.zeros(n, np.int32), np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        nn, npt = np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.L.port_trace_closest(self.h, rays.ctypes.data, n, geom.ctypes.data, prim.ctypes.data,
                                  t.ctypes.data, nn.ctypes.data, npt.ctypes.data)
        return {"geom": geom, "prim": prim, "t": t, "n_nodes": nn, "n_prim_tests": npt}

    def trace_any(self, rays):
        rays = _f(rays).reshape(-1, 8)
        n = len(rays)
        occ = np.zeros(n, np.uint8)
        nn, npt = np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.L.port_trace_any(self.h, rays.ctypes.data, n, occ.ctypes.data, nn.ctypes.data, npt.ctypes.data)
        return {"occluded": occ, "n_nodes": nn, "n_prim_tests": npt}

    def bvh_tree(self, which=-1):
        root, npr = C.c_int32(), C.c_int32()
        n = self.L.port_bvh_info(self.h, which, C.byref(root), C.byref(npr))
        aabbs = np.zeros((n, 12), np.float32)
        children = np.zeros((n, 2), np.int32)
        leaf_range = np.zeros((n, 2), np.int32)
        leaf_prims = np.zeros(max(npr.value, 1), np.int32)
        if n:
            self.L.port_bvh_nodes(self.h, which, aabbs.ctypes.data, children.ctypes.data,
                                  leaf_range.ctypes.data, leaf_prims.ctypes.data)
        return {"root": root.value, "aabbs": aabbs, "children": children, "leaf_range": leaf_range,
                "leaf_prims": leaf_prims[:npr.value]}

    def gen_rays(self, pfilm_u2):
        a = _f(pfilm_u2).reshape(-1, 4)
        rays = np.zeros((len(a), 8), np.float32)
        self.L.port_gen_rays(self.cam.ctypes.data, a.ctypes.data, len(a), rays.ctypes.data)
        return rays

    def primary_rays(self, si=0):
        w, h = self.size
        rays = np.zeros((w * h, 8), np.float32)
        self.L.port_primary_rays(self.cam.ctypes.data, w, h, si, rays.ctypes.data)
        return rays


def hash_pixel(px, py, si):
    return int(lib().port_hash_pixel(px, py, si))


def rng_u64(seed, n):
    out = np.zeros(n, np.uint64)
    lib().port_rng_u64(C.c_uint64(seed), n, out.ctypes.data)
    return out


def bluenoise(tables, px, py, index, dim):
    sob, scr, rnk = tables
    return lib().port_bluenoise(sob.ctypes.data, scr.ctypes.data, rnk.ctypes.data, px, py, index, dim)


def sampler_stream(tables, px, py, sample_index, n_rand, pattern, n_samples):
    sob, scr, rnk = tables
    width = sum(1 if c == "1" else 2 for c in pattern)
    r = np.zeros(max(n_rand, 1), np.float32)
    d = np.zeros(max(width * n_samples, 1), np.float32)
    lib().port_sampler_stream(sob.ctypes.data, scr.ctypes.data, rnk.ctypes.data, px, py, sample_index,
                              n_rand, r.ctypes.data, pattern.encode(), n_samples, d.ctypes.data)
    return r[:n_rand], d[:width * n_samples].reshape(n_samples, width)


def spawn_rays(p_n_wo_dist):
    a = _f(p_n_wo_dist).reshape(-1, 10)
    rays = np.zeros((len(a), 8), np.float32)
    lib().port_spawn_rays(a.ctypes.data, len(a), rays.ctypes.data)
    return rays
