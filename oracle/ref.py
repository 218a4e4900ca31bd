"""oracle/ref.py -- TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_ref/libpine_ref.so: the UNMODIFIED reference renderer (wicstas/pine)
compiled in place by oracle/Makefile, behind the C harness oracle/ref_harness.cpp.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module, and only as the checker or the timed CPU baseline -- never on the product path.

`RefScene` exposes the same builder methods as `pine_b200.Scene` (add_material / add_rect / ...),
so one scene-description function (pine_b200/scenes.py) feeds both sides identical numbers.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libpine_ref.so")
# The same harness with the reference's real PRL-default accel (EmbreeAccel over the vendored Embree 4.3.1,
# `make -C oracle embree`): a TIMING baseline only, selected per process with PINE_REF_EMBREE=1 (the two
# libraries export the same symbols, so one process loads one of them).
EMBREE_LIB_PATH = os.path.join(_HERE, "_ref", "libpine_ref_embree.so")
if os.environ.get("PINE_REF_EMBREE") == "1":
    LIB_PATH = EMBREE_LIB_PATH

SHIM_PATH = os.path.join(_HERE, "_ref", "libpine_shim.so")
_lib = None
_shim = None


def shim():
    """oracle/_ref/libpine_shim.so: pine::CudaPathIntegrator (integration/) + its C entry points. The only
    library on the oracle side that links the product; loaded by the drop-in tests alone."""
    global _shim
    if _shim is None:
        lib()  # libpine_ref.so first: the shim resolves Pine's classes from it
        S = C.CDLL(SHIM_PATH)
        S.ref_render_cuda_kind.restype = C.c_double
        S.ref_render_cuda_kind.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        S.ref_render_cuda_rgba8.restype = C.c_double
        S.ref_render_cuda_rgba8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _shim = S
    return _shim


def available():
    return os.path.exists(LIB_PATH)


def embree_available():
    return os.path.exists(EMBREE_LIB_PATH)


def use_embree_accel(on=True):
    """bounded_prepare / bounded_render_rows then run PathIntegrator(EmbreeAccel(), ...) (needs PINE_REF_EMBREE=1)."""
    L = lib()
    if on and not L.ref_has_embree():
        raise RuntimeError("this process loaded libpine_ref.so (pine::BVH only): set PINE_REF_EMBREE=1")
    L.ref_set_bounded_accel(1 if on else 0)


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_scene_new.restype = C.c_void_p
        L.ref_scene_free.argtypes = [C.c_void_p]
        L.ref_render.restype = C.c_double
        L.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ref_accel_build.restype = C.c_double
        L.ref_render_sample_range.restype = C.c_double
        L.ref_render_sample_range.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ref_render_moments.restype = C.c_double
        L.ref_render_moments.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_bounded_prepare.restype = C.c_double
        L.ref_bounded_prepare.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_bounded_render_rows.restype = C.c_double
        L.ref_bounded_render_rows.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ref_accel_build.argtypes = [C.c_void_p]
        L.ref_bluenoise.restype = C.c_float
        L.ref_hash_pixel.restype = C.c_uint64
        for name in ("ref_add_material", "ref_add_rect", "ref_add_box", "ref_add_aabb",
                     "ref_add_cone", "ref_add_sphere", "ref_add_disk", "ref_add_triangle", "ref_add_plane", "ref_add_cylinder",
                     "ref_add_mesh", "ref_num_geometries", "ref_num_lights", "ref_bvh_info",
                     "ref_bvh_num_lbvh", "ref_bvh_num_indices", "ref_load_gltf"):
            getattr(L, name).restype = C.c_int
        _lib = L
    return _lib


MATERIAL_KINDS = {"emissive": 0, "diffuse": 1, "metal": 2, "glossy": 3, "glass": 4, "uber": 5,
                  "subsurface": 6}


def pack_material(color, roughness=0.0, ior=1.4, metallic=0.0, transmission=0.0,
                  sigma_s=(0.0, 0.0, 0.0)):
    return _f([*color, roughness, ior, metallic, transmission, *sigma_s])


def _program_arrays(prog):
    ops = np.ascontiguousarray([i[0] for i in prog], np.int32)
    args = _f([c for i in prog for c in i[1:]])
    return ops, args


def eval_node(node, ctx8):
    """The reference's own evaluation, Mnode<T>::operator()(NodeEvalCtx{p, n, uv}), of a pine_b200.nodes graph at
    the points ctx8 = n x {p.xyz, n.xyz, uv} -> n x 3 (a Nodef's value in all three components)."""
    prog = node.program() + [(0, 0.0, 0.0, 0.0)]
    ops, args = _program_arrays(prog)
    c = _f(ctx8).reshape(-1, 8)
    out = np.zeros((len(c), 3), np.float32)
    rc = lib().ref_eval_node(ops.ctypes.data_as(C.POINTER(C.c_int)), _fp(args), len(prog), 1, len(c),
                             _fp(c), _fp(out))
    if rc:
        raise ValueError(f"ref_eval_node failed ({rc})")
    return out


def mat4(op, args):
    out = np.zeros(16, np.float32)
    a = _f(args)
    lib().ref_mat4(C.c_int(op), _fp(a), _fp(out))
    return out.reshape(4, 4)


def mat4_mul(a, b):
    out = np.zeros(16, np.float32)
    a, b = _f(a), _f(b)
    lib().ref_mat4_mul(_fp(a), _fp(b), _fp(out))
    return out.reshape(4, 4)


class RefScene:
    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.ref_scene_new())
        self.size = None

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_scene_free(self.h)
            self.h = None

    # ---- builder interface (mirrors pine_b200.Scene) ----
    def add_material(self, kind, color, roughness=0.0, ior=1.4, metallic=0.0, transmission=0.0,
                     sigma_s=(0.0, 0.0, 0.0)):
        from pine_b200 import nodes as N
        params = dict(color=color, roughness=roughness, ior=ior, metallic=metallic, transmission=transmission)
        prog, starts = [], []
        for slot in ("color", "roughness", "ior", "metallic", "transmission"):   # pb2_material.*_node order
            if N.is_graph(params[slot]):
                # the reference builds its OWN node objects from the typed program (ref_harness.cpp build_node)
                starts.append(len(prog) + 1)
                prog += params[slot].program() + [(0, 0.0, 0.0, 0.0)]
                params[slot] = (0.0, 0.0, 0.0) if slot == "color" else 0.0
            else:
                starts.append(0)
                params[slot] = N.constant_of(params[slot], slot != "color")
        p = pack_material(params["color"], params["roughness"], params["ior"], params["metallic"],
                          params["transmission"], sigma_s)
        if not prog:
            return self.L.ref_add_material(self.h, MATERIAL_KINDS[kind], _fp(p))
        ops, args = _program_arrays(prog)
        st = (C.c_int * 5)(*starts)
        rc = self.L.ref_add_material_nodes(self.h, MATERIAL_KINDS[kind], _fp(p), ops.ctypes.data_as(C.POINTER(C.c_int)),
                                           _fp(args), len(prog), st)
        if rc < 0:
            raise ValueError(f"ref_add_material_nodes failed ({rc})")
        return rc

    def add_rect(self, pos, ex, ey, mat, flip=False):
        a, b, c = _f(pos), _f(ex), _f(ey)
        return self.L.ref_add_rect(self.h, _fp(a), _fp(b), _fp(c), int(flip), mat)

    def add_box(self, lo, hi, m, mat):
        a, b, c = _f(lo), _f(hi), _f(m).reshape(-1)
        return self.L.ref_add_box(self.h, _fp(a), _fp(b), _fp(c), mat)

    def add_aabb(self, lo, hi, mat):
        a, b = _f(lo), _f(hi)
        return self.L.ref_add_aabb(self.h, _fp(a), _fp(b), mat)

    def add_cone(self, p, n, r, h, mat):
        a, b = _f(p), _f(n)
        return self.L.ref_add_cone(self.h, _fp(a), _fp(b), C.c_float(r), C.c_float(h), mat)

    def add_sphere(self, c, r, mat):
        a = _f(c)
        return self.L.ref_add_sphere(self.h, _fp(a), C.c_float(r), mat)

    def add_disk(self, p, n, r, mat):
        a, b = _f(p), _f(n)
        return self.L.ref_add_disk(self.h, _fp(a), _fp(b), C.c_float(r), mat)

    def add_plane(self, p, n, mat):
        a, b = _f(p), _f(n)
        return self.L.ref_add_plane(self.h, _fp(a), _fp(b), mat)

    def add_cylinder(self, p0, p1, r, mat):
        a, b = _f(p0), _f(p1)
        return self.L.ref_add_cylinder(self.h, _fp(a), _fp(b), C.c_float(r), mat)

    def add_triangle(self, a, b, c, mat):
        a, b, c = _f(a), _f(b), _f(c)
        return self.L.ref_add_triangle(self.h, _fp(a), _fp(b), _fp(c), mat)

    def add_mesh(self, verts, idx, mat, normals=None, uvs=None):
        v = _f(verts).reshape(-1, 3)
        i = np.ascontiguousarray(idx, dtype=np.uint32).reshape(-1, 3)
        n = _f(normals) if normals is not None else None
        t = _f(uvs) if uvs is not None else None
        return self.L.ref_add_mesh(self.h, _fp(v), len(v), i.ctypes.data_as(C.c_void_p), len(i),
                                   _fp(n) if n is not None else None,
                                   _fp(t) if t is not None else None, mat)

    def set_camera(self, size, frm, to, fov, len_radius=0.0, focus=1.0, tonemapper="uncharted2"):
        self.size = (int(size[0]), int(size[1]))
        a, b = _f(frm), _f(to)
        self.L.ref_set_camera(self.h, self.size[0], self.size[1],
                              0 if tonemapper == "uncharted2" else 1, _fp(a), _fp(b),
                              C.c_float(fov), C.c_float(len_radius), C.c_float(focus))

    def add_point_light(self, position, color):
        a, b = _f(position), _f(color)
        self.L.ref_add_point_light(self.h, _fp(a), _fp(b))

    def add_directional_light(self, direction, color):
        a, b = _f(direction), _f(color)
        self.L.ref_add_directional_light(self.h, _fp(a), _fp(b))

    def add_spot_light(self, position, direction, color, falloff_radian, cutoff_additional_radian=0.0):
        a, b, c = _f(position), _f(direction), _f(color)
        self.L.ref_add_spot_light(self.h, _fp(a), _fp(b), _fp(c), C.c_float(falloff_radian),
                                  C.c_float(cutoff_additional_radian))

    def set_sky(self, sun_color):
        a = _f(sun_color)
        self.L.ref_set_sky(self.h, _fp(a))

    def load_gltf(self, path, has_camera=False):
        """scene_from_gltf(scene, path) (fileio.cpp:146-327); `has_camera`: the file carries a camera node,
        so the film size is read back from the scene it set."""
        n = self.L.ref_load_gltf(self.h, path.encode())
        if has_camera:
            wh = (C.c_int * 2)()
            self.L.ref_camera_size(self.h, wh)
            self.size = (int(wh[0]), int(wh[1]))
        return n

    # ---- reference operations ----
    def render(self, spp, depth, sampler="blue"):
        """PathIntegrator(BVH(), sampler(spp), UniformLightSampler(), depth).render(scene)
        -> (film float32 [h, w, 4] pre-tonemap, wall seconds incl. BVH build)."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32)
        kind = {"blue": 0, "sobol": 1, "uniform": 2, "halton": 3}[sampler]
        secs = self.L.ref_render(self.h, kind, spp, depth, film.ctypes.data_as(C.c_void_p))
        return film, secs

    def render_cuda(self, spp, depth, device=0, sampler="blue"):
        """The drop-in test: THIS pine::Scene through pine::CudaPathIntegrator (integration/) ->
        (film, seconds, stats). The product library does the work; the harness only hosts the scene.
        device = -1: every visible GPU (the shim's default), else that one device."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32)
        stats = np.zeros(5, np.float64)
        kind = {"blue": 0, "sobol": 1, "uniform": 2, "halton": 3}[sampler]
        if device < 0:  # the shim's pb2_multi binds NCCL at run time: keep it the one a later `import torch` expects
            from pine_b200.integrator import _preload_bundled_nccl
            _preload_bundled_nccl()
        secs = shim().ref_render_cuda_kind(self.h, kind, spp, depth, device, film.ctypes.data, stats.ctypes.data)
        return film, secs, {"closest_rays": int(stats[0]), "shadow_rays": int(stats[1]),
                            "kernel_launches": int(stats[2]), "render_ms": float(stats[3]), "n_devices": int(stats[4])}

    def render_sample_range(self, spp, depth, s0, s1):
        """Samples [s0, s1) of a BlueSampler(spp) render with the sampler started at sample index s0: one
        rank's film of a sample-sharded render (sum / spp, alpha = (s1 - s0) / spp)."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32)
        self.L.ref_render_sample_range(self.h, spp, depth, s0, s1, film.ctypes.data)
        return film

    def render_moments(self, spp, depth):
        """BlueSampler render that also returns the per-pixel mean and the standard error of that mean, from
        the reference's own per-sample radiances -> (film, mean [h, w, 3] f64, sigma_of_mean [h, w, 3] f64)."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32)
        s1, s2 = np.zeros((h, w, 3), np.float64), np.zeros((h, w, 3), np.float64)
        self.L.ref_render_moments(self.h, spp, depth, film.ctypes.data, s1.ctypes.data, s2.ctypes.data)
        mean = s1 / spp
        var = np.maximum(s2 / spp - mean ** 2, 0.0) * (spp / max(spp - 1, 1))
        return film, mean, np.sqrt(var / spp)

    def render_cuda_rgba8(self, spp, depth, device=0):
        """pine::CudaPathIntegrator::render + ::finalize_rgba8 (the device epilogue of Film::save) on THIS scene
        -> (film float32, rgba8 uint8 [h, w, 4], y-flipped)."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32)
        rgba8 = np.zeros((h, w, 4), np.uint8)
        shim().ref_render_cuda_rgba8(self.h, spp, depth, device, film.ctypes.data, rgba8.ctypes.data)
        return film, rgba8

    def bounded_prepare(self, spp, depth):
        """RTIntegrator::render prologue (sampler clones, BVH::build, light list) of a
        PathIntegrator(BVH(), BlueSobolSampler(spp), UniformLightSampler(), depth); -> seconds."""
        return self.L.ref_bounded_prepare(self.h, spp, depth)

    def bounded_render_rows(self, row_begin, row_end, row_step=1, want_film=False):
        """The reference pixel loop (path.cpp:31-40) over rows row_begin::row_step < row_end only,
        all host threads; -> (wall seconds, film or None). Needs bounded_prepare first."""
        w, h = self.size
        film = np.zeros((h, w, 4), np.float32) if want_film else None
        secs = self.L.ref_bounded_render_rows(self.h, row_begin, row_end, row_step,
                                              film.ctypes.data_as(C.c_void_p) if want_film else None)
        if secs < 0:
            raise RuntimeError("bounded_prepare was not called")
        return secs, film

    def finalize(self, film):
        out = np.zeros_like(film)
        f = _f(film)
        self.L.ref_film_finalize(self.h, _fp(f), _fp(out))
        return out

    def accel_build(self):
        return self.L.ref_accel_build(self.h)

    def trace_closest(self, rays, surf=False, count_tests=False):
        rays = _f(rays).reshape(-1, 8)
        n = len(rays)
        geom = np.zeros(n, np.int32)
        prim = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        s = np.zeros((n, 8), np.float32) if surf else None
        k = np.zeros(n, np.int32) if count_tests else None
        self.L.ref_trace_closest(self.h, _fp(rays), C.c_int64(n), geom.ctypes.data_as(C.c_void_p),
                                 prim.ctypes.data_as(C.c_void_p), _fp(t),
                                 _fp(s) if surf else None,
                                 k.ctypes.data_as(C.c_void_p) if count_tests else None)
        out = {"geom": geom, "prim": prim, "t": t}
        if surf:
            out["surf"] = s
        if count_tests:
            out["n_prim_tests"] = k
        return out

    def trace_any(self, rays):
        rays = _f(rays).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint8)
        self.L.ref_trace_any(self.h, _fp(rays), C.c_int64(len(rays)),
                             occ.ctypes.data_as(C.c_void_p))
        return occ

    def primary_rays(self, si=0):
        w, h = self.size
        rays = np.zeros((h * w, 8), np.float32)
        self.L.ref_primary_rays(self.h, si, _fp(rays))
        return rays

    def gen_rays(self, pfilm_u2):
        a = _f(pfilm_u2).reshape(-1, 4)
        rays = np.zeros((len(a), 8), np.float32)
        self.L.ref_gen_rays(self.h, _fp(a), C.c_int64(len(a)), _fp(rays))
        return rays

    def geometry_aabbs(self):
        n = self.L.ref_num_geometries(self.h)
        out = np.zeros((n, 6), np.float32)
        self.L.ref_geometry_aabbs(self.h, _fp(out))
        return out

    def shape_sample(self, g, p_u2_u1):
        a = _f(p_u2_u1).reshape(-1, 6)
        out = np.zeros((len(a), 12), np.float32)
        self.L.ref_shape_sample(self.h, g, _fp(a), C.c_int64(len(a)), _fp(out))
        return out

    def bvh_tree(self, which=-1):
        """Reference BVHImpl node array (which=-1 top level, k>=0 per-mesh tree k)."""
        root = C.c_int32()
        npr = C.c_int32()
        n = self.L.ref_bvh_info(self.h, which, C.byref(root), C.byref(npr))
        aabbs = np.zeros((n, 12), np.float32)
        children = np.zeros((n, 2), np.int32)
        leaf_range = np.zeros((n, 2), np.int32)
        leaf_prims = np.zeros(max(npr.value, 1), np.int32)
        self.L.ref_bvh_nodes(self.h, which, _fp(aabbs), children.ctypes.data_as(C.c_void_p),
                             leaf_range.ctypes.data_as(C.c_void_p),
                             leaf_prims.ctypes.data_as(C.c_void_p))
        return {"root": root.value, "aabbs": aabbs, "children": children,
                "leaf_range": leaf_range, "leaf_prims": leaf_prims[:npr.value]}

    def bvh_indices(self):
        n = self.L.ref_bvh_num_indices(self.h)
        out = np.zeros(n, np.int32)
        self.L.ref_bvh_indices(self.h, out.ctypes.data_as(C.c_void_p))
        return out, self.L.ref_bvh_num_lbvh(self.h)


def bluenoise(spp, px, py, index, dim):
    return lib().ref_bluenoise(spp, px, py, index, dim)


def sampler_stream(spp, px, py, sample_index, n_rand, pattern, n_samples):
    width = sum(1 if c == "1" else 2 for c in pattern)
    r = np.zeros(max(n_rand, 1), np.float32)
    d = np.zeros(max(width * n_samples, 1), np.float32)
    lib().ref_sampler_stream(spp, px, py, sample_index, n_rand, _fp(r), pattern.encode(),
                             n_samples, _fp(d))
    return r[:n_rand], d[:width * n_samples].reshape(n_samples, width)


def sampler_stream_kind(sampler, spp, size, px, py, sample_index, pattern, n_samples):
    """get1d/get2d draws of Sampler(<sampler>(spp)) after init(size), start_pixel((px, py), sample_index)."""
    width = sum(1 if c == "1" else 2 for c in pattern)
    d = np.zeros(max(width * n_samples, 1), np.float32)
    kind = {"blue": 0, "sobol": 1, "uniform": 2, "halton": 3}[sampler]
    lib().ref_sampler_stream_kind(kind, spp, size[0], size[1], px, py, sample_index, pattern.encode(),
                                  n_samples, _fp(d))
    return d[:width * n_samples].reshape(n_samples, width)


def spawn_rays(p_n_wo_dist):
    a = _f(p_n_wo_dist).reshape(-1, 10)
    rays = np.zeros((len(a), 8), np.float32)
    lib().ref_spawn_rays(_fp(a), C.c_int64(len(a)), _fp(rays))
    return rays
