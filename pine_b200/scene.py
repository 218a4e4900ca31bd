"""Host-side scene model: the subset of Pine's scripting surface that feeds PathIntegrator.

`Scene` offers the builder methods `add_material / add_rect / add_box / add_cone / ... /
set_camera` (identical signatures on `oracle.ref.RefScene`, so one description feeds both sides)
and the PRL-flavoured front (`scene.add("floor", Diffuse(...))`, `scene.add(Rect(...), "floor")`,
`scene.set(ThinLenCamera(Film(size), frm, to, fov))`) mirroring program_context.cpp:23-125 and
scene.cpp:8-44 of the reference. All derived shape members are computed by the C++ host layer
(include/pine_b200_host.h), never in Python, so they carry the reference constructors' float ops.
Matrices are 4x4 numpy float32 in math (row, column) convention; `mat4.*` helpers call the C++ layer.
"""
import ctypes as C

import numpy as np

from . import _abi as A
from . import nodes as N


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class mat4:
    """translate / scale / rotate_* / inverse / look_at of vecmath.h:1088-1180, via the C++ host."""

    @staticmethod
    def _op(op, args):
        out = np.zeros(16, np.float32)
        a = _f(args).reshape(-1)
        A.lib().pb2h_mat4(op, _fp(a), _fp(out))
        return out.reshape(4, 4).T.copy()  # column-major storage -> math layout

    @staticmethod
    def translate(v): return mat4._op(0, v)
    @staticmethod
    def scale(v): return mat4._op(1, v)
    @staticmethod
    def rotate_x(r): return mat4._op(2, [r])
    @staticmethod
    def rotate_y(r): return mat4._op(3, [r])
    @staticmethod
    def rotate_z(r): return mat4._op(4, [r])
    @staticmethod
    def inverse(m): return mat4._op(5, _f(m).T.reshape(-1))
    @staticmethod
    def look_at(frm, to): return mat4._op(6, [*frm, *to])

    @staticmethod
    def mul(a, b):
        out = np.zeros(16, np.float32)
        ca, cb = _f(_f(a).T).reshape(-1), _f(_f(b).T).reshape(-1)
        A.lib().pb2h_mat4_mul(_fp(ca), _fp(cb), _fp(out))
        return out.reshape(4, 4).T.copy()


# ---- PRL-flavoured value types ---------------------------------------------------------------------
class _Mat:
    def __init__(self, kind, color, roughness=0.0, ior=1.4, metallic=0.0, transmission=0.0,
                 sigma_s=(0.0, 0.0, 0.0)):
        self.kind, self.roughness, self.ior = kind, roughness, ior
        self.color = color if isinstance(color, N.Node) else tuple(color)   # a Node3f graph stays a graph
        self.metallic, self.transmission, self.sigma_s = metallic, transmission, tuple(sigma_s)


def Emissive(color): return _Mat("emissive", color)
def Diffuse(albedo): return _Mat("diffuse", albedo)
def Metal(albedo, roughness): return _Mat("metal", albedo, roughness)
def Glossy(albedo, roughness, ior=1.4): return _Mat("glossy", albedo, roughness, ior)
def Glass(albedo, roughness, ior=1.4): return _Mat("glass", albedo, roughness, ior)
def Uber(albedo, roughness, metallic=0.0, transmission=0.0, ior=1.45):
    return _Mat("uber", albedo, roughness, ior, metallic, transmission)
def Subsurface(albedo, roughness, sigma_s):
    s = (sigma_s,) * 3 if np.isscalar(sigma_s) else tuple(sigma_s)
    return _Mat("subsurface", albedo, roughness, 1.4, sigma_s=s)


class _Shape:
    def __init__(self, kind, *args, **kw):
        self.kind, self.args, self.kw = kind, args, kw


def Rect(pos, ex, ey, flip=False): return _Shape("rect", pos, ex, ey, flip=flip)
def AABB(lo, hi): return _Shape("aabb", lo, hi)
def Box(a, b=None, m=None):
    """Box(lower, upper) -> AABB; Box(AABB, m) / Box(lower, upper, m) -> OBB (geometry.cpp:917-921)."""
    if isinstance(a, _Shape) and a.kind == "aabb":
        return _Shape("box", a.args[0], a.args[1], b)
    if m is None:
        return _Shape("aabb", a, b)
    return _Shape("box", a, b, m)
def Cone(p, n, r, h): return _Shape("cone", p, n, r, h)
def Sphere(c, r): return _Shape("sphere", c, r)
def Disk(p, n, r): return _Shape("disk", p, n, r)
def Triangle(a, b, c): return _Shape("triangle", a, b, c)
def Plane(p, n): return _Shape("plane", p, n)
def Cylinder(p0, p1, r): return _Shape("cylinder", p0, p1, r)
def Mesh(verts, idx, normals=None, uvs=None): return _Shape("mesh", verts, idx, normals=normals, uvs=uvs)


class Film:
    def __init__(self, size, tonemapper="uncharted2"):
        self.size = (int(size[0]), int(size[1]))
        self.tonemapper = tonemapper
        self.pixels = None  # float32 [h, w, 4], filled by PathIntegrator.render

    def save(self, path, scene=None):
        from .integrator import save_film
        save_film(self, path, scene)


def Uncharted2(): return "uncharted2"
def ACES(): return "aces"


class ThinLenCamera:
    def __init__(self, film, frm, to, fov, len_radius=0.0, focus_distance=1.0):
        self._film, self.frm, self.to, self.fov = film, frm, to, fov
        self.len_radius, self.focus_distance = len_radius, focus_distance

    def film(self):
        return self._film


class Scene:
    def __init__(self):
        self.materials = []          # A.Material
        self.named = {}
        self.geometries = []         # (shape_kind, shape_index, material)
        self.shapes = {k: [] for k in A.SHAPE_KINDS}
        self.mesh_data = []          # keeps numpy arrays alive
        self.textures = []           # uint8 [h, w, 4] arrays (pine::Image of vec4u8)
        self.node_instrs = []        # A.NodeInstr: the materials' node programs, each ending in `end`
        self.extra_lights = []       # (position in Scene::lights order = n geometries added before, A.Light)
        self.env_light = None        # A.Light of kind "sky"
        self.camera = None
        self._cam = None             # A.Camera
        self.size = None

    # ---- builder interface (shared with oracle.ref.RefScene) ----
    def add_texture(self, rgba):
        """An 8-bit image (uint8 [h, w, 3 or 4], row 0 first) for Uber materials' image nodes -> index."""
        a = np.asarray(rgba)
        if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] not in (3, 4) or a.size == 0:
            raise ValueError("texture must be a non-empty uint8 [h, w, 3|4] array")
        if a.shape[2] == 3:
            a = np.concatenate([a, np.full(a.shape[:2] + (1,), 255, np.uint8)], axis=2)
        self.textures.append(np.ascontiguousarray(a))
        return len(self.textures) - 1

    def add_material(self, kind, color, roughness=0.0, ior=1.4, metallic=0.0, transmission=0.0,
                     sigma_s=(0.0, 0.0, 0.0), color_texture=None, roughness_texture=None, metallic_texture=None):
        """`*_texture`: indices from add_texture (Uber only): NodeImage(NodeUV(), image) as albedo, its
        component 1 as roughness, its component 2 as metallic (fileio.cpp:285-292)."""
        m = A.Material()
        # a parameter given as a node graph (pine_b200.nodes) becomes a program; its constant stays 0
        params = dict(color=color, roughness=roughness, ior=ior, metallic=metallic, transmission=transmission)
        for slot in A.NODE_SLOTS:
            if N.is_graph(params[slot]):
                if kind == "emissive":
                    raise ValueError("Emissive color must be a constant on the CUDA path")
                setattr(m, slot + "_node", len(self.node_instrs) + 1)
                for ins in params[slot].program() + [(0, 0.0, 0.0, 0.0)]:
                    self.node_instrs.append(A.NodeInstr(ins[0] & 0xff, A.f3(*ins[1:])))
                params[slot] = (0.0, 0.0, 0.0) if slot == "color" else 0.0
            else:
                params[slot] = N.constant_of(params[slot], slot != "color")
        color, roughness, ior = params["color"], params["roughness"], params["ior"]
        metallic, transmission = params["metallic"], params["transmission"]
        m.color_texture = 0 if color_texture is None else int(color_texture) + 1
        m.roughness_texture = 0 if roughness_texture is None else int(roughness_texture) + 1
        m.metallic_texture = 0 if metallic_texture is None else int(metallic_texture) + 1
        m.kind = A.MATERIAL_KINDS[kind]
        m.color = A.f3(*[float(c) for c in color])
        m.roughness, m.ior, m.metallic, m.transmission = roughness, ior, metallic, transmission
        m.sigma_s = A.f3(*[float(c) for c in sigma_s])
        self.materials.append(m)
        return len(self.materials) - 1

    def _add(self, kind, rec, mat):
        self.shapes[kind].append(rec)
        self.geometries.append((A.SHAPE_KINDS[kind], len(self.shapes[kind]) - 1, mat))
        return len(self.geometries) - 1

    def add_rect(self, pos, ex, ey, mat, flip=False):
        r = A.Rect()
        a, b, c = _f(pos), _f(ex), _f(ey)
        A.lib().pb2h_make_rect(_fp(a), _fp(b), _fp(c), int(flip), C.byref(r))
        return self._add("rect", r, mat)

    def add_box(self, lo, hi, m, mat):
        o = A.Obb()
        a, b = _f(lo), _f(hi)
        cm = _f(_f(m).reshape(4, 4).T).reshape(-1)
        A.lib().pb2h_make_obb(_fp(a), _fp(b), _fp(cm), C.byref(o))
        return self._add("obb", o, mat)

    def add_aabb(self, lo, hi, mat):
        o = A.Aabb()
        o.lower, o.upper = A.f3(*map(float, lo)), A.f3(*map(float, hi))
        return self._add("aabb", o, mat)

    def add_cone(self, p, n, r, h, mat):
        o = A.Cone()
        a, b = _f(p), _f(n)
        A.lib().pb2h_make_cone(_fp(a), _fp(b), r, h, C.byref(o))
        return self._add("cone", o, mat)

    def add_sphere(self, c, r, mat):
        o = A.Sphere()
        o.c, o.r = A.f3(*map(float, c)), r
        return self._add("sphere", o, mat)

    def add_disk(self, p, n, r, mat):
        o = A.Disk()
        a, b = _f(p), _f(n)
        A.lib().pb2h_make_disk(_fp(a), _fp(b), r, C.byref(o))
        return self._add("disk", o, mat)

    def add_plane(self, p, n, mat):
        o = A.Plane()
        a, b = _f(p), _f(n)
        if not np.any(b):
            raise ValueError("`Plane` can't have degenerated normal")  # geometry.cpp:32
        A.lib().pb2h_make_plane(_fp(a), _fp(b), C.byref(o))
        return self._add("plane", o, mat)

    def add_cylinder(self, p0, p1, r, mat):
        o = A.Cylinder()
        a, b = _f(p0), _f(p1)
        A.lib().pb2h_make_cylinder(_fp(a), _fp(b), C.c_float(r), C.byref(o))
        return self._add("cylinder", o, mat)

    def add_triangle(self, a, b, c, mat):
        o = A.Triangle()
        a, b, c = _f(a), _f(b), _f(c)
        A.lib().pb2h_make_triangle(_fp(a), _fp(b), _fp(c), C.byref(o))
        return self._add("triangle", o, mat)

    def add_mesh(self, verts, idx, mat, normals=None, uvs=None):
        v = _f(verts).reshape(-1, 3)
        i = np.ascontiguousarray(idx, dtype=np.uint32).reshape(-1, 3)
        n = _f(normals).reshape(-1, 3) if normals is not None else None
        t = _f(uvs).reshape(-1, 2) if uvs is not None else None
        # pb2_upload_scene copies n_vertices normals / texcoords: shorter arrays would be read out of bounds
        if n is not None and len(n) != len(v):
            raise ValueError(f"mesh has {len(v)} vertices but {len(n)} normals")
        if t is not None and len(t) != len(v):
            raise ValueError(f"mesh has {len(v)} vertices but {len(t)} texture coordinates")
        if len(i) and int(i.max()) >= len(v):
            raise ValueError("mesh index out of range")
        self.mesh_data.append((v, i, n, t))
        m = A.Mesh()
        m.vertices, m.indices = v.ctypes.data, i.ctypes.data
        m.normals = n.ctypes.data if n is not None else None
        m.texcoords = t.ctypes.data if t is not None else None
        m.n_vertices, m.n_triangles = len(v), len(i)
        return self._add("mesh", m, mat)

    # ---- lights other than Emissive geometry (light.h:21-66; Scene::add_light / set_env_light) ----
    def _light(self, kind, position=(0, 0, 0), direction=(0, 0, 0), color=(0, 0, 0), falloff_cos=0.0,
               cutoff_cos=0.0):
        l = A.Light()
        l.kind, l.geometry = A.LIGHT_KINDS[kind], -1
        l.position = A.f3(*[float(np.float32(c)) for c in position])
        l.direction = A.f3(*[float(np.float32(c)) for c in direction])
        l.color = A.f3(*[float(np.float32(c)) for c in color])
        l.falloff_cos, l.cutoff_cos = falloff_cos, cutoff_cos
        return l

    @staticmethod
    def _unit(v):  # normalize(vec3) (vecmath.h): v / sqrt(x*x + y*y + z*z) in binary32, zero stays zero
        v = _f(v)
        n = np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2], dtype=np.float32)
        return v if n == 0 else (v / n).astype(np.float32)

    def add_point_light(self, position, color):
        self.extra_lights.append((len(self.geometries), self._light("point", position=position, color=color)))

    def add_directional_light(self, direction, color):
        self.extra_lights.append((len(self.geometries),
                                  self._light("directional", direction=self._unit(direction), color=color)))

    def add_spot_light(self, position, direction, color, falloff_radian, cutoff_additional_radian=0.0):
        f, c = np.float32(falloff_radian), np.float32(cutoff_additional_radian)
        if f <= 0 or f > np.float32(np.pi * 2):
            raise ValueError("`SpotLight` invalid falloff angle")  # light.cpp:26-29
        if c < 0 or f + c > np.float32(np.pi * 2):
            raise ValueError("`SpotLight` invalid cutoff angle")   # light.cpp:30-33
        import math
        fc = A.lib().pb2h_cosf(float(f))
        cc = A.lib().pb2h_cosf(float(np.float32(f + c)))
        self.extra_lights.append((len(self.geometries),
                                  self._light("spot", position, self._unit(direction), color, fc, cc)))

    def set_sky(self, sun_color):
        """scene.set(Sky(sun_color)): the environment light (light.cpp:71-84)"""
        self.env_light = self._light("sky", color=sun_color)

    def light_records(self):
        """UniformLightSampler::lights (lightsampler.cpp:6-10): Scene::lights in insertion order -- an AreaLight
        per Emissive geometry at the point it was added, the explicit lights where they were added -- then
        the environment light."""
        emissive = A.MATERIAL_KINDS["emissive"]
        out, extra = [], list(self.extra_lights)
        for g, (_, _, m) in enumerate(self.geometries):
            while extra and extra[0][0] <= g:
                out.append(extra.pop(0)[1])
            if self.materials[m].kind == emissive:
                l = A.Light()
                l.kind, l.geometry = A.LIGHT_KINDS["area"], g
                out.append(l)
        out += [l for _, l in extra]
        if self.env_light is not None:
            out.append(self.env_light)
        return out

    def set_camera(self, size, frm, to, fov, len_radius=0.0, focus=1.0, tonemapper="uncharted2"):
        self.size = (int(size[0]), int(size[1]))
        cam = A.Camera()
        a, b = _f(frm), _f(to)
        A.lib().pb2h_make_camera(self.size[0], self.size[1], 0 if tonemapper == "uncharted2" else 1,
                                 _fp(a), _fp(b), fov, len_radius, focus, C.byref(cam))
        self._cam = cam
        if self.camera is None:
            self.camera = ThinLenCamera(Film(self.size, tonemapper), frm, to, fov, len_radius, focus)

    # ---- PRL-flavoured front (scene.cpp:8-44) ----
    def add(self, a, b=None):
        if isinstance(a, str):
            self.named[a] = self.add_material(b.kind, b.color, b.roughness, b.ior, b.metallic,
                                              b.transmission, b.sigma_s)
            return self.named[a]
        if isinstance(b, str):
            if b not in self.named:
                raise KeyError(f"Can't find material `{b}`")  # scene.cpp:52-57
            mat = self.named[b]
        else:
            mat = self.add_material(b.kind, b.color, b.roughness, b.ior, b.metallic,
                                    b.transmission, b.sigma_s)
        s = a
        fn = {"rect": lambda: self.add_rect(*s.args, mat, **s.kw),
              "aabb": lambda: self.add_aabb(*s.args, mat),
              "box": lambda: self.add_box(*s.args, mat),
              "cone": lambda: self.add_cone(*s.args, mat),
              "sphere": lambda: self.add_sphere(*s.args, mat),
              "disk": lambda: self.add_disk(*s.args, mat),
              "triangle": lambda: self.add_triangle(*s.args, mat),
              "plane": lambda: self.add_plane(*s.args, mat),
              "cylinder": lambda: self.add_cylinder(*s.args, mat),
              "mesh": lambda: self.add_mesh(*s.args, mat, **s.kw)}[s.kind]
        return fn()

    def load(self, filename, m=None):
        """scene.load("file.glb"[, mat4]) (fileio.cpp:584-589)"""
        from .gltf import scene_from_gltf
        return scene_from_gltf(self, filename, m)

    def set(self, camera):
        self.camera = camera
        f = camera.film()
        self.set_camera(f.size, camera.frm, camera.to, camera.fov, camera.len_radius,
                        camera.focus_distance, f.tonemapper)

    def lights(self):
        """Scene::lights: one AreaLight per Emissive geometry, in insertion order (scene.cpp:19-20)."""
        return [g for g, (_, _, m) in enumerate(self.geometries)
                if self.materials[m].kind == A.MATERIAL_KINDS["emissive"]]

    def desc(self):
        """Flatten into the C-ABI pb2_scene_desc; returns (desc, keepalive)."""
        if self._cam is None:
            raise ValueError("scene has no camera")
        d = A.SceneDesc()
        keep = []

        def put(field, typ, items):
            arr = (typ * max(len(items), 1))(*items)
            keep.append(arr)
            setattr(d, field, C.cast(arr, C.POINTER(typ)))
            setattr(d, "n_" + field, len(items))

        geoms = []
        for k, i, m in self.geometries:
            g = A.Geometry()
            g.shape_kind, g.shape_index, g.material = k, i, m
            geoms.append(g)
        put("geometries", A.Geometry, geoms)
        put("materials", A.Material, self.materials)
        put("aabbs", A.Aabb, self.shapes["aabb"])
        put("obbs", A.Obb, self.shapes["obb"])
        put("spheres", A.Sphere, self.shapes["sphere"])
        put("triangles", A.Triangle, self.shapes["triangle"])
        put("cones", A.Cone, self.shapes["cone"])
        put("rects", A.Rect, self.shapes["rect"])
        put("disks", A.Disk, self.shapes["disk"])
        if self.shapes["plane"]:
            put("planes", A.Plane, self.shapes["plane"])
        if self.shapes["cylinder"]:
            put("cylinders", A.Cylinder, self.shapes["cylinder"])
        put("meshes", A.Mesh, self.shapes["mesh"])
        texs = []
        for a in self.textures:
            t = A.Texture()
            t.rgba, t.width, t.height = a.ctypes.data, a.shape[1], a.shape[0]
            texs.append(t)
        put("textures", A.Texture, texs)
        if self.node_instrs:
            put("node_instrs", A.NodeInstr, self.node_instrs)
        if self.extra_lights or self.env_light is not None:
            put("light_records", A.Light, self.light_records())
        ls = self.lights()
        larr = (C.c_int32 * max(len(ls), 1))(*ls)
        keep.append(larr)
        d.light_geometries = C.cast(larr, C.POINTER(C.c_int32))
        d.n_lights = len(ls)
        d.camera = self._cam
        return d, keep
