"""A loader for the subset of PRL (Pine Rendering Language) that Pine's scene scripts use, so that the
scene FILES themselves (README.md:7-35, scenes/cbox.pine, scenes/cbox-indirect.pine) drive the CUDA
backend -- and, through the same builder interface, the reference harness.

The real front end is an LLVM-JIT compiler (src/pine/core/jit.cpp, needs libLLVM-18, absent here);
this is not it. It interprets exactly the grammar those files use:
  name := expr;   expr;   for x in a ~ step ~ b { ... }   # comments
  numbers, strings, true/false, [a, b, c] vector literals, + - * /, calls, method calls,
and binds the names the scripts call to the builder interface of `pine_b200.Scene` /
`oracle.ref.RefScene` (program_context.cpp / *_context functions register the same names):
  Scene, Rect, Box, AABB, Cone, Sphere, Disk, Triangle, Diffuse, Metal, Glossy, Glass, Subsurface,
  Emissive, PointLight, SpotLight, DirectionalLight, Sky, translate, rotate_x/y/z, scale, ThinLenCamera, Film, Uncharted2, ACES, BlueSampler, SobolSampler, UniformSampler,
  PathIntegrator, X/Y/Z.
Arithmetic follows the JIT's types: int stays int, anything with a float is binary32, the inclusive
float loop `a ~ s ~ b` accumulates in binary32 (jit.cpp:1534-1541), matrices multiply through the
C++ host mirror of vecmath.h. Anything outside the subset raises PrlError -- nothing is guessed.
"""
import re

import numpy as np

from . import nodes as N
from .scene import mat4


class PrlError(Exception):
    pass


_TOKEN = re.compile(r"""
    (?P<ws>\s+|\#[^\n]*)
  | (?P<num>\d+\.\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+|\d+)
  | (?P<str>"[^"\n]*"|'[^'\n]*')
  | (?P<id>[A-Za-z_][A-Za-z_0-9]*)
  | (?P<op>:=|[-+*/~.,;(){}\[\]=])
""", re.X)


def tokenize(text):
    out, pos = [], 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            line = text.count("\n", 0, pos) + 1
            raise PrlError(f"line {line}: unexpected character {text[pos]!r}")
        pos = m.end()
        kind = m.lastgroup
        if kind != "ws":
            out.append((kind, m.group(kind), text.count("\n", 0, m.start()) + 1))
    out.append(("eof", "", text.count("\n") + 1))
    return out


class Parser:
    """text -> AST of tuples: ("assign", name, e) ("expr", e) ("for", var, a, step, b, body) and
    expressions ("num", v) ("str", s) ("bool", b) ("id", n) ("vec", [e]) ("neg", e)
    ("bin", op, l, r) ("call", callee_expr, [args]) ("attr", obj_expr, name)."""

    def __init__(self, text):
        self.t, self.i = tokenize(text), 0

    def peek(self, k=0):
        return self.t[self.i + k]

    def take(self, value=None, kind=None):
        tok = self.t[self.i]
        if (value is not None and tok[1] != value) or (kind is not None and tok[0] != kind):
            raise PrlError(f"line {tok[2]}: expected {value or kind}, got {tok[1]!r}")
        self.i += 1
        return tok

    def program(self):
        body = []
        while self.peek()[0] != "eof":
            body.append(self.statement())
        return body

    def statement(self):
        tok = self.peek()
        if tok == ("id", "for", tok[2]):
            self.take()
            var = self.take(kind="id")[1]
            self.take("in")
            a = self.expr(); self.take("~"); step = self.expr(); self.take("~"); b = self.expr()
            self.take("{")
            body = []
            while self.peek()[1] != "}":
                body.append(self.statement())
            self.take("}")
            return ("for", var, a, step, b, body)
        if tok[0] == "id" and self.peek(1)[1] in (":=", "="):  # `:=` declares, `=` assigns (minimal.pine)
            name = self.take()[1]
            self.take()
            e = self.expr()
            self.take(";")
            return ("assign", name, e)
        e = self.expr()
        self.take(";")
        return ("expr", e)

    def expr(self):
        left = self.term()
        while self.peek()[1] in ("+", "-"):
            op = self.take()[1]
            left = ("bin", op, left, self.term())
        return left

    def term(self):
        left = self.unary()
        while self.peek()[1] in ("*", "/"):
            op = self.take()[1]
            left = ("bin", op, left, self.unary())
        return left

    def unary(self):
        if self.peek()[1] == "-":
            self.take()
            return ("neg", self.unary())
        return self.postfix()

    def args(self, close):
        out = []
        while self.peek()[1] != close:
            out.append(self.expr())
            if self.peek()[1] == ",":
                self.take()
        self.take(close)
        return out

    def postfix(self):
        e = self.primary()
        while True:
            v = self.peek()[1]
            if v == ".":
                self.take()
                e = ("attr", e, self.take(kind="id")[1])
            elif v == "(":
                self.take()
                e = ("call", e, self.args(")"))
            else:
                return e

    def primary(self):
        kind, v, line = self.take()
        if kind == "num":
            return ("num", int(v) if re.fullmatch(r"\d+", v) else np.float32(v))
        if kind == "str":
            return ("str", v[1:-1])
        if kind == "id":
            if v in ("true", "false"):
                return ("bool", v == "true")
            return ("id", v)
        if v == "[":
            return ("vec", self.args("]"))
        if v == "(":
            e = self.expr()
            self.take(")")
            return e
        raise PrlError(f"line {line}: unexpected {v!r}")


# ---- values -------------------------------------------------------------------------------------------------
class _Shape:
    def __init__(self, kind, *a):
        self.kind, self.a = kind, a


class _Material:
    def __init__(self, kind, color, **kw):
        self.kind, self.color, self.kw = kind, color, kw


class _Film:
    def __init__(self, size, tonemapper="uncharted2"):
        self.size, self.tonemapper = size, tonemapper


class _Camera:
    def __init__(self, film, frm, to, fov, len_radius=0.0, focus=1.0):
        self.film_, self.frm, self.to, self.fov, self.len_radius, self.focus = film, frm, to, fov, len_radius, focus

    def film(self):
        return self.film_


class _Light:
    def __init__(self, kind, *a):
        self.kind, self.a = kind, a


class _Sampler:
    name = "BlueSampler"

    def __init__(self, spp, seed=0):
        self.spp, self.seed = int(spp), int(seed)


class _SobolSampler(_Sampler):
    name = "SobolSampler"


class _UniformSampler(_Sampler):
    name = "UniformSampler"


class _HaltonSampler(_Sampler):
    name = "HaltonSampler"


def _f32(v):
    return np.float32(v)


def _vec(v, n=3):
    """A PRL vector literal / scalar as n binary32 components (scalars broadcast, as the C++
    converting constructors do: Subsurface(..., 40.0) takes a vec3)."""
    if isinstance(v, (list, tuple, np.ndarray)):
        if len(v) != n:
            raise PrlError(f"expected a {n}-vector, got {len(v)} components")
        return tuple(float(_f32(x)) for x in v)
    return tuple(float(_f32(v)) for _ in range(n))


def _node3(v):
    """A Node3f parameter: a graph stays a graph, anything else becomes a vec3 constant."""
    return v if isinstance(v, N.Node) else _vec(v)


def _nodef(v):
    return v if isinstance(v, N.Node) else float(_f32(v))


def _arith(op, a, b):
    """int op int stays int (jit.cpp integer ops); otherwise binary32, element-wise over vectors,
    4x4 * 4x4 through the C++ host mirror of vecmath.h."""
    if isinstance(a, N.Node) or isinstance(b, N.Node):  # ctx("+") ... ctx("/") on Nodef / Node3f (node.cpp:72-92)
        try:
            x, y = N.as_node(a), N.as_node(b)
            return {"+": x + y, "-": x - y, "*": x * y, "/": x / y}[op]
        except TypeError as e:
            raise PrlError(str(e))
    if isinstance(a, str) or isinstance(b, str):  # 'images/pine' + f + '.png' (jit.cpp string concatenation)
        if op != "+":
            raise PrlError("only + is defined on strings")
        fmt = lambda v: v if isinstance(v, str) else (str(v) if isinstance(v, int) else repr(float(v)))
        return fmt(a) + fmt(b)
    is_mat = lambda x: isinstance(x, np.ndarray) and x.shape == (4, 4)
    if is_mat(a) or is_mat(b):
        if op == "*" and is_mat(a) and is_mat(b):
            return mat4.mul(a, b)
        raise PrlError("only mat4 * mat4 is supported on matrices")
    is_vec = lambda x: isinstance(x, list)
    if is_vec(a) or is_vec(b):
        n = len(a) if is_vec(a) else len(b)
        av = a if is_vec(a) else [a] * n
        bv = b if is_vec(b) else [b] * n
        if len(av) != len(bv):
            raise PrlError("vector size mismatch")
        return [_arith(op, x, y) for x, y in zip(av, bv)]
    if isinstance(a, int) and isinstance(b, int) and not isinstance(a, bool):
        if op == "/":
            if b == 0:
                raise PrlError("integer division by zero")
            q = abs(a) // abs(b)
            return q if (a >= 0) == (b >= 0) else -q  # C semantics: truncation toward zero
        return {"+": a + b, "-": a - b, "*": a * b}[op]
    x, y = _f32(a), _f32(b)
    with np.errstate(all="ignore"):
        return {"+": x + y, "-": x - y, "*": x * y, "/": x / y}[op]


class SceneProxy:
    """`scene` of a script: forwards to a builder with the interface of pine_b200.Scene."""

    def __init__(self, target, on_render=None, on_save=None):
        self.target, self.on_render, self.on_save = target, on_render, on_save
        self.named, self.camera, self.config = {}, None, {}
        self.search_dirs = []

    def _material(self, m):
        if isinstance(m, str):
            if m not in self.named:
                raise PrlError(f"unknown material {m!r}")  # Scene::find_material aborts (scene.cpp)
            return self.named[m]
        if not isinstance(m, _Material):
            raise PrlError("expected a material or a material name")
        return self.target.add_material(m.kind, m.color, **m.kw)

    def add(self, a, b=None):
        if isinstance(a, _Light) and b is None:  # scene.add(Light) (scene.cpp:29-31)
            if a.kind == "point":
                self.target.add_point_light(_vec(a.a[0]), _vec(a.a[1]))
            elif a.kind == "directional":
                self.target.add_directional_light(_vec(a.a[0]), _vec(a.a[1]))
            elif a.kind == "spot":
                self.target.add_spot_light(_vec(a.a[0]), _vec(a.a[1]), _vec(a.a[2]), float(_f32(a.a[3])),
                                           float(_f32(a.a[4])) if len(a.a) > 4 else 0.0)
            else:
                raise PrlError("scene.add: an environment light is set with scene.set(...)")
            return
        if isinstance(a, str):  # scene.add("name", Material)
            self.named[a] = self._material(b)
            return
        if not isinstance(a, _Shape):
            raise PrlError("scene.add expects (name, material) or (shape, material)")
        mat = self._material(b)
        k, p = a.kind, a.a
        if k == "rect":
            self.target.add_rect(_vec(p[0]), _vec(p[1]), _vec(p[2]), mat, flip=bool(p[3]) if len(p) > 3 else False)
        elif k == "box":
            self.target.add_box(_vec(p[0].a[0]), _vec(p[0].a[1]), p[1], mat)
        elif k == "aabb":
            self.target.add_aabb(_vec(p[0]), _vec(p[1]), mat)
        elif k == "cone":
            self.target.add_cone(_vec(p[0]), _vec(p[1]), float(_f32(p[2])), float(_f32(p[3])), mat)
        elif k == "sphere":
            self.target.add_sphere(_vec(p[0]), float(_f32(p[1])), mat)
        elif k == "disk":
            self.target.add_disk(_vec(p[0]), _vec(p[1]), float(_f32(p[2])), mat)
        elif k == "triangle":
            self.target.add_triangle(_vec(p[0]), _vec(p[1]), _vec(p[2]), mat)
        elif k == "plane":
            self.target.add_plane(_vec(p[0]), _vec(p[1]), mat)
        elif k == "cylinder":
            self.target.add_cylinder(_vec(p[0]), _vec(p[1]), float(_f32(p[2])), mat)
        else:
            raise PrlError(f"shape {k} is not supported by the CUDA path")

    def load(self, filename, m=None):
        """scene.load("file.glb"[, mat4]) (fileio.cpp:584-589): scene_from_gltf into this scene. The name is
        tried as given (the reference resolves it against the working directory), then against the
        script's directory and its parent."""
        import os
        if not isinstance(filename, str):
            raise PrlError("scene.load expects a file name")
        tried = [filename] + [os.path.join(d, filename) for d in self.search_dirs]
        path = next((t for t in tried if os.path.isfile(t)), None)
        if path is None:
            raise PrlError(f"scene.load: no such file {filename!r}")
        if hasattr(self.target, "load_gltf"):  # oracle.ref.RefScene: the reference's own loader
            if m is not None:
                raise PrlError("scene.load with a transform is not wired to the reference harness")
            self.target.load_gltf(path)
        else:
            from . import gltf
            gltf.scene_from_gltf(self.target, path, m)

    def quick_render(self, frm, to, filename):
        """quick_render(scene, from, to, file) (program_context.cpp:120-124): 640x480 film, fov 0.5,
        PathIntegrator(BlueSampler(4), 4), save."""
        self.set(_Camera(_Film([640, 480]), frm, to, 0.5))
        _Integrator(_Sampler(4), 4).render(self)
        self.config["save"] = filename
        if self.on_save:
            self.on_save(self, filename)

    def set(self, cam):
        if isinstance(cam, _Light) and cam.kind == "sky":  # scene.set(Sky(sun_color)) (scene.cpp:44-46)
            self.target.set_sky(_vec(cam.a[0]))
            return
        if not isinstance(cam, _Camera):
            raise PrlError("scene.set expects a camera")
        self.camera = cam
        size = tuple(int(x) for x in cam.film_.size)
        self.target.set_camera(size, _vec(cam.frm), _vec(cam.to), float(_f32(cam.fov)),
                               float(_f32(cam.len_radius)), float(_f32(cam.focus)), cam.film_.tonemapper)


class _Integrator:
    def __init__(self, sampler, depth):
        if not isinstance(sampler, _Sampler):
            raise PrlError("PathIntegrator expects a sampler")
        self.sampler, self.depth = sampler, int(depth)

    def render(self, scene):
        scene.config = {"spp": self.sampler.spp, "depth": self.depth}
        if self.sampler.name != "BlueSampler":
            scene.config["sampler"] = self.sampler.name
            scene.config["sampler_seed"] = self.sampler.seed
        if scene.on_render:
            scene.on_render(scene, self.sampler.spp, self.depth)


class _FilmHandle:
    def __init__(self, scene):
        self.scene = scene

    def save(self, path):
        self.scene.config["save"] = path
        if self.scene.on_save:
            self.scene.on_save(self.scene, path)


def _subsurface(albedo, roughness, sigma_s):
    return _Material("subsurface", _node3(albedo), roughness=_nodef(roughness), sigma_s=_vec(sigma_s))


def _lerp(t, a, b):
    """ctx("lerp"): on nodes the graph form (node.cpp:94-110), on plain values psl::lerp(t, a, b) = (1 - t) a + t b."""
    if any(isinstance(v, N.Node) for v in (t, a, b)):
        wrap = lambda v: v if isinstance(v, N.Node) or not isinstance(v, list) else tuple(v)
        try:
            return N.lerp(wrap(t), wrap(a), wrap(b))
        except TypeError as e:
            raise PrlError(str(e))
    return _arith("+", _arith("*", _arith("-", 1.0, t), a), _arith("*", t, b))


def _node_or(node_fn, plain_fn):
    return lambda v: node_fn(v) if isinstance(v, N.Node) else plain_fn(v)


def _globals(make_scene):
    shape = lambda kind: (lambda *a: _Shape(kind, *a))
    return {
        "Scene": make_scene,
        "Rect": shape("rect"), "Box": shape("box"), "AABB": shape("aabb"), "Cone": shape("cone"),
        "Sphere": shape("sphere"), "Disk": shape("disk"), "Triangle": shape("triangle"),
        "Plane": shape("plane"), "Cylinder": shape("cylinder"),
        "Diffuse": lambda c: _Material("diffuse", _node3(c)),
        "Emissive": lambda c: _Material("emissive", _vec(c)),
        "Metal": lambda c, r: _Material("metal", _node3(c), roughness=_nodef(r)),
        "Glossy": lambda c, r, ior=1.4: _Material("glossy", _node3(c), roughness=_nodef(r), ior=_nodef(ior)),
        "Glass": lambda c, r, ior=1.4: _Material("glass", _node3(c), roughness=_nodef(r), ior=_nodef(ior)),
        "Subsurface": _subsurface,
        # UberMaterial(albedo, roughness, metallic = 0, transmission = 0, ior = 1.45) (material.h:80-87, .cpp:53-57)
        "Uber": lambda c, r, metallic=0.0, transmission=0.0, ior=1.45: _Material(
            "uber", _node3(c), roughness=_nodef(r), metallic=_nodef(metallic),
            transmission=_nodef(transmission), ior=float(_f32(ior))),
        # material parameter graphs (node_context, node.cpp:29-132); noise and Function nodes have no device form
        "Position": N.Position, "Normal": N.Normal, "UV": N.UV, "Checkerboard": N.Checkerboard,
        "Comp": N.Comp, "Vec3": N.Vec3, "lerp": _lerp,
        "abs": _node_or(N.abs_, lambda v: abs(v)), "sqr": _node_or(N.sqr, lambda v: _arith("*", v, v)),
        "sqrt": _node_or(N.sqrt, lambda v: np.sqrt(_f32(v))),
        "fract": _node_or(N.fract, lambda v: _f32(v) - np.floor(_f32(v))),
        "Pi": np.float32(3.14159265358979323846),  # ctx.var("Pi") (math.cpp:7)
        "print": lambda *a: None, "println": lambda *a: None,  # LOG output only (program_context.cpp:41)
        "translate": lambda v: mat4.translate(_vec(v)), "scale": lambda v: mat4.scale(_vec(v)),
        "rotate_x": lambda r: mat4.rotate_x(float(_f32(r))), "rotate_y": lambda r: mat4.rotate_y(float(_f32(r))),
        "rotate_z": lambda r: mat4.rotate_z(float(_f32(r))),
        "Film": lambda size, tm="uncharted2": _Film(size, tm),
        "Uncharted2": lambda: "uncharted2", "ACES": lambda: "aces",
        "PointLight": lambda p, c: _Light("point", p, c),
        "DirectionalLight": lambda d, c: _Light("directional", d, c),
        "SpotLight": lambda *a: _Light("spot", *a), "Sky": lambda c: _Light("sky", c),
        "ThinLenCamera": _Camera, "BlueSampler": _Sampler, "SobolSampler": _SobolSampler,
        "UniformSampler": _UniformSampler, "HaltonSampler": _HaltonSampler, "PathIntegrator": _Integrator,
        "X": [1, 0, 0], "Y": [0, 1, 0], "Z": [0, 0, 1],
    }


class Interpreter:
    def __init__(self, target_factory, on_render=None, on_save=None):
        self.scenes = []
        self.search_dirs = []

        def make_scene():
            s = SceneProxy(target_factory(), on_render, on_save)
            s.search_dirs = self.search_dirs
            self.scenes.append(s)
            return s
        self.env = _globals(make_scene)

    def run(self, text):
        for st in Parser(text).program():
            self.exec(st)
        return self.scenes

    def exec(self, st):
        if st[0] == "assign":
            self.env[st[1]] = self.eval(st[2])
        elif st[0] == "expr":
            self.eval(st[1])
        else:  # for x in a ~ step ~ b: x = a; x <= b; x += step, in the operands' type
            _, var, a, step, b, body = st
            x, step, b = self.eval(a), self.eval(step), self.eval(b)
            if not all(isinstance(v, int) for v in (x, step, b)):
                x, step, b = _f32(x), _f32(step), _f32(b)
            if not step > 0:
                raise PrlError("for: step must be positive")
            while x <= b:
                self.env[var] = x
                for s in body:
                    self.exec(s)
                x = x + step if isinstance(x, int) else _f32(x + step)

    def eval(self, e):
        k = e[0]
        if k in ("num", "str", "bool"):
            return e[1]
        if k == "id":
            if e[1] not in self.env:
                raise PrlError(f"unknown name {e[1]!r} (outside the supported PRL subset)")
            return self.env[e[1]]
        if k == "vec":
            return [self.eval(x) for x in e[1]]
        if k == "neg":
            return _arith("-", 0, self.eval(e[1]))
        if k == "bin":
            return _arith(e[1], self.eval(e[2]), self.eval(e[3]))
        if k == "attr":
            obj = self.eval(e[1])
            if isinstance(obj, SceneProxy) and e[2] == "camera":
                if obj.camera is None:
                    raise PrlError("scene.camera read before scene.set")
                return _CameraHandle(obj)
            if not hasattr(obj, e[2]) or e[2].startswith("_"):
                raise PrlError(f"no member {e[2]!r}")
            return getattr(obj, e[2])
        if k == "call":
            fn = self.eval(e[1])
            if not callable(fn):
                raise PrlError("call of a non-function")
            return fn(*[self.eval(a) for a in e[2]])
        raise PrlError(f"bad node {k}")


class _CameraHandle:
    def __init__(self, scene):
        self.scene = scene

    def film(self):
        return _FilmHandle(self.scene)


def load(text, target_factory, on_render=None, on_save=None, search_dirs=()):
    """Run a script against builders made by `target_factory()`; -> list of SceneProxy (each with
    .target, .config = {"spp", "depth", "save"} as the script asked). `search_dirs`: where scene.load
    looks for a file it cannot find relative to the working directory."""
    it = Interpreter(target_factory, on_render, on_save)
    it.search_dirs = list(search_dirs)
    return it.run(text)


def run_file(path, device=0):
    """`pine <file>` for the supported subset: builds the scene, renders it on the GPU when the script
    calls PathIntegrator(...).render(scene), saves the PNG when it calls film().save()."""
    from . import integrator as I
    from .scene import Scene
    films = {}

    def on_render(sp, spp, depth):
        kind = sp.config.get("sampler", "BlueSampler")
        sampler = (I.SobolSampler(spp) if kind == "SobolSampler" else
                   I.HaltonSampler(spp) if kind == "HaltonSampler" else
                   I.UniformSampler(spp, sp.config.get("sampler_seed", 0)) if kind == "UniformSampler" else
                   I.BlueSampler(spp))
        integ = I.PathIntegrator(sampler, depth, device=device)
        films[id(sp)] = integ.render(sp.target)

    def on_save(sp, out):
        I.save_film(sp.target.camera.film(), out, scene=sp.target)
    import os
    here = os.path.dirname(os.path.abspath(path))
    with open(path) as f:
        return load(f.read(), Scene, on_render, on_save, search_dirs=(here, os.path.dirname(here)))
