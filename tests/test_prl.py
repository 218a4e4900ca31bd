"""pine_b200/prl.py: the PRL-subset loader that lets Pine's scene FILES drive the backend.
CPU tests: grammar, binary32 arithmetic rules, error behaviour, and (where /root/reference is
mounted) that the reference's own scripts produce scene descriptions byte-identical to the builder
functions the benchmarks use. The GPU test renders a script and the equivalent builder scene."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import pine_b200 as pb
from pine_b200 import _abi as A
from pine_b200 import prl, scenes
from common import relmse

REFERENCE = "/root/reference"
needs_reference_tree = pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "scenes")),
                                          reason="reference checkout not mounted (GPU box)")


def desc_bytes(scene):
    """Every array of the pb2_scene_desc a Scene hands to pb2_upload_scene, as bytes."""
    d, keep = scene.desc()
    out = {}
    for name, typ in (("geometries", A.Geometry), ("materials", A.Material), ("rects", A.Rect),
                      ("obbs", A.Obb), ("cones", A.Cone), ("spheres", A.Sphere), ("disks", A.Disk)):
        n = getattr(d, "n_" + name)
        p = getattr(d, name)
        out[name] = bytes((C.c_char * (n * C.sizeof(typ))).from_address(C.addressof(p.contents))) if n else b""
    out["lights"] = [d.light_geometries[i] for i in range(d.n_lights)]
    out["camera"] = bytes(d.camera)
    return out


SCRIPT = """
# a small room, written the way Pine's scene files are
scene := Scene();
size := [16, 8] * 4;
scene.add("white", Diffuse([0.9, 0.9, 0.9]));
scene.add("mirror", Metal([1, 1, 1], 0.0));
scene.add(Rect([0, 0, 1], [2, 0, 0], [0, 0, 2]), "white");
scene.add(Rect([0, 1, 2], [2, 0, 0], [0, 2, 0], true), "mirror");
for x in -0.5 ~ 0.5 ~ 0.5 {
    scene.add(Cone([x, 0, 1], Y, 0.1, 0.3), "white");
}
scene.add(Sphere([0, 0.5, 1.2], 0.25), Glossy([1, 1, 1] / 4.0, 0.05));
scene.add(Box(AABB([0, 0, 0], [1, 1, 1]), translate([0.3, 0.0, 0.6]) * rotate_y(0.4) * scale([0.2, 0.4, 0.2])), "white");
scene.add(Rect([0.0, 1.9, 1], [0.3, 0, 0], [0, 0, 0.3]), Emissive(60 * [1.0, 0.64, 0.185]));
scene.set(ThinLenCamera(Film(size, ACES()), [0, 1, -4], [0, 1, 0], 0.25));
PathIntegrator(BlueSampler(4), 3).render(scene);
scene.camera.film().save("out/room.png");
"""


def build_by_hand(B):
    white = B.add_material("diffuse", (0.9, 0.9, 0.9))
    mirror = B.add_material("metal", (1, 1, 1), roughness=0.0)
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), white)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), mirror, flip=True)
    for x in (-0.5, 0.0, 0.5):
        B.add_cone((x, 0, 1), (0, 1, 0), 0.1, 0.3, white)
    g = B.add_material("glossy", (0.25, 0.25, 0.25), roughness=0.05)
    B.add_sphere((0, 0.5, 1.2), 0.25, g)
    m = pb.mat4.mul(pb.mat4.mul(pb.mat4.translate((0.3, 0.0, 0.6)), pb.mat4.rotate_y(0.4)), pb.mat4.scale((0.2, 0.4, 0.2)))
    B.add_box((0, 0, 0), (1, 1, 1), m, white)
    e = B.add_material("emissive", tuple(np.float32(60) * np.float32(c) for c in (1.0, 0.64, 0.185)))
    B.add_rect((0.0, 1.9, 1), (0.3, 0, 0), (0, 0, 0.3), e)
    B.set_camera((64, 32), (0, 1, -4), (0, 1, 0), 0.25, tonemapper="aces")


def test_script_equals_builder_calls():
    renders, saves = [], []
    sp, = prl.load(SCRIPT, pb.Scene, on_render=lambda s, spp, d: renders.append((spp, d)),
                   on_save=lambda s, p: saves.append(p))
    assert renders == [(4, 3)] and saves == ["out/room.png"]
    assert sp.config == {"spp": 4, "depth": 3, "save": "out/room.png"}
    S = pb.Scene()
    build_by_hand(S)
    assert desc_bytes(sp.target) == desc_bytes(S)


def test_arithmetic_follows_the_jit_types():
    ev = lambda src: prl.Interpreter(pb.Scene).eval(prl.Parser(src + ";").statement()[1])
    assert ev("[256, 256] * 4") == [1024, 1024] and all(isinstance(v, int) for v in ev("[256, 256] * 4"))
    assert ev("7 / 2") == 3 and ev("-7 / 2") == -3                      # integer division truncates
    v = ev("600 * [1.0, 0.64, 0.185]")
    assert [type(x) for x in v] == [np.float32] * 3
    assert v[1] == np.float32(600) * np.float32(0.64)                   # binary32, not double
    assert ev("1 + 2 * 3 - 4 / 2") == 5 and ev("(1 + 2) * 3") == 9
    # the inclusive float loop accumulates in binary32: -1 + 20 * 0.1f overshoots 1, so 20 values
    it = prl.Interpreter(pb.Scene)
    it.run("n := 0; for x in -1.0 ~ 0.1 ~ 1.0 { n := n + 1; }")
    assert it.env["n"] == 20
    it.run("k := 0; for i in 0 ~ 1 ~ 3 { k := k + i; }")
    assert it.env["k"] == 6


@pytest.mark.parametrize("src, msg", [
    ("scene := Scene(); scene.add(Rect([0,0,0],[1,0,0],[0,1,0]), \"nope\");", "unknown material"),
    ("x := foo(1);", "unknown name"),
    ("scene := Scene(); scene.add(Line([0,0,0],[0,1,0], 0.1), Diffuse([1,1,1]));", "unknown name"),
    ("x := [1, 2] + [1, 2, 3];", "size mismatch"),
    ("x := 1 $ 2;", "unexpected character"),
    ("x := (1 + 2;", "expected"),
    ("scene := Scene(); y := scene.camera;", "before scene.set"),
])
def test_errors_are_loud(src, msg):
    with pytest.raises(prl.PrlError, match=msg):
        prl.load(src, pb.Scene)


@needs_reference_tree
@pytest.mark.parametrize("source, builder, kw, want", [
    ("README.md", "cbox_readme", dict(size=(640, 640)), {"spp": 16, "depth": 4}),
    ("scenes/cbox.pine", "cbox_file", dict(size=(1024, 1024)), {"spp": 256, "depth": 8}),
    ("scenes/cbox-indirect.pine", "cbox_indirect", dict(size=(720, 360)), {"spp": 64, "depth": 5}),
])
def test_reference_scene_files_load_byte_identically(source, builder, kw, want):
    text = open(os.path.join(REFERENCE, source)).read()
    if source.endswith(".md"):
        text = re.search(r"```\n(scene := Scene\(\);.*?)```", text, re.S).group(1)
    sp, = prl.load(text, pb.Scene)
    assert {k: sp.config[k] for k in ("spp", "depth")} == want
    S = pb.Scene()
    scenes.SCENES[builder](S, **kw)
    assert desc_bytes(sp.target) == desc_bytes(S)


@pytest.mark.gpu
def test_script_renders_like_the_builder_scene(tmp_path):
    path = tmp_path / "room.pine"
    path.write_text(SCRIPT.replace("out/room.png", str(tmp_path / "room.png")))
    sp, = prl.run_file(str(path))
    S = pb.Scene()
    build_by_hand(S)
    film = pb.PathIntegrator(pb.BlueSampler(4), 3).render(S)
    assert np.array_equal(sp.target.camera.film().pixels.view(np.uint32), film.view(np.uint32))
    assert (tmp_path / "room.png").stat().st_size > 100


def test_lights_and_sky_from_a_script():
    # scene.add(PointLight | SpotLight | DirectionalLight) and scene.set(Sky) (light.cpp:174-185, scene.cpp:29-46)
    src = """
scene := Scene();
scene.add(Rect([0, 0, 0], [8, 0, 0], [0, 0, 8]), Diffuse([0.6, 0.6, 0.55]));
scene.add(DirectionalLight([1.0, 2.0, -1.5], [2.5, 2.4, 2.2]));
scene.add(Rect([1.6, 1.8, 0.6], [0.5, 0, 0], [0, 0, 0.5]), Emissive([30, 24, 16]));
scene.add(PointLight([-1.5, 1.2, -0.8], [3, 4, 6]));
scene.add(SpotLight([0.0, 2.6, -1.2], [0.1, -1.0, 0.45], [40, 40, 36], 0.35, 0.2));
scene.set(Sky([0.9, 1.0, 1.2]));
scene.set(ThinLenCamera(Film([32, 16], Uncharted2()), [0.4, 1.5, -4.5], [0.1, 0.7, 0], 0.3));
"""
    sp, = prl.load(src, pb.Scene)
    S = pb.Scene()
    S.add_rect((0, 0, 0), (8, 0, 0), (0, 0, 8), S.add_material("diffuse", (0.6, 0.6, 0.55)))
    S.add_directional_light((1.0, 2.0, -1.5), (2.5, 2.4, 2.2))
    S.add_rect((1.6, 1.8, 0.6), (0.5, 0, 0), (0, 0, 0.5), S.add_material("emissive", (30, 24, 16)))
    S.add_point_light((-1.5, 1.2, -0.8), (3, 4, 6))
    S.add_spot_light((0.0, 2.6, -1.2), (0.1, -1.0, 0.45), (40, 40, 36), 0.35, 0.2)
    S.set_sky((0.9, 1.0, 1.2))
    a, b = sp.target.light_records(), S.light_records()
    assert [l.kind for l in a] == [A.LIGHT_KINDS[k] for k in ("directional", "area", "point", "spot", "sky")]
    assert [bytes(l) for l in a] == [bytes(l) for l in b]
    d = np.array(list(a[0].direction), np.float32)
    assert abs(float(np.linalg.norm(d)) - 1) < 1e-6  # DirectionalLight normalises (light.h:42)
    assert np.float32(a[3].falloff_cos) == np.float32(np.cos(np.float32(0.35)))


MINIMAL = """
scene = Scene();
scene.add(Sphere([0, 0, 0], 1), Diffuse([1, 1, 1]));
scene.add(DirectionalLight([1, 1, -1], [1, 1, 1]));
scene.quick_render([0, 0, -4], [0, 0, 0], "images/minimal.png");
"""


def test_minimal_scene_with_quick_render():
    # the shape of scenes/minimal.pine: `=` assignment, a delta light, quick_render (program_context.cpp:120-124)
    saves = []
    sp, = prl.load(MINIMAL, pb.Scene, on_save=lambda s, p: saves.append(p))
    assert sp.config == {"spp": 4, "depth": 4, "save": "images/minimal.png"} and saves == ["images/minimal.png"]
    assert sp.target.size == (640, 480) and len(sp.target.geometries) == 1
    S = pb.Scene()
    S.add_sphere((0, 0, 0), 1.0, S.add_material("diffuse", (1, 1, 1)))
    S.add_directional_light((1, 1, -1), (1, 1, 1))
    S.set_camera((640, 480), (0, 0, -4), (0, 0, 0), 0.5)
    assert desc_bytes(sp.target) == desc_bytes(S)
    assert [bytes(l) for l in sp.target.light_records()] == [bytes(l) for l in S.light_records()]
    ref_file = "/root/reference/scenes/minimal.pine"
    if os.path.isfile(ref_file):  # the reference's own file, where the checkout is present
        sp2, = prl.load(open(ref_file).read(), pb.Scene)
        assert desc_bytes(sp2.target) == desc_bytes(S) and sp2.config == sp.config


@pytest.mark.gpu
def test_minimal_scene_renders_like_the_reference(golden, tmp_path):
    g = golden("minimal")
    path = tmp_path / "minimal.pine"
    path.write_text(MINIMAL.replace("images/minimal.png", str(tmp_path / "minimal.png")))
    sp, = prl.run_file(str(path))
    film = sp.target.camera.film().pixels
    assert film.shape == (480, 640, 4)
    assert abs(float(film[..., :3].astype(np.float64).sum()) / float(g["film_sum"][0]) - 1) < 1e-5
    assert relmse(film, g["film"].astype(np.float32)) <= 1e-4  # the fixture is stored as float16
    assert (tmp_path / "minimal.png").stat().st_size > 1000


@needs_reference_tree
def test_pine_pine_loads_through_the_subset():
    """scenes/pine.pine: single-quoted strings, a `for f in 0.0~0.2~Pi` loop that builds a scene per iteration,
    string + number concatenation for the file name, println -- 16 PathIntegrator(BlueSampler(4), 3) renders of
    a Cone under a Rect emitter, the second scene.set(camera) winning."""
    calls = []
    text = open(os.path.join(REFERENCE, "scenes", "pine.pine")).read()
    sps = prl.load(text, pb.Scene, on_render=lambda sp, spp, depth: calls.append(("render", spp, depth)),
                   on_save=lambda *a: calls.append(("save", a[-1])))
    assert len(sps) == 16
    assert [c for c in calls if c[0] == "render"] == [("render", 4, 3)] * 16
    saves = [c[1] for c in calls if c[0] == "save"]
    assert saves[0] == "images/pine0.0.png" and len(set(saves)) == 16
    S = sps[-1].target
    d, keep = S.desc()
    assert d.n_cones == 1 and d.n_rects == 1 and d.n_geometries == 2 and d.n_lights == 1
    assert tuple(d.camera.position) == (0.0, 0.0, -8.0) and (d.camera.width, d.camera.height) == (640, 360)


def test_uber_material_and_string_rules():
    sp, = prl.load("""
        scene = Scene();
        scene.add('m', Uber([0.98, 0.55, 0.02], 0.25, 1.0));
        scene.add(Sphere([0, 1, 0], 1), 'm');
        name = 'a' + 2 + "b";
    """, pb.Scene)
    m = sp.target.materials[0]
    assert m.kind == A.MATERIAL_KINDS["uber"] and abs(m.metallic - 1.0) < 1e-7 and abs(m.ior - 1.45) < 1e-6
    with pytest.raises(prl.PrlError):
        prl.load("x = 'a' * 2;", pb.Scene)


def test_plane_and_cylinder_in_a_script():
    """Plane(position, normal) and Cylinder(p0, p1, r) (program_context / geometry_context names) build the records
    of the builder scene."""
    text = """
scene = Scene();
scene.add(Plane([0, 0, 0], [0, 1, 0]), Diffuse([0.8, 0.8, 0.8]));
scene.add(Cylinder([-2, 0, 2], [-2, 2.5, 2], 0.5), Diffuse([0.9, 0.2, 0.1]));
scene.set(ThinLenCamera(Film([32, 24]), [0, 2.5, -6], [0, 1, 2], 0.35));
"""
    it = prl.Interpreter(pb.Scene)
    it.run(text)
    S = it.scenes[0].target
    W = pb.Scene()
    W.add_plane((0, 0, 0), (0, 1, 0), W.add_material("diffuse", (0.8, 0.8, 0.8)))
    W.add_cylinder((-2, 0, 2), (-2, 2.5, 2), 0.5, W.add_material("diffuse", (0.9, 0.2, 0.1)))
    for k in ("plane", "cylinder"):
        assert [bytes(memoryview(x)) for x in S.shapes[k]] == [bytes(memoryview(x)) for x in W.shapes[k]] and S.shapes[k]
    B = pb.Backend(-1)
    B.upload_scene(S)   # host-only context: validation, BVH over the +-100 / radius-r boxes
    assert B.counters()["n_primitives"] == 2
    B.close()
