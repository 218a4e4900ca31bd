"""Material parameter graphs (pine's Nodef / Node3f, src/pine/core/node.h:105-250): flattening to the postfix
programs of include/pine_b200.h, the reference's OWN evaluation of the same graphs (oracle/_ref builds pine's node
objects from the typed program), the PRL names, and what pb2_upload_scene rejects. No GPU needed; the device
evaluation and the films are in tests/test_gpu_materials.py."""
import os

import numpy as np
import pytest

import pine_b200 as pb
from conftest import needs_ref
from pine_b200 import _abi as A
from pine_b200 import nodes as N
from pine_b200 import prl, scenes

REFERENCE = "/root/reference"


def _points(n, seed=7):
    rng = np.random.RandomState(seed)
    c = rng.uniform(-3, 3, (n, 8)).astype(np.float32)
    c[:, 6:] = rng.uniform(-1.5, 2.5, (n, 2)).astype(np.float32)
    return c


GRAPHS = {
    "classic_albedo": lambda: N.lerp(N.Checkerboard(N.UV(), 0.95), (0.5, 0.7, 1.0), (0.01, 0.02, 0.03)),
    "classic_rough": lambda: N.lerp(N.Checkerboard(N.UV(), 0.95), 0.0, 0.4),
    "checker_p": lambda: N.Checkerboard(N.Position() * 2.0, 0.5),
    "arith3": lambda: N.fract(N.Position() * (0.5, 2.0, 1.5)) / (N.abs_(N.Normal()) + (1.0, 2.0, 3.0)) - N.sqr(N.UV()),
    "arithf": lambda: N.sqrt(N.abs_(N.Position()[1])) * 0.25 + N.fract(N.UV()[0] * 5.0) - (-N.Normal()[2]),
    "vec3": lambda: N.Vec3(N.Position()[2], N.UV()[1], N.Checkerboard(N.UV() * 3.0)) * N.Vec3(N.UV()[0]),
    "lerp3": lambda: N.lerp(N.fract(N.Position()), N.Normal(), (1.0, 0.0, 0.5)),
}


def test_classic_pine_flattens_to_the_expected_program():
    prog = GRAPHS["classic_albedo"]().program()
    names = {v: k for k, v in N.OPS.items()}
    assert [names[i[0] & 0xff] for i in prog] == ["uv", "checkerboard", "const", "*", "const", "uv", "checkerboard",
                                                   "-", "const", "*", "+"]
    assert prog[1][1] == pytest.approx(0.95) and prog[1][0] & N.FLOAT_FLAG
    assert prog[2][1:] == tuple(float(np.float32(c)) for c in (0.01, 0.02, 0.03))   # t * b first, as ctx("lerp") builds it
    with pytest.raises(TypeError):
        N.Position() + 1.0                       # no `+` between a Node3f and a Nodef (node.cpp:72-92)
    with pytest.raises(ValueError):
        N.Comp(N.Position(), 3)                  # node.h:183-184
    with pytest.raises(TypeError):
        N.lerp(N.UV()[0], 0.5, (1.0, 1.0, 1.0))


@needs_ref
@pytest.mark.parametrize("name", list(GRAPHS))
def test_reference_evaluates_the_graphs_as_numpy_does(name):
    """The harness rebuilds pine's own node objects; their values equal the plain binary32 restatement here."""
    from oracle import ref
    c = _points(257)
    got = ref.eval_node(GRAPHS[name](), c)
    f32 = np.float32
    p, n, uv = c[:, 0:3], c[:, 3:6], np.concatenate([c[:, 6:8], np.zeros((len(c), 1), f32)], axis=1)
    fract = lambda x: (x - np.floor(x)).astype(f32)
    check = lambda q, r: ((fract(q) - f32(r)).prod(axis=1, dtype=f32) > 0).astype(f32)[:, None]
    one = f32(1.0)
    if name == "classic_albedo":
        t = check(uv, 0.95)
        want = t * np.array([0.01, 0.02, 0.03], f32) + (one - t) * np.array([0.5, 0.7, 1.0], f32)
    elif name == "classic_rough":
        t = check(uv, 0.95)
        want = np.repeat(t * f32(0.4) + (one - t) * f32(0.0), 3, axis=1)
    elif name == "checker_p":
        want = np.repeat(check(p * f32(2.0), 0.5), 3, axis=1)
    elif name == "arith3":
        want = fract(p * np.array([0.5, 2.0, 1.5], f32)) / (np.abs(n) + np.array([1, 2, 3], f32)) - uv * uv
    elif name == "arithf":
        want = np.repeat((np.sqrt(np.abs(p[:, 1])) * f32(0.25) + fract(uv[:, 0] * f32(5.0)) - (-n[:, 2]))[:, None], 3, axis=1)
    elif name == "vec3":
        want = np.stack([p[:, 2], uv[:, 1], check(uv * f32(3.0), 0.5)[:, 0]], axis=1) * uv[:, 0:1]
    else:
        t = fract(p)
        want = t * np.array([1.0, 0.0, 0.5], f32) + (one - t) * n
    assert np.array_equal(got.view(np.uint32), want.astype(f32).view(np.uint32)), name


def _upload_host_only(S):
    B = pb.Backend(-1)   # host-only context: validation + BVH build, no device
    try:
        B.upload_scene(S)
    finally:
        B.close()


def test_upload_validates_node_programs():
    def scene_with(instrs, kind="diffuse", slot="color_node", start=1):
        S = pb.Scene()
        m = S.add_material(kind, (0.5, 0.5, 0.5))
        setattr(S.materials[m], slot, start)
        S.node_instrs = [A.NodeInstr(op, A.f3(*a)) for op, *a in instrs]
        S.add_sphere((0, 0, 0), 1.0, m)
        S.set_camera((8, 8), (0, 0, -4), (0, 0, 0), 0.3)
        return S
    ok = [(N.OPS["uv"], 0, 0, 0), (N.OPS["checkerboard"], 0.5, 0, 0), (0, 0, 0, 0)]
    _upload_host_only(scene_with(ok))
    bad = {
        "no end": ok[:2],
        "underflow": [(N.OPS["+"], 0, 0, 0), (0, 0, 0, 0)],
        "two values left": [(N.OPS["uv"], 0, 0, 0), (N.OPS["uv"], 0, 0, 0), (0, 0, 0, 0)],
        "unknown op (a noise node)": [(99, 0, 0, 0), (0, 0, 0, 0)],
        "component 3": [(N.OPS["uv"], 0, 0, 0), (N.OPS["component"], 3, 0, 0), (0, 0, 0, 0)],
        "texture out of range": [(N.OPS["uv"], 0, 0, 0), (N.OPS["image"], 1, 0, 0), (0, 0, 0, 0)],
        "too deep": [(N.OPS["uv"], 0, 0, 0)] * 9 + [(N.OPS["+"], 0, 0, 0)] * 8 + [(0, 0, 0, 0)],
    }
    for why, prog in bad.items():
        with pytest.raises(pb.PineB200Error):
            _upload_host_only(scene_with(prog))
    with pytest.raises(pb.PineB200Error):
        _upload_host_only(scene_with(ok, start=7))                       # start beyond the array
    with pytest.raises(pb.PineB200Error):
        _upload_host_only(scene_with(ok, kind="emissive"))               # Emissive::color stays a constant
    with pytest.raises(ValueError):
        pb.Scene().add_material("emissive", N.Checkerboard(N.UV()) * (1.0, 1.0, 1.0))


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference checkout not present")
def test_classic_pine_loads_with_its_path_integrator_line():
    """scenes/classic.pine with the PathIntegrator line it keeps commented out instead of MicroRenderGI: the script
    builds the scene of scenes.classic(skin=True) -- same materials, node programs and geometry records."""
    text = open(os.path.join(REFERENCE, "scenes", "classic.pine")).read()
    text = text.replace("MicroRenderGI(1024 * 2).render(scene);", "").replace("#PathIntegrator", "PathIntegrator")
    renders = []
    it = prl.Interpreter(pb.Scene, on_render=lambda sc, spp, depth: renders.append((sc, spp, depth)))
    it.run(text)
    assert len(renders) == 1 and renders[0][1:] == (64, 6)
    S = renders[0][0].target
    W = pb.Scene()
    scenes.classic(W, skin=True)
    assert bytes(memoryview((A.NodeInstr * len(S.node_instrs))(*S.node_instrs))) == \
        bytes(memoryview((A.NodeInstr * len(W.node_instrs))(*W.node_instrs)))
    assert len(S.materials) == len(W.materials) and len(S.geometries) == len(W.geometries)
    used = lambda X: [bytes(memoryview(X.materials[g[2]])) for g in X.geometries]
    assert used(S) == used(W)
