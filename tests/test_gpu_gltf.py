"""A .glb scene through the product loader (pine_b200/gltf.py) onto the CUDA path, on a B200: the hits and
the film must be the ones the unmodified reference gets from ITS loader on the same file
(tests/golden/room_like.npz, made by make_golden.py gltf_kats from tests/golden/room_like.glb)."""
import os

import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32

pytestmark = pytest.mark.gpu
GLB = os.path.join(common.GOLDEN, "room_like.glb")


@pytest.fixture(scope="module")
def room():
    S = pb.load_scene(GLB)
    B = pb.Backend(0)
    B.upload_scene(S)
    yield S, B
    B.close()


@pytest.mark.parametrize("accel", [0, 1], ids=["wide", "ref_order"])
def test_glb_hits_bit_exact(golden, room, accel):
    g, (S, B) = golden("room_like"), room
    assert np.array_equal(u32(B.primary_rays(0)), u32(g["primary0"]))
    for tag, rays in (("p", g["primary0"]), ("f", g["fan"])):
        h = B.trace_closest(rays, accel=accel, surf=True)
        assert np.array_equal(h["geom"], g[f"{tag}_geom"]) and np.array_equal(h["prim"], g[f"{tag}_prim"])
        assert np.array_equal(u32(h["t"]), u32(g[f"{tag}_t"]))
        hit = h["geom"] >= 0
        # hit point and (interpolated, transformed) shading normal: Mesh::apply + compute_surface_info
        assert np.array_equal(u32(h["surf"][hit][:, :6]), u32(g[f"{tag}_surf"][hit][:, :6]))


def test_glb_film_matches_the_reference(golden, room):
    g, (S, B) = golden("room_like"), room
    film = B.render(4, 5)
    assert film.shape == (640, 96, 4) and np.isfinite(film).all()
    # Uber materials pick their lobe with the per-pixel RNG (sampler.h:317-324) and the glass cube refracts:
    # last-bit libm differences send a few samples down another path (as in test_gpu_materials), so: nearly
    # every pixel agrees closely, the rest is as far from a converged film as the reference itself is
    a, b = film[..., :3].astype(np.float64), g["film"][..., :3].astype(np.float64)
    close = (np.abs(a - b) <= 1e-3 * np.abs(b) + 1e-5).all(-1)
    assert close.mean() >= 0.97, close.mean()
    assert abs(a.sum() / b.sum() - 1) < 0.01
    hi = g["film_sobol256"]
    e_gpu, e_ref = relmse(film, hi), relmse(g["film"], hi)
    assert e_gpu <= 1.1 * e_ref + 1e-4, (e_gpu, e_ref)
    sob = B.render(pb.SobolSampler(256), 5)
    a, b = sob[..., :3].astype(np.float64), hi[..., :3].astype(np.float64)
    assert abs(a.sum() / b.sum() - 1) < 0.005 and relmse(sob, hi) <= 0.05 * e_ref + 1e-4, (relmse(sob, hi), e_ref)
    assert np.array_equal(u32(film), u32(B.render(4, 5, accel=1)))


def test_glb_through_path_integrator_and_prl(tmp_path, golden):
    g = golden("room_like")
    S = pb.Scene()
    S.load(GLB)
    film = pb.PathIntegrator(pb.BlueSampler(4), 5).render(S)
    assert relmse(film, g["film"]) <= 1e-3
    from pine_b200 import prl
    script = tmp_path / "room.pine"
    script.write_text(f'scene := Scene();\nscene.load("{GLB}");\nPathIntegrator(BlueSampler(4), 5).render(scene);\n')
    sp, = prl.run_file(str(script))
    assert np.array_equal(u32(sp.target.camera.film().pixels), u32(film))


def test_textured_glb_film_matches_the_reference(golden):
    """Image-node Uber materials: texel addressing (fract / wrap / nearest), gamma decode and the
    roughnessFactor-overrides-texture quirk all show in the film; the fixture keeps every lobe choice
    deterministic (metallic texels 0 or 255), so the equal-spp tolerance of the plain scenes applies."""
    from pine_b200 import _abi as A
    g = golden("room_textured")
    S = pb.load_scene(os.path.join(common.GOLDEN, "room_textured.glb"))
    S.set_camera((96, 64), (0.3, 1.6, 3.6), (0.0, 0.8, 0.0), 0.45)
    B = pb.Backend(0)
    B.upload_scene(S)
    for depth, key in ((2, "film_depth2"), (4, "film")):
        film = B.render(8, depth)
        assert np.isfinite(film).all()
        assert relmse(film, g[key]) <= 1e-4, (depth, relmse(film, g[key]))
    assert relmse(B.render(pb.SobolSampler(256), 4), g["film_sobol256"]) <= 1e-4
    # texture indices are validated, and only Uber materials may carry image nodes
    S.materials[0].color_texture = 9
    with pytest.raises(pb.PineB200Error) as e:
        B.upload_scene(S)
    assert e.value.code == A.PB2_E_INVALID
    S.materials[0].color_texture = 1
    S.materials[0].kind = A.MATERIAL_KINDS["diffuse"]
    with pytest.raises(pb.PineB200Error) as e:
        B.upload_scene(S)
    assert e.value.code == A.PB2_E_UNSUPPORTED
    B.close()
