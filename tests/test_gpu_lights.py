"""Lights beyond Emissive geometry on the CUDA path, on a B200: PointLight, SpotLight,
DirectionalLight (delta lights: no MIS, path.cpp:104-106) and the Sky environment light (seen by rays that
leave the scene, sampled uniformly over the sphere, light.cpp:71-84), in the reference's light order.
Films are compared with tests/golden/lights_yard.npz, rendered by the unmodified reference
(make_golden.py lights_kats); the scene has no RNG-driven material, so the equal-spp tolerance applies."""
import sys
import os

import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32
from pine_b200 import _abi as A
from pine_b200 import scenes

sys.path.insert(0, os.path.join(common.GOLDEN))
pytestmark = pytest.mark.gpu
VARIANTS = {"all": {}, "no_sky": dict(sky=False), "sky_only": dict(delta=False, area=False),
            "delta_only": dict(sky=False, area=False), "area_sky": dict(delta=False)}


def _render(kw, spp=None, depth=None, sampler=None):
    S = pb.Scene()
    cfg = scenes.lights_yard(S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    film = B.render(sampler or (spp or cfg["spp"]), depth or cfg["depth"])
    B.close()
    return film, S


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_light_subsets_match_the_reference(golden, name):
    g = golden("lights_yard")
    film, S = _render(VARIANTS[name])
    assert np.isfinite(film).all() and (film[..., 3] == 1).all()
    n_ref = int(g["n_lights_" + name][0])  # Scene::lights; the sampler appends the environment light
    assert len(S.light_records()) == n_ref + (1 if S.env_light is not None else 0)
    assert relmse(film, g["film_" + name]) <= 1e-4, relmse(film, g["film_" + name])


def test_camera_rays_see_the_sky_and_sobol_converges(golden):
    g = golden("lights_yard")
    film, _ = _render({}, spp=4, depth=1)   # depth 1: emitted / environment radiance only (path.cpp:76-89)
    assert relmse(film, g["film_depth1"]) <= 1e-6
    assert (film[-1, :, :3] > 0).all()      # the last film row looks over the wall into the sky (film y runs up)
    film, _ = _render({}, sampler=pb.SobolSampler(1024))
    assert relmse(film, g["film_sobol1024"]) <= 1e-4


def test_light_record_validation():
    S = pb.Scene()
    scenes.lights_yard(S)
    B = pb.Backend(0)
    recs = S.light_records()
    S.light_records = lambda: [recs[-1]] + recs[:-1]  # the environment light must come last
    with pytest.raises(pb.PineB200Error) as e:
        B.upload_scene(S)
    assert e.value.code == A.PB2_E_INVALID
    bad = A.Light()
    bad.kind = 9
    S.light_records = lambda: recs[:-1] + [bad]
    with pytest.raises(pb.PineB200Error) as e:
        B.upload_scene(S)
    assert e.value.code == A.PB2_E_UNSUPPORTED
    with pytest.raises(ValueError):
        S.add_spot_light((0, 1, 0), (0, -1, 0), (1, 1, 1), 7.0)  # light.cpp:28-29: radians, not degrees
    B.close()
