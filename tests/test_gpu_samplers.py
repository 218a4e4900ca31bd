"""The other alternatives of pine::Sampler on the CUDA path, on a B200.

  * SobolSampler (sampler.h:83-164, sampler.cpp:82-113) is integer arithmetic up to the final scale: the
    device draws are BIT-EXACT against streams recorded from the compiled reference
    (tests/golden/sobol_kats.npz), and PathIntegrator(SobolSampler(4096), D) reproduces the reference's
    converged film (the `film_converged` fixture, rendered by the reference with that very sampler)
    within the equal-spp film tolerance;
  * UniformSampler (sampler.h:13-39) draws from a per-thread RNG in the reference, so only its
    distribution can be matched: the film must converge to the same image.
"""
import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32
from pine_b200 import _abi as A

pytestmark = pytest.mark.gpu
FILM_RELMSE_TOL = 1e-4


def _backend(name):
    S = pb.Scene()
    common.build(S, name)
    B = pb.Backend(0)
    B.upload_scene(S)
    return B, S


def test_sobol_device_draws_bit_exact(golden):
    k = golden("sobol_kats")
    for spp, w, h in k["cases"]:
        spp = int(spp)
        S = pb.Scene()
        m = S.add_material("diffuse", (0.5, 0.5, 0.5))
        S.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), m)
        S.set_camera((int(w), int(h)), (0, 1, -4), (0, 1, 0), 0.25)
        B = pb.Backend(0)
        B.upload_scene(S)  # Sampler::init(film size) fixes nbase4_digits
        for i, (px, py) in enumerate(k["pixels"]):
            d = B.test_sampler_kind(pb.SobolSampler(spp), int(px), int(py), 0, "21221", min(6, spp))
            assert np.array_equal(u32(d), u32(k[f"draws_{spp}_{i}"])), (spp, i)
        d = B.test_sampler_kind(pb.SobolSampler(spp), 5, 7, 0, "2" * 60 + "1" * 10, 1)
        assert np.array_equal(u32(d), u32(k[f"long_{spp}"])), spp
        B.close()


@pytest.mark.parametrize("name", ["c1_64", "c3_48", "mesh_48"])
def test_sobol_4096_reproduces_the_converged_reference_film(golden, name):
    g = golden(name)
    _, _, _, depth = common.SMALL_SCENES[name]
    B, _ = _backend(name)
    film = B.render(pb.SobolSampler(4096), depth)
    assert np.isfinite(film).all() and (film[..., 3] == 1).all()
    assert relmse(film, g["film_converged"]) <= FILM_RELMSE_TOL
    # sample sharding: two half-ranges summed converge to the same film (the per-pixel RNG that jitters the
    # camera sample is re-seeded per sample range, SURVEY 8e caveat 1, so not bit for bit)
    a = B.render(pb.SobolSampler(4096), depth, samples=(0, 2048))
    b = B.render(pb.SobolSampler(4096), depth, samples=(2048, 4096))
    both = a + b
    assert relmse(both, g["film_converged"]) <= 2e-3
    assert abs(float(both[..., :3].sum()) / float(film[..., :3].sum()) - 1) < 5e-3
    B.close()


def test_sobol_non_power_of_two_spp(golden):
    g = golden("c1_64")
    B, _ = _backend("c1_64")
    film = B.render(pb.SobolSampler(100), 4)
    e = relmse(film, g["film_converged"])
    e4 = relmse(g["film"], g["film_converged"])  # the reference at 4 spp (BlueSampler)
    assert e < e4, (e, e4)
    B.close()


def test_halton_device_draws_and_film(golden):
    k = golden("halton_kats")
    B, _ = _backend("c1_64")
    for spp in k["spps"]:
        spp = int(spp)
        for i, (px, py) in enumerate(k["pixels"]):
            d = B.test_sampler_kind(pb.HaltonSampler(spp), int(px), int(py), 0, "21221", min(6, spp))
            assert np.array_equal(u32(d), u32(k[f"draws_{spp}_{i}"])), (spp, i)
    d = B.test_sampler_kind(pb.HaltonSampler(4), 5, 7, 0, "2" * 520 + "1" * 3, 2)
    assert np.array_equal(u32(d), u32(k["wrap"]))
    film = B.render(pb.HaltonSampler(16), 4)
    assert relmse(film, k["film_c1_64_spp16"]) <= FILM_RELMSE_TOL
    B.close()


def test_uniform_sampler_converges_to_the_same_image(golden):
    g = golden("c1_64")
    B, _ = _backend("c1_64")
    f256 = B.render(pb.UniformSampler(256), 4)
    f4096 = B.render(pb.UniformSampler(4096, seed=3), 4)
    e256, e4096 = relmse(f256, g["film_converged"]), relmse(f4096, g["film_converged"])
    assert e4096 < 2e-3 and e4096 < e256 / 6, (e256, e4096)  # ~1/N: 16x more samples
    # a different seed is a different stream with the same mean
    other = B.render(pb.UniformSampler(256, seed=9), 4)
    assert not np.array_equal(u32(other), u32(f256))
    assert abs(float(other[..., :3].mean()) / float(f256[..., :3].mean()) - 1) < 0.02
    B.close()


def test_sampler_argument_errors():
    B, _ = _backend("c3_48")
    with pytest.raises(pb.PineB200Error) as e:
        B.render(pb.SobolSampler(1 << 20), 4)
    assert e.value.code == A.PB2_E_UNSUPPORTED

    class Unknown:
        kind, seed = 7, 0

        def spp(self):
            return 4
    with pytest.raises(pb.PineB200Error) as e:
        B.render(Unknown(), 4)
    assert e.value.code == A.PB2_E_UNSUPPORTED and "sampler kind" in str(e.value)
    with pytest.raises(ValueError):
        pb.UniformSampler(0)
    B.close()
