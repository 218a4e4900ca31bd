"""The drop-in boundary, end to end: ONE live `pine::Scene` (built through the reference's own C++
API inside oracle/_ref) is rendered by the reference's `PathIntegrator` and by
`pine::CudaPathIntegrator` (integration/cuda_path_integrator.cpp -- the Pine-side shim that flattens
the scene and calls the C ABI), both through `Integrator::render(Scene&)`, and the films written to
`scene.camera.film()` are compared. Also: the shim's flatten must feed the kernels the same numbers
the Python host layer does (films bit-identical), which pins `pb2_*` struct layouts against the
reference structs themselves."""
import numpy as np
import pytest

from common import SMALL_SCENES, build, relmse
from conftest import needs_ref

pytestmark = [pytest.mark.gpu, needs_ref]


@pytest.mark.parametrize("name", sorted(SMALL_SCENES))
def test_cuda_path_integrator_is_a_drop_in(name):
    import pine_b200 as pb
    from oracle import ref

    _, _, spp, depth = SMALL_SCENES[name]
    R = ref.RefScene()
    build(R, name)
    film_ref, _ = R.render(spp, depth)                  # pine::PathIntegrator, CPU
    film_cuda, _, stats = R.render_cuda(spp, depth)     # pine::CudaPathIntegrator on the same Scene
    assert stats["kernel_launches"] > 0 and stats["closest_rays"] > 0
    assert np.isfinite(film_cuda).all() and (film_cuda[..., 3] == 1).all()
    # same tolerance as test_film_matches_reference_at_equal_spp: device libm differs in last bits
    assert relmse(film_cuda, film_ref) < 2e-3

    # the shim's flatten == the Python host layer's description, bit for bit
    S = pb.Scene()
    build(S, name)
    B = pb.Backend(0)
    B.upload_scene(S)
    film_py = B.render(spp, depth)
    assert np.array_equal(film_py.view(np.uint32), film_cuda.view(np.uint32))


def test_shim_carries_lights_sky_and_the_other_samplers():
    """Scene::lights of every kind + the Sky environment light + SobolSampler through the C++ shim: the film
    the shim's flatten produces equals the Python host layer's bit for bit, and both match the reference."""
    import pine_b200 as pb
    from oracle import ref
    from pine_b200 import scenes
    R, S = ref.RefScene(), pb.Scene()
    scenes.lights_yard(R)
    cfg = scenes.lights_yard(S)
    film_ref, _ = R.render(cfg["spp"], cfg["depth"])
    film_cuda, _, stats = R.render_cuda(cfg["spp"], cfg["depth"])
    assert stats["shadow_rays"] > 0 and relmse(film_cuda, film_ref) <= 1e-4
    B = pb.Backend(0)
    B.upload_scene(S)
    assert np.array_equal(B.render(cfg["spp"], cfg["depth"]).view(np.uint32), film_cuda.view(np.uint32))
    sob_ref, _ = R.render(64, cfg["depth"], sampler="sobol")
    sob_cuda, _, _ = R.render_cuda(64, cfg["depth"], sampler="sobol")
    assert relmse(sob_cuda, sob_ref) <= 1e-4
    assert np.array_equal(B.render(pb.SobolSampler(64), cfg["depth"]).view(np.uint32), sob_cuda.view(np.uint32))
    B.close()


def test_reference_loader_plus_shim_equals_product_loader():
    """A .glb loaded by the REFERENCE's scene_from_gltf and rendered through the shim == the same file loaded by
    pine_b200/gltf.py and rendered through the Python host layer, bit for bit: the two loaders build the same
    scene (vertices, normals, Uber parameters, camera) and the two flatten paths the same records."""
    import os
    import common
    import pine_b200 as pb
    from oracle import ref
    path = os.path.join(common.GOLDEN, "room_like.glb")
    R = ref.RefScene()
    R.load_gltf(path, has_camera=True)
    film_cuda, _, _ = R.render_cuda(4, 5)
    S = pb.load_scene(path)
    B = pb.Backend(0)
    B.upload_scene(S)
    assert np.array_equal(B.render(4, 5).view(np.uint32), film_cuda.view(np.uint32))
    B.close()


def test_shim_keeps_the_reference_error_behaviour():
    """max_path_length <= 0 is SEVERE -> abort in the reference (path.cpp:12-13); the shim does the
    same, so it is checked in a child process."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); from oracle import ref; from common import build;"
            "R = ref.RefScene(); build(R, 'c3_48'); R.render_cuda(4, 0)")
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code % root], cwd=os.path.join(root, "tests"),
                       capture_output=True, text=True, timeout=120)
    assert r.returncode != 0
    assert "max_path_length" in (r.stdout + r.stderr)
