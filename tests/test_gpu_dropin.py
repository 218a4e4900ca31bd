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


def test_textured_glb_through_the_reference_loader_and_the_shim():
    """Image-node Uber materials (baseColorTexture / metallicRoughnessTexture) loaded by the REFERENCE's
    scene_from_gltf (fileio.cpp:285-292) reach the device through the shim's flatten as pb2_texture records:
    the film equals the product loader's bit for bit and the reference's own render to 1e-4."""
    import os
    import common
    import pine_b200 as pb
    from oracle import ref
    path = os.path.join(common.GOLDEN, "room_textured.glb")
    cam = ((96, 64), (0.3, 1.6, 3.6), (0.0, 0.8, 0.0), 0.45)
    R = ref.RefScene()
    R.load_gltf(path)
    R.set_camera(*cam)
    film_cuda, _, _ = R.render_cuda(8, 4)
    film_ref, _ = R.render(8, 4)
    assert relmse(film_cuda, film_ref) <= 1e-4
    S = pb.load_scene(path)
    S.set_camera(*cam)
    assert any(m.color_texture or m.roughness_texture or m.metallic_texture for m in S.materials)
    B = pb.Backend(0)
    B.upload_scene(S)
    assert np.array_equal(B.render(8, 4).view(np.uint32), film_cuda.view(np.uint32))
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


@pytest.mark.parametrize("tonemapper", ["uncharted2", "aces"])
def test_shim_device_epilogue_is_film_save(tonemapper):
    """CudaPathIntegrator::finalize_rgba8 -- Film::finalize + gamma / x256 / clamp / y-flip on the film still on
    the device, 4 bytes per pixel back (save_film_as_image, film.cpp:12-27, fileio.cpp:42-54) -- against the
    reference's own Film::finalize of the same film; the shim finds the Film's private tone mapper by itself."""
    from oracle import ref
    from pine_b200 import scenes
    R = ref.RefScene()
    cfg = scenes.cbox_readme(R, size=(96, 64))
    R.set_camera((96, 64), (0, 1, -4), (0, 1, 0), 0.25, tonemapper=tonemapper)
    film, rgba8 = R.render_cuda_rgba8(8, cfg["depth"])
    want = R.finalize(film)  # the reference's tone mapping of that film (host)
    q = np.clip(np.power(want[..., :3].astype(np.float32), np.float32(1 / 2.2)) * 256, 0, 255).astype(np.uint8)
    assert (np.abs(rgba8[::-1, :, :3].astype(int) - q.astype(int)) <= 1).all()
    assert (rgba8[..., 3] == 255).all()
