"""Generates tests/golden/*.npz from the compiled reference (oracle/_ref, i.e. the unmodified
wicstas/pine sources under /root/reference built by oracle/Makefile). Run in the build container:

    python tests/golden/make_golden.py

The reference ships no known-answer vectors for this path, so these files ARE the
pin: every vector is an output of the reference's own code on inputs described here.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402
from oracle import ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def kats():
    out = {}
    pixels = [(0, 0), (37, 411), (639, 12), (1000, 999)]
    for spp in (1, 2, 16, 64, 256):
        for i, (px, py) in enumerate(pixels):
            r, d = ref.sampler_stream(spp, px, py, i, 8, "21221", 6)
            out[f"rand_{spp}_{i}"] = r
            out[f"draws_{spp}_{i}"] = d
        # wrap-around of the dimension counter (sampler.h:182-193): 140 get2d draws in one sample
        _, d = ref.sampler_stream(spp, 5, 7, 0, 0, "2" * 140 + "1", 1)
        out[f"wrap_{spp}"] = d
    out["pixels"] = np.array(pixels, np.int32)
    hp = [(x, y, s) for x in (0, 1, 77, 3839) for y in (0, 5, 2159) for s in (0, 3)]
    out["hash_in"] = np.array(hp, np.int32)
    out["hash_out"] = np.array([ref.lib().ref_hash_pixel(*p) for p in hp], np.uint64)
    seeds = np.array([0, 1, 0xdeadbeef, 2 ** 63 + 5], np.uint64)
    rng = np.zeros((len(seeds), 8), np.uint64)
    for i, s in enumerate(seeds):
        import ctypes as C
        ref.lib().ref_rng_u64(C.c_uint64(int(s)), 8, rng[i].ctypes.data_as(C.c_void_p))
    out["rng_seeds"], out["rng_out"] = seeds, rng
    grid = np.zeros((5, 8, 8, 4, 8), np.float32)
    for a, spp in enumerate((1, 4, 16, 64, 256)):
        for px in range(8):
            for py in range(8):
                for idx in range(4):
                    for dim in range(8):
                        grid[a, px, py, idx, dim] = ref.bluenoise(spp, px * 17, py * 29, idx * 5, dim * 3)
    out["bluenoise_grid"] = grid
    out["nee_draw_order_2d_first"] = np.array([ref.lib().ref_nee_draw_order_2d_first()], np.int32)
    np.savez_compressed(os.path.join(OUT, "kats.npz"), **out)


SOBOL_CASES = [(1, (64, 64)), (3, (640, 640)), (16, (640, 640)), (100, (1024, 768)), (256, (720, 360)),
               (4096, (3840, 2160))]
SOBOL_PIXELS = [(0, 0), (37, 411), (639, 12), (3839, 2159)]


def sobol_kats():
    """SobolSampler(spp) after init(size), start_pixel(p, 0): 6 samples of the draw pattern 21221, plus a
    70-draw sample (dimension > 128), for power-of-two, odd-log2 and non-power-of-two spp."""
    out = {"cases": np.array([(s, w, h) for s, (w, h) in SOBOL_CASES], np.int32),
           "pixels": np.array(SOBOL_PIXELS, np.int32)}
    for spp, size in SOBOL_CASES:
        for i, (px, py) in enumerate(SOBOL_PIXELS):
            out[f"draws_{spp}_{i}"] = ref.sampler_stream_kind("sobol", spp, size, px, py, 0, "21221", min(6, spp))
        out[f"long_{spp}"] = ref.sampler_stream_kind("sobol", spp, size, 5, 7, 0, "2" * 60 + "1" * 10, 1)
    np.savez_compressed(os.path.join(OUT, "sobol_kats.npz"), **out)


def halton_kats():
    """HaltonSampler(spp) (sampler.h:40-81): 6 samples of the pattern 21221 at four pixels, a 1043-draw sample
    (the dimension counter wraps at PrimeTablesize = 1000), and its film of the small README scene."""
    out = {"pixels": np.array(SOBOL_PIXELS, np.int32), "spps": np.array([1, 4, 100, 4096], np.int32)}
    for spp in (1, 4, 100, 4096):
        for i, (px, py) in enumerate(SOBOL_PIXELS):
            out[f"draws_{spp}_{i}"] = ref.sampler_stream_kind("halton", spp, (640, 640), px, py, 0, "21221", min(6, spp))
    out["wrap"] = ref.sampler_stream_kind("halton", 4, (64, 64), 5, 7, 0, "2" * 520 + "1" * 3, 2)
    R = ref.RefScene()
    common.build(R, "c1_64")
    out["film_c1_64_spp16"], _ = R.render(16, 4, sampler="halton")
    np.savez_compressed(os.path.join(OUT, "halton_kats.npz"), **out)


def gltf_kats():
    """scene_from_gltf (fileio.cpp:146-327) on tests/golden/room_like.glb (written by tests/gltf_fixture.py):
    camera rays, closest hits of the camera rays and of a spherical ray fan from inside the room, the film
    at 4 spp and a SobolSampler(256) film, all from the reference's own loader + integrator."""
    path = os.path.join(OUT, "room_like.glb")
    R = ref.RefScene()
    out = {"n_geometries": np.array([R.load_gltf(path, has_camera=True)], np.int32),
           "size": np.array(R.size, np.int32)}
    rays = R.primary_rays(0)
    out["primary0"] = rays
    fan = common.ray_fan((0.05, 1.2, 0.1), 20000, 5)
    out["fan"] = fan
    for tag, r in (("p", rays), ("f", fan)):
        h = R.trace_closest(r, surf=True)
        for k in ("geom", "prim", "t", "surf"):
            out[f"{tag}_{k}"] = h[k]
    out["geometry_aabbs"] = R.geometry_aabbs()
    out["film"], _ = R.render(4, 5)
    out["film_sobol256"], _ = R.render(256, 5, sampler="sobol")
    np.savez_compressed(os.path.join(OUT, "room_like.npz"), **out)


TEXTURED_CAMERA = dict(size=(96, 64), frm=(0.3, 1.6, 3.6), to=(0.0, 0.8, 0.0), fov=0.45)


def gltf_textured_kats():
    """tests/golden/room_textured.glb (image-node materials, fileio.cpp:285-292) through the reference's loader
    (tinygltf + stb_image decode the embedded PNGs): films at 8 spp (BlueSampler) and 256 spp (SobolSampler)."""
    R = ref.RefScene()
    R.load_gltf(os.path.join(OUT, "room_textured.glb"))
    R.set_camera(**TEXTURED_CAMERA)
    out = {}
    out["film"], _ = R.render(8, 4)
    out["film_depth2"], _ = R.render(8, 2)
    out["film_sobol256"], _ = R.render(256, 4, sampler="sobol")
    np.savez_compressed(os.path.join(OUT, "room_textured.npz"), **out)


LIGHT_VARIANTS = {"all": {}, "no_sky": dict(sky=False), "sky_only": dict(delta=False, area=False),
                  "delta_only": dict(sky=False, area=False), "area_sky": dict(delta=False)}


def lights_kats():
    """scenes.lights_yard (Directional / Point / Spot lights, an Emissive rect between them, Sky) rendered by the
    reference: BlueSampler(16) films of every light subset, plus a SobolSampler(1024) film of the full scene."""
    from pine_b200 import scenes
    out = {}
    for name, kw in LIGHT_VARIANTS.items():
        R = ref.RefScene()
        cfg = scenes.lights_yard(R, **kw)
        out["film_" + name], _ = R.render(cfg["spp"], cfg["depth"])
        out["n_lights_" + name] = np.array([R.L.ref_num_lights(R.h)], np.int32)
    R = ref.RefScene()
    cfg = scenes.lights_yard(R)
    out["film_depth1"], _ = R.render(4, 1)
    out["film_sobol1024"], _ = R.render(1024, cfg["depth"], sampler="sobol")
    np.savez_compressed(os.path.join(OUT, "lights_yard.npz"), **out)


def minimal_kats():
    """What `scenes/minimal.pine` asks for (a unit sphere, one DirectionalLight, quick_render =
    640x480, fov 0.5, PathIntegrator(BlueSampler(4), 4), program_context.cpp:120-124) rendered by the reference."""
    R = ref.RefScene()
    R.add_sphere((0, 0, 0), 1.0, R.add_material("diffuse", (1, 1, 1)))
    R.add_directional_light((1, 1, -1), (1, 1, 1))
    R.set_camera((640, 480), (0, 0, -4), (0, 0, 0), 0.5)
    film, _ = R.render(4, 4)
    np.savez_compressed(os.path.join(OUT, "minimal.npz"), film=film.astype(np.float16), film_sum=np.array([film[..., :3].astype(np.float64).sum()]))


def scene(name):
    fn, kw, spp, depth = common.SMALL_SCENES[name]
    R = ref.RefScene()
    common.build(R, name)
    out = {}
    rays = R.primary_rays(0)
    out["primary0"] = rays
    out["primary2"] = R.primary_rays(2)
    h = R.trace_closest(rays, surf=True, count_tests=True)
    for k in ("geom", "prim", "t", "surf", "n_prim_tests"):
        out["p_" + k] = h[k]
    inp = common.secondary_inputs(h["surf"], h["geom"], 1)
    sec = ref.spawn_rays(inp)
    out["spawn_in"], out["secondary"] = inp, sec
    h2 = R.trace_closest(sec, surf=True)
    for k in ("geom", "prim", "t", "surf"):
        out["s_" + k] = h2[k]
    sh = common.shadow_from(sec, 2)
    out["shadow"], out["shadow_occ"] = sh, R.trace_any(sh)
    film, _ = R.render(spp, depth)
    out["film"] = film
    conv, _ = R.render(4096, depth, sampler="sobol")   # converged reference for relMSE
    out["film_converged"] = conv
    out["geometry_aabbs"] = R.geometry_aabbs()
    R.accel_build()
    t = R.bvh_tree(-1)
    for k, v in t.items():
        out["top_" + k] = np.asarray(v)
    idx, n_lbvh = R.bvh_indices()
    out["bvh_indices"], out["n_lbvh"] = idx, np.array([n_lbvh], np.int32)
    if n_lbvh:
        t = R.bvh_tree(0)
        for k, v in t.items():
            out["mesh0_" + k] = np.asarray(v)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k in ("primary0", "secondary", "film")})


def camera_lens():
    R = ref.RefScene()
    m = R.add_material("diffuse", (0.5, 0.5, 0.5))
    R.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), m)
    R.set_camera((32, 24), (0.3, 1.1, -4), (0.1, 0.9, 0.2), 0.3, len_radius=0.05, focus=3.5)
    rng = np.random.RandomState(3)
    inp = rng.rand(512, 4).astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "camera_lens.npz"), inp=inp, rays=R.gen_rays(inp))


if __name__ == "__main__":
    if not ref.available():
        raise SystemExit("oracle/_ref/libpine_ref.so missing: make -C oracle ref")
    kats()
    sobol_kats()
    halton_kats()
    gltf_kats()
    gltf_textured_kats()
    lights_kats()
    minimal_kats()
    camera_lens()
    for name in common.SMALL_SCENES:
        scene(name)
