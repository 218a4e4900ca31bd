"""Parity of the CUDA path with the reference, through the C ABI (include/pine_b200.h), on a B200.

Bars (BASELINE.json north_star):
  * integer / index / bit-pattern work -- RNG, sampler, camera rays, closest-hit (geom, prim, t),
    any-hit, spawn_ray, surface point and normal -- is BIT-EXACT against the golden vectors, which are
    outputs of the unmodified reference (tests/golden/make_golden.py);
  * the film is compared in float: relMSE(gpu, reference at equal spp) <= 1e-4 (device libm
    sin/cos/pow/log differ from glibc in the last bits), the error against the converged
    4096-spp reference film matches the reference's own error within 2 %, and every pixel mean lies
    within 3 sigma of the reference's.
"""
import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32
from conftest import needs_ref
from pine_b200 import scenes

pytestmark = pytest.mark.gpu
SCENES = list(common.SMALL_SCENES)
FILM_RELMSE_TOL = 1e-4


@pytest.fixture(scope="module")
def gpu():
    return pb.Backend(0)  # raises without a device: no CPU fallback


@pytest.fixture(scope="module")
def backends():
    cache = {}

    def get(name):
        if name not in cache:
            S = pb.Scene()
            common.build(S, name)
            B = pb.Backend(0)
            B.upload_scene(S)
            cache[name] = B
        return cache[name]
    yield get
    for b in cache.values():
        b.close()


def test_native_library_is_the_one_running(gpu):
    from pine_b200 import _abi
    assert _abi.lib().pb2_device_count() >= 1
    with open("/proc/self/maps") as f:
        assert "libpine_b200.so" in f.read()


def test_rng_and_sampler_known_answers(golden, gpu):
    k = golden("kats")
    for spp in (1, 2, 16, 64, 256):
        for i, (px, py) in enumerate(k["pixels"]):
            r, d = gpu.test_sampler(spp, int(px), int(py), i, 8, "21221", 6)
            assert np.array_equal(u32(r), u32(k[f"rand_{spp}_{i}"]))
            assert np.array_equal(u32(d), u32(k[f"draws_{spp}_{i}"]))
        _, d = gpu.test_sampler(spp, 5, 7, 0, 0, "2" * 140 + "1", 1)
        assert np.array_equal(u32(d), u32(k[f"wrap_{spp}"]))


@pytest.mark.parametrize("name", SCENES)
def test_primary_rays_bit_exact(golden, backends, name):
    g, B = golden(name), backends(name)
    assert np.array_equal(u32(B.primary_rays(0)), u32(g["primary0"]))
    assert np.array_equal(u32(B.primary_rays(2)), u32(g["primary2"]))


@pytest.mark.parametrize("accel", [0, 1], ids=["wide", "ref_order"])
@pytest.mark.parametrize("name", SCENES)
def test_closest_hit_bit_exact(golden, backends, name, accel):
    g, B = golden(name), backends(name)
    for rays, pre in ((g["primary0"], "p_"), (g["secondary"], "s_")):
        h = B.trace_closest(rays, accel=accel, surf=True)
        assert np.array_equal(h["geom"], g[pre + "geom"])
        assert np.array_equal(h["prim"], g[pre + "prim"])
        assert np.array_equal(u32(h["t"]), u32(g[pre + "t"]))
        # surface point and normal: same float ops => same bits. uv of a Sphere goes through
        # atan2f/acosf (geometry.cpp:94-96), libm => tolerance there only.
        assert np.array_equal(u32(h["surf"][:, :6]), u32(g[pre + "surf"][:, :6]))
        np.testing.assert_allclose(h["surf"][:, 6:], g[pre + "surf"][:, 6:], rtol=0, atol=2e-6)


@pytest.mark.parametrize("accel", [0, 1], ids=["wide", "ref_order"])
@pytest.mark.parametrize("name", SCENES)
def test_any_hit_exact(golden, backends, name, accel):
    g, B = golden(name), backends(name)
    assert np.array_equal(B.trace_any(g["shadow"], accel=accel), g["shadow_occ"])


@pytest.mark.parametrize("name", SCENES)
def test_film_matches_reference_at_equal_spp(golden, backends, name):
    g, B = golden(name), backends(name)
    _, _, spp, depth = common.SMALL_SCENES[name]
    film = B.render(spp, depth)
    assert np.isfinite(film).all() and (film[..., 3] == 1).all()
    assert relmse(film, g["film"]) <= FILM_RELMSE_TOL
    # the production (wide) and the verification (reference-order) traversal give the same film bits
    assert np.array_equal(u32(film), u32(B.render(spp, depth, accel=1)))
    # error against the converged reference equals the reference's own error at this spp
    e_gpu, e_ref = relmse(film, g["film_converged"]), relmse(g["film"], g["film_converged"])
    assert abs(e_gpu - e_ref) <= 0.02 * e_ref + 1e-6


@pytest.mark.parametrize("name", SCENES)
def test_every_ray_of_a_render_matches_the_oracle(backends, name):
    """All closest-hit and shadow rays a production render actually traces (pb2_set_ray_record):
    the results the render used equal, bit for bit, the oracle's reference-order walk (CPU
    restatement, oracle/oracle_port.cpp) and the device's own reference-order walk."""
    from oracle import port
    B = backends(name)
    S = pb.Scene()
    common.build(S, name)
    _, _, spp, depth = common.SMALL_SCENES[name]
    film, rec = B.render_recorded(spp, depth, accel=0)
    assert np.array_equal(u32(film), u32(B.render(spp, depth)))
    assert len(rec["closest_t"]) >= film.shape[0] * film.shape[1] * spp and len(rec["shadow_occluded"]) > 0
    P = port.PortScene(S)
    want = P.trace_closest(rec["closest_rays"])
    assert np.array_equal(rec["closest_geom"], want["geom"])
    assert np.array_equal(rec["closest_prim"], want["prim"])
    assert np.array_equal(u32(rec["closest_t"]), u32(want["t"]))
    assert np.array_equal(rec["shadow_occluded"], P.trace_any(rec["shadow_rays"])["occluded"])
    dev = B.trace_closest(rec["closest_rays"], accel=1)
    assert np.array_equal(rec["closest_geom"], dev["geom"]) and np.array_equal(u32(rec["closest_t"]), u32(dev["t"]))
    assert np.array_equal(rec["shadow_occluded"], B.trace_any(rec["shadow_rays"], accel=1))


@pytest.mark.parametrize("name", ["c1_64", "mesh_48"])
def test_pixel_means_within_three_sigma(golden, backends, name):
    """|mean_gpu - mean_ref| <= 3 sigma per pixel and channel, sigma = standard error of the mean from
    the GPU's own per-sample radiance (one single-sample render per sample index)."""
    g, B = golden(name), backends(name)
    _, _, spp, depth = common.SMALL_SCENES[name]
    film = B.render(spp, depth)[..., :3].astype(np.float64)
    per = np.stack([B.render(spp, depth, samples=(s, s + 1))[..., :3].astype(np.float64) * spp
                    for s in range(spp)])
    sigma = per.std(axis=0, ddof=1) / np.sqrt(spp)
    diff = np.abs(film - g["film"][..., :3].astype(np.float64))
    assert (diff <= 3 * sigma + 1e-4 * (1 + np.abs(film))).all()


def test_edge_cases_of_the_ray_hooks(backends):
    B = backends("c1_64")
    h = B.trace_closest(np.zeros((0, 8), np.float32))
    assert len(h["geom"]) == 0 and len(B.trace_any(np.zeros((0, 8), np.float32))) == 0
    fmax = np.float32(3.4028235e38)
    rays = np.array([
        [0, 1, -4, 0, 0, 1, 0, fmax],     # axis-aligned direction: two zero components (safe_rcp)
        [0, 1, -4, 0, 0, -1, 0, fmax],    # points away from the box: miss
        [0, 1, -4, 0, 0, 1, 0, 0.0],      # tmax = 0: nothing can be accepted
        [0, 1, -4, 0, 0, 1, 100.0, fmax], # tmin beyond the scene
        [0, 5, 1, 0, -1, 0, 0, fmax],     # from above through the ceiling Rect
    ], np.float32)
    for accel in (0, 1):
        h = B.trace_closest(rays, accel=accel)
        assert h["geom"][0] >= 0 and h["geom"][1] == -1 and h["geom"][2] == -1 and h["geom"][3] == -1
        assert h["t"][1] == fmax and h["t"][2] == 0.0
    assert np.array_equal(B.trace_closest(rays, accel=0)["t"].view(np.uint32),
                          B.trace_closest(rays, accel=1)["t"].view(np.uint32))
    assert np.array_equal(B.trace_any(rays, accel=0), B.trace_any(rays, accel=1))


def test_render_edge_cases(gpu):
    # empty scene: every path misses, film = 0 with alpha 1 (path.cpp:75-81 without an env light)
    S = pb.Scene()
    S.set_camera((16, 8), (0, 0, -3), (0, 0, 0), 0.3)
    B = pb.Backend(0)
    B.upload_scene(S)
    film = B.render(4, 4)
    assert (film[..., :3] == 0).all() and (film[..., 3] == 1).all()
    # depth 1: only directly visible emission survives (path.cpp:83-89)
    S = pb.Scene()
    scenes.cbox_indirect(S, size=(32, 32))
    B.upload_scene(S)
    f1 = B.render(1, 1)
    assert set(np.unique(f1[..., 0])) <= {np.float32(0.0), f1[..., 0].max()}
    # empty row range and empty sample range render nothing
    assert (B.render(4, 3, rows=(5, 5))[..., :3] == 0).all()
    assert (B.render(4, 3, samples=(2, 2))[..., :3] == 0).all()
    # the largest table set
    f256 = B.render(256, 2)
    assert np.isfinite(f256).all()
    with pytest.raises(pb.PineB200Error):
        B.render(3, 4)   # not a BlueSobolSampler::spp() value
    with pytest.raises(pb.PineB200Error):
        B.render(4, 0)   # path.cpp:12-13
    B.close()


def test_full_size_properties_c1(gpu):
    """BASELINE config 1 at full size (640x640, 16 spp, depth 4): size-independent properties."""
    S = pb.Scene()
    cfg = scenes.cbox_readme(S, size=(640, 640))
    B = pb.Backend(0)
    B.upload_scene(S)
    spp, depth = cfg["spp"], cfg["depth"]
    full = B.render(spp, depth)
    c = B.counters()
    assert c["last_closest_rays"] <= 640 * 640 * spp * depth
    assert c["last_shadow_rays"] <= 640 * 640 * spp * (depth - 1)
    # determinism: one writer per pixel, fixed sample order (path.cpp:36-38)
    assert np.array_equal(u32(full), u32(B.render(spp, depth)))
    # tile sharding (rows) is bit-identical to the full render
    parts = np.zeros_like(full)
    for r0 in range(0, 640, 160):
        parts[r0:r0 + 160] = B.render(spp, depth, rows=(r0, r0 + 160))[r0:r0 + 160]
    assert np.array_equal(u32(parts), u32(full))
    # wide traversal + fallback == reference-order traversal, ray for ray => same film bits
    assert np.array_equal(u32(full), u32(B.render(spp, depth, accel=1)))
    # sample sharding: the shards sum to a film statistically equal to the full one (per-rank RNG
    # reseed): same total energy within 1 %, alpha sums to 1
    acc = sum(B.render(spp, depth, samples=(s, s + spp // 4)).astype(np.float64) for s in range(0, spp, spp // 4))
    assert np.allclose(acc[..., 3], 1.0)
    assert abs(acc[..., :3].sum() / full[..., :3].astype(np.float64).sum() - 1) < 0.01
    # closest hits of all 409600 camera rays: production walk == verification walk, bit for bit
    rays = B.primary_rays(0)
    a, b = B.trace_closest(rays, accel=0), B.trace_closest(rays, accel=1)
    assert np.array_equal(a["geom"], b["geom"]) and np.array_equal(a["prim"], b["prim"])
    assert np.array_equal(u32(a["t"]), u32(b["t"]))
    B.close()


def test_large_mesh_wide_equals_reference_order(gpu):
    """~1.3 M triangles (C5's generator, fewer instances): 2 M incoherent rays through the wide tree
    and through the reference-order tree agree bit for bit (closest) and exactly (any)."""
    S = pb.Scene()
    scenes.mesh_field(S, n_instances=256, subdiv=4, seed=42, size=(512, 512), with_shapes=True)
    B = pb.Backend(0)
    B.upload_scene(S)
    rays = B.primary_rays(0)
    h = B.trace_closest(rays, accel=0, surf=True)
    sec = common.secondary_inputs(h["surf"], h["geom"], 5)
    from oracle import port  # spawn_ray (interaction.cpp:6-13) restated on the CPU
    sec_rays = port.spawn_rays(sec)
    allrays = np.concatenate([rays, sec_rays])
    a, b = B.trace_closest(allrays, accel=0), B.trace_closest(allrays, accel=1)
    assert np.array_equal(a["geom"], b["geom"]) and np.array_equal(a["prim"], b["prim"])
    assert np.array_equal(u32(a["t"]), u32(b["t"]))
    sh = common.shadow_from(sec_rays, 9)
    assert np.array_equal(B.trace_any(sh, accel=0), B.trace_any(sh, accel=1))
    B.close()


@needs_ref
def test_live_reference_c1_160(gpu):
    """The compiled reference run on this box (oracle/_ref): README scene at 160x160, 16 spp."""
    from oracle import ref
    R, S = ref.RefScene(), pb.Scene()
    scenes.cbox_readme(R, size=(160, 160))
    cfg = scenes.cbox_readme(S, size=(160, 160))
    B = pb.Backend(0)
    B.upload_scene(S)
    rays = R.primary_rays(1)
    assert np.array_equal(u32(rays), u32(B.primary_rays(1)))
    hr, hg = R.trace_closest(rays, surf=True), B.trace_closest(rays, surf=True)
    assert np.array_equal(hr["geom"], hg["geom"]) and np.array_equal(u32(hr["t"]), u32(hg["t"]))
    assert np.array_equal(u32(hr["surf"][:, :6]), u32(hg["surf"][:, :6]))
    film_ref, _ = R.render(cfg["spp"], cfg["depth"])
    film = B.render(cfg["spp"], cfg["depth"])
    assert relmse(film, film_ref) <= FILM_RELMSE_TOL
    # Film::finalize (Uncharted2) + 8-bit quantisation: device epilogue vs the reference's
    rgba8, mapped = B.finalize(film_ref, flip_y=False)
    want = R.finalize(film_ref)
    np.testing.assert_allclose(mapped[..., :3], want[..., :3], rtol=2e-6, atol=1e-7)
    q = np.clip(np.power(want[..., :3].astype(np.float32), np.float32(1 / 2.2)) * 256, 0, 255).astype(np.uint8)
    assert (np.abs(rgba8[..., :3].astype(int) - q.astype(int)) <= 1).all()
    # the same epilogue on the film the render left on the device (no upload, 4 B per pixel back) gives the
    # bytes the host-film entry point gives for that film
    r8_dev, mapped_dev = B.finalize_device(flip_y=False, want_mapped=True)
    r8_host, mapped_host = B.finalize(film, flip_y=False)
    assert np.array_equal(r8_dev, r8_host) and np.array_equal(u32(mapped_dev), u32(mapped_host))
    assert np.array_equal(B.finalize_device(flip_y=True), r8_host[::-1])
    B.close()
