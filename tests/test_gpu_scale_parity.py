"""The CUDA path against the LIVE reference (oracle/_ref, the unmodified C++ compiled in place; it travels to
the GPU box) at the sizes the benchmark runs: BASELINE config 1 at 640x640, the 10.24 M-triangle geometry of
config 5, and the room-proc interior of config 4. Closest hits (geom, prim, t) and shadow-ray occlusion are
compared BIT FOR BIT on every camera ray of a sample plus every bounce and shadow ray a production render
(wide tree, certify-or-fall-back) actually traced (pb2_set_ray_record); films at equal spp to relMSE <= 1e-4.
Reference functions matched: BVH::intersect / BVH::hit (src/pine/impl/accel/bvh.cpp:497-548) and
PathIntegrator::render (src/pine/impl/integrator/path.cpp:26-41).
"""
import os

import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32
from conftest import needs_ref
from pine_b200 import scenes

pytestmark = [pytest.mark.gpu, needs_ref]
FILM_RELMSE_TOL = 1e-4


def _pair(fn, **kw):
    from oracle import ref
    R, S = ref.RefScene(), pb.Scene()
    scenes.SCENES[fn](R, **kw)
    cfg = scenes.SCENES[fn](S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    return R, B, cfg


def _check_rays(R, B, rec_spp, depth, min_rays):
    """camera rays of sample 0 + every ray of a recorded `rec_spp`-sample production render vs the reference"""
    rays = R.primary_rays(0)
    assert np.array_equal(u32(rays), u32(B.primary_rays(0)))
    film, rec = B.render_recorded(rec_spp, depth, accel=0)
    closest = np.concatenate([rays, rec["closest_rays"]])
    n_total = len(closest) + len(rec["shadow_rays"])
    assert n_total >= min_rays, n_total
    got = B.trace_closest(rays, accel=0)
    want = R.trace_closest(closest)
    g = np.concatenate([got["geom"], rec["closest_geom"]])
    p = np.concatenate([got["prim"], rec["closest_prim"]])
    t = np.concatenate([got["t"], rec["closest_t"]])
    assert np.array_equal(g, want["geom"]), int((g != want["geom"]).sum())
    assert np.array_equal(p, want["prim"]), int((p != want["prim"]).sum())
    assert np.array_equal(u32(t), u32(want["t"])), int((u32(t) != u32(want["t"])).sum())
    assert np.array_equal(rec["shadow_occluded"], R.trace_any(rec["shadow_rays"]))
    # the film of the recorded render is the film of an unrecorded one
    assert np.array_equal(u32(film), u32(B.render(rec_spp, depth)))
    return n_total, B.counters()


def test_c1_640_every_ray_and_film_vs_live_reference():
    """BASELINE config 1 at full size: all 409 600 camera rays + >= 1 M bounce and shadow rays, and the 16-spp film."""
    R, B, cfg = _pair("cbox_readme", size=(640, 640))
    n, c = _check_rays(R, B, 4, cfg["depth"], 5_000_000)
    assert c["last_closest_fallbacks"] > 0   # the two Box shapes send rays through the reference-order walk
    film_ref, _ = R.render(cfg["spp"], cfg["depth"])
    film = B.render(cfg["spp"], cfg["depth"])
    assert relmse(film, film_ref) <= FILM_RELMSE_TOL
    B.close()


def test_c5_geometry_rays_and_film_vs_live_reference():
    """The 10.24 M-triangle geometry of config 5 (2000 x icosphere(4) under one top-level tree): >= 2 M rays of
    every kind and the c5s film (960x540, 16 spp, depth 6)."""
    R, B, cfg = _pair("mesh_field", n_instances=2000, subdiv=4, seed=42, size=(960, 540), spp=16, depth=6,
                      with_shapes=False)
    assert B.counters()["n_primitives"] >= 10_240_000
    n, c = _check_rays(R, B, 4, cfg["depth"], 2_000_000)
    film_ref, _ = R.render(cfg["spp"], cfg["depth"])
    film = B.render(cfg["spp"], cfg["depth"])
    assert relmse(film, film_ref) <= FILM_RELMSE_TOL
    B.close()


def test_c4_room_proc_rays_and_film_vs_live_reference():
    """Config 4's stand-in interior (2.0 M triangles, closed room): >= 2 M rays and the film at 480x270, 16 spp."""
    R, B, cfg = _pair("room_proc", size=(480, 270), spp=16, depth=6)
    n, c = _check_rays(R, B, 4, cfg["depth"], 2_000_000)
    film_ref, _ = R.render(cfg["spp"], cfg["depth"])
    film = B.render(cfg["spp"], cfg["depth"])
    assert relmse(film, film_ref) <= FILM_RELMSE_TOL
    B.close()


@pytest.mark.parametrize("name", ["c1_64", "mesh_48"])
def test_pixel_means_within_three_sigma_of_the_oracle(name):
    """|mean_gpu - mean_ref| <= 3 sigma per pixel and channel at equal spp, sigma = standard error of the mean
    from the ORACLE's per-sample radiances; the absolute slack covers pixels whose samples
    are all equal (sigma = 0) against device-libm rounding."""
    from oracle import ref
    R, S = ref.RefScene(), pb.Scene()
    common.build(R, name)
    common.build(S, name)
    _, _, spp, depth = common.SMALL_SCENES[name]
    spp = 16
    B = pb.Backend(0)
    B.upload_scene(S)
    film_ref, mean, sigma = R.render_moments(spp, depth)
    film = B.render(spp, depth)[..., :3].astype(np.float64)
    assert np.abs(mean - film_ref[..., :3]).max() <= 1e-5 * (1 + np.abs(mean).max())
    assert (np.abs(film - mean) <= 3 * sigma + 1e-4 * (1 + np.abs(mean))).all()
    B.close()


@pytest.mark.parametrize("name", ["c1_64", "cones_48"])
def test_sample_shards_equal_the_references_own_shards(golden, name):
    """Four sample shards: every shard's film equals the film the REFERENCE renders for the same
    sample range with its sampler started at the shard's first sample (Sampler::start_pixel(p, sample_index),
    sampler.h:286-290) to relMSE <= 1e-4, and the summed film's error against the converged 4096-spp film is
    the reference's own sharded error within 2 %. (Re-seeding per shard changes the estimate -- on cones_48 the
    shard starting at sample 48 carries a firefly in BOTH implementations -- which is why the comparison is
    with the reference's shards, not with its unsharded film.)"""
    from oracle import ref
    spp = 64
    g = golden(name)
    R, S = ref.RefScene(), pb.Scene()
    common.build(R, name)
    common.build(S, name)
    depth = common.SMALL_SCENES[name][3]
    B = pb.Backend(0)
    B.upload_scene(S)
    acc, acc_ref = 0, 0
    for s in range(0, spp, spp // 4):
        f, fr = B.render(spp, depth, samples=(s, s + spp // 4)), R.render_sample_range(spp, depth, s, s + spp // 4)
        assert relmse(f * 4, fr * 4) <= FILM_RELMSE_TOL, s
        assert np.array_equal(f[..., 3], fr[..., 3])
        acc, acc_ref = acc + f.astype(np.float64), acc_ref + fr.astype(np.float64)
    assert np.allclose(acc[..., 3], 1.0)
    e_shard, e_ref = relmse(acc, g["film_converged"]), relmse(acc_ref, g["film_converged"])
    assert abs(e_shard - e_ref) <= 0.02 * e_ref + 1e-6, (e_shard, e_ref)
    B.close()


def _with_env(env, fn):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return fn()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.parametrize("fn,kw", [("cbox_readme", dict(size=(96, 80))), ("materials_room", dict(size=(64, 48)))])
def test_wave_plan_splits_are_bit_identical(fn, kw):
    """Pixel waves (pixel0 offsets into the RNG / accumulator arrays), sample waves (RNG state reload, sample0 of
    the sampler index), one or two lanes in flight: every plan renders the same film bits as the single wave.
    materials_room has Uber / Subsurface materials, whose randf() draws force one sample per wave."""
    S = pb.Scene()
    cfg = scenes.SCENES[fn](S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    w = kw["size"][0]
    spp, depth = 8, cfg["depth"]
    base = _with_env({"PB2_LANES": 1}, lambda: B.render(spp, depth))
    for env in ({"PB2_MAX_SLOTS": w * 7, "PB2_LANES": 1},                      # many pixel waves x 8 sample waves
                {"PB2_MAX_SLOTS": w * 7 * 3, "PB2_LANES": 1},                  # s_batch = 3: ragged last sample wave
                {"PB2_MAX_SLOTS": w * 7 * 3, "PB2_LANES": 2, "PB2_MIN_LANE_SLOTS": 1},
                {"PB2_LANES": 2, "PB2_MIN_LANE_SLOTS": 1},                      # two half-film waves in flight
                {"PB2_LANES": 4, "PB2_MIN_LANE_SLOTS": 1, "PB2_MAX_SLOTS": w * 5 * 2},
                {"PB2_LANES": 3, "PB2_MIN_LANE_SLOTS": 1}):
        film = _with_env(env, lambda: B.render(spp, depth))
        assert np.array_equal(u32(film), u32(base)), env
    rows = _with_env({"PB2_LANES": 2, "PB2_MIN_LANE_SLOTS": 1}, lambda: B.render(spp, depth, rows=(5, 38)))
    assert np.array_equal(u32(rows[5:38]), u32(base[5:38])) and (rows[:5] == 0).all()
    B.close()


def test_progress_callback_follows_the_completed_waves():
    """pb2_set_progress_callback (path.cpp:39's set_progress): one call per completed wave, fractions k / n_waves
    each reported once, ending at 1; the film is the film rendered without a callback."""
    S = pb.Scene()
    cfg = scenes.cbox_readme(S, size=(64, 64))
    B = pb.Backend(0)
    B.upload_scene(S)
    base = B.render(8, cfg["depth"])
    seen = []
    B.set_progress_callback(seen.append)
    for env, n_waves in (({"PB2_LANES": 1}, 1),
                         ({"PB2_MAX_SLOTS": 64 * 8 * 2, "PB2_LANES": 1}, 8 * 4),      # 4 pixel waves x 8 sample waves
                         ({"PB2_MAX_SLOTS": 64 * 16 * 8, "PB2_LANES": 4, "PB2_MIN_LANE_SLOTS": 1}, 4)):
        del seen[:]
        film = _with_env(env, lambda: B.render(8, cfg["depth"]))
        assert np.array_equal(u32(film), u32(base)), env
        assert len(seen) == n_waves and max(seen) == 1.0, (env, seen)
        assert sorted(seen) == [np.float32(k) / np.float32(n_waves) for k in range(1, n_waves + 1)]
        if env.get("PB2_LANES") == 1:
            assert seen == sorted(seen)   # one stream: waves complete in order
    B.set_progress_callback(None)
    del seen[:]
    B.render(8, cfg["depth"])
    assert seen == []
    B.close()



@pytest.mark.parametrize("fn,kw,masks", [
    ("cbox_indirect", dict(size=(64, 64)), (32, 160, 236, 51, 252, 255, 1023)),                 # Rects only
    ("mesh_field", dict(n_instances=8, subdiv=2, size=(64, 36), spp=4, depth=5, with_shapes=True), (236, 252, 255, 1023)),
    ("cbox_readme", dict(size=(64, 64)), (51, 255, 1023)),                                       # Rect + Cone + OBB
])
def test_every_shape_mask_instantiation_renders_the_same_film(fn, kw, masks):
    """The traversal kernels are instantiated per shape-kind mask (shapes.cuh kShapes*) and a render picks the first
    that covers the scene; every OTHER instantiation that covers it (forced with PB2_TRACE_SHAPES /
    PB2_SHADOW_SHAPES) gives the same film bits, and a mask that does not cover the scene is ignored."""
    S = pb.Scene()
    cfg = scenes.SCENES[fn](S, **kw)
    B = pb.Backend(0)
    B.upload_scene(S)
    spp, depth = 4, cfg["depth"]
    base = B.render(spp, depth)
    for m in masks:
        film = _with_env({"PB2_TRACE_SHAPES": m, "PB2_SHADOW_SHAPES": m}, lambda: B.render(spp, depth))
        assert np.array_equal(u32(film), u32(base)), m
    film = _with_env({"PB2_TRACE_SHAPES": 0, "PB2_SHADOW_SHAPES": 4}, lambda: B.render(spp, depth))  # not a cover
    assert np.array_equal(u32(film), u32(base))
    B.close()
