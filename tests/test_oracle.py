"""The oracle against the golden vectors (no GPU).

tests/golden/*.npz are outputs of the unmodified reference compiled in place (oracle/_ref, made by
tests/golden/make_golden.py). Here the plain-C++ restatement oracle/oracle_port.cpp is pinned
against them bit for bit: hash / RNG / BlueSobol sampler, camera rays, spawn_ray, the binned-SAH
builder and the reference-order closest-hit / any-hit walks. Where oracle/_ref itself is present it is
re-run and must reproduce the committed vectors (the fixtures are not stale).
"""
import os

import numpy as np
import pytest

import common
import pine_b200 as pb
from common import u32
from conftest import needs_ref
from oracle import port
from pine_b200.integrator import sampler_tables

SCENES = list(common.SMALL_SCENES)


@pytest.fixture(scope="module")
def kats(golden):
    return golden("kats")


def test_hash_and_rng_known_answers(kats):
    # murmur_hash64A over (px, py, sample_index) and the xoroshiro step (rng.h:9-45,97-144)
    for (px, py, si), want in zip(kats["hash_in"], kats["hash_out"]):
        assert port.hash_pixel(int(px), int(py), int(si)) == int(want)
    for seed, want in zip(kats["rng_seeds"], kats["rng_out"]):
        assert np.array_equal(port.rng_u64(int(seed), 8), want)


def test_bluenoise_table_lookup(kats):
    # bluenoise_Nspp(px, py, index, dim) (bluenoise_16spp.cpp:14-34) on a strided grid
    tabs = sampler_tables()
    for a, spp in enumerate((1, 4, 16, 64, 256)):
        got = np.zeros((8, 8, 4, 8), np.float32)
        for px in range(8):
            for py in range(8):
                for idx in range(4):
                    for dim in range(8):
                        got[px, py, idx, dim] = port.bluenoise(tabs[spp], px * 17, py * 29, idx * 5, dim * 3)
        assert np.array_equal(u32(got), u32(kats["bluenoise_grid"][a])), spp


def test_sampler_streams_and_dimension_wrap(kats):
    # Sampler::start_pixel + rand draws + get1d/get2d/start_next_sample (sampler.h:166-201,286-309)
    tabs = sampler_tables()
    for spp in (1, 2, 16, 64, 256):
        for i, (px, py) in enumerate(kats["pixels"]):
            r, d = port.sampler_stream(tabs[spp], int(px), int(py), i, 8, "21221", 6)
            assert np.array_equal(u32(r), u32(kats[f"rand_{spp}_{i}"]))
            assert np.array_equal(u32(d), u32(kats[f"draws_{spp}_{i}"]))
        _, d = port.sampler_stream(tabs[spp], 5, 7, 0, 0, "2" * 140 + "1", 1)
        assert np.array_equal(u32(d), u32(kats[f"wrap_{spp}"]))


def test_sobol_sampler_host_draws_match_reference_vectors(golden):
    # SobolSampler (sampler.h:83-164, sampler.cpp:82-113): the library's __host__ __device__ draw functions,
    # run on the host, against streams recorded from the compiled reference (tests/golden/sobol_kats.npz)
    from pine_b200.integrator import host_sampler_draws
    k = golden("sobol_kats")
    for spp, w, h in k["cases"]:
        spp, size = int(spp), (int(w), int(h))
        for i, (px, py) in enumerate(k["pixels"]):
            d = host_sampler_draws(pb.SobolSampler(spp), size, int(px), int(py), 0, "21221", min(6, spp))
            assert np.array_equal(u32(d), u32(k[f"draws_{spp}_{i}"])), (spp, size, i)
        d = host_sampler_draws(pb.SobolSampler(spp), size, 5, 7, 0, "2" * 60 + "1" * 10, 1)
        assert np.array_equal(u32(d), u32(k[f"long_{spp}"])), spp


@needs_ref
def test_sobol_sampler_host_draws_match_live_reference():
    from oracle import ref
    from pine_b200.integrator import host_sampler_draws
    rng = np.random.RandomState(11)
    for spp in (1, 2, 5, 8, 64, 1000, 65536):
        for _ in range(6):
            size = (int(rng.randint(1, 4000)), int(rng.randint(1, 2200)))
            px, py = int(rng.randint(size[0])), int(rng.randint(size[1]))
            want = ref.sampler_stream_kind("sobol", spp, size, px, py, 0, "1221", min(spp, 9))
            got = host_sampler_draws(pb.SobolSampler(spp), size, px, py, 0, "1221", min(spp, 9))
            assert np.array_equal(u32(got), u32(want)), (spp, size, px, py)


def test_halton_sampler_host_draws_match_reference_vectors(golden):
    # HaltonSampler (sampler.h:40-81, sampler.cpp:20-81, lowdiscrepancy.h:26-60): primes, digit permutations
    # (rebuilt with the reference's generator), CRT pixel offset and the dimension wrap at 1000
    from pine_b200.integrator import host_sampler_draws
    k = golden("halton_kats")
    for spp in k["spps"]:
        spp = int(spp)
        for i, (px, py) in enumerate(k["pixels"]):
            d = host_sampler_draws(pb.HaltonSampler(spp), (640, 640), int(px), int(py), 0, "21221", min(6, spp))
            assert np.array_equal(u32(d), u32(k[f"draws_{spp}_{i}"])), (spp, i)
    d = host_sampler_draws(pb.HaltonSampler(4), (64, 64), 5, 7, 0, "2" * 520 + "1" * 3, 2)
    assert np.array_equal(u32(d), u32(k["wrap"]))
    with pytest.raises(ValueError):
        pb.HaltonSampler(0)


def test_uniform_sampler_host_draws_are_uniform_and_keyed():
    # UniformSampler (sampler.h:13-39) is statistical parity only: check the counter-based stream's
    # moments, its range (nextf < 1, rng.h:140-143) and that pixel / sample / dimension / seed all key it
    from pine_b200.integrator import host_sampler_draws
    d = host_sampler_draws(pb.UniformSampler(4096), (64, 64), 3, 9, 0, "2" * 8, 4096).astype(np.float64)
    assert d.min() >= 0 and d.max() < 1
    assert abs(d.mean() - 0.5) < 4 / np.sqrt(12 * d.size) and abs(d.var() - 1 / 12) < 2e-3
    c = np.corrcoef(d.T)
    assert np.abs(c - np.eye(16)).max() < 0.08
    a = host_sampler_draws(pb.UniformSampler(4), (64, 64), 3, 9, 0, "21", 4)
    assert np.array_equal(a, host_sampler_draws(pb.UniformSampler(4), (64, 64), 3, 9, 0, "21", 4))
    assert not np.array_equal(a, host_sampler_draws(pb.UniformSampler(4), (64, 64), 4, 9, 0, "21", 4))
    assert not np.array_equal(a, host_sampler_draws(pb.UniformSampler(4, seed=1), (64, 64), 3, 9, 0, "21", 4))
    assert len(np.unique(a)) == a.size


def test_nee_draw_order_is_pinned(kats):
    # light_sampler.sample(p, get1d(), get2d()) under g++: get2d is drawn first (lightsampler.h:27)
    assert int(kats["nee_draw_order_2d_first"][0]) == 1


def test_thin_lens_camera(golden):
    g = golden("camera_lens")
    S = pb.Scene()
    m = S.add_material("diffuse", (0.5, 0.5, 0.5))
    S.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), m)
    S.set_camera((32, 24), (0.3, 1.1, -4), (0.1, 0.9, 0.2), 0.3, len_radius=0.05, focus=3.5)
    P = port.PortScene(S)
    got = P.gen_rays(g["inp"])
    # the lens sample goes through libm sin/cos (sampling.h:13-18): same glibc => same bits here
    assert np.array_equal(u32(got), u32(g["rays"]))


@pytest.mark.parametrize("name", SCENES)
def test_port_traversal_matches_reference_vectors(golden, name):
    g = golden(name)
    S = pb.Scene()
    common.build(S, name)
    P = port.PortScene(S)
    assert np.array_equal(u32(P.primary_rays(0)), u32(g["primary0"]))
    assert np.array_equal(u32(P.primary_rays(2)), u32(g["primary2"]))
    h = P.trace_closest(g["primary0"])
    assert np.array_equal(h["geom"], g["p_geom"]) and np.array_equal(h["prim"], g["p_prim"])
    assert np.array_equal(u32(h["t"]), u32(g["p_t"]))
    assert np.array_equal(h["n_prim_tests"], g["p_n_prim_tests"])  # same leaves visited, same order
    assert np.array_equal(u32(port.spawn_rays(g["spawn_in"])), u32(g["secondary"]))
    h = P.trace_closest(g["secondary"])
    assert np.array_equal(h["geom"], g["s_geom"]) and np.array_equal(h["prim"], g["s_prim"])
    assert np.array_equal(u32(h["t"]), u32(g["s_t"]))
    assert np.array_equal(P.trace_any(g["shadow"])["occluded"], g["shadow_occ"])


def _same_tree(a, g, prefix):
    assert a["root"] == int(g[prefix + "root"])
    assert np.array_equal(u32(a["aabbs"]), u32(g[prefix + "aabbs"]))
    assert np.array_equal(a["children"], g[prefix + "children"])
    lr, want = a["leaf_range"], g[prefix + "leaf_range"]
    assert np.array_equal(lr[:, 1], want[:, 1])  # primitives per node (0 = internal)
    leaf = want[:, 1] > 0
    assert np.array_equal(lr[leaf, 0], want[leaf, 0])  # offset is meaningful for leaves only
    assert np.array_equal(a["leaf_prims"], g[prefix + "leaf_prims"])


@pytest.mark.parametrize("name", SCENES)
def test_sah_builders_reproduce_reference_trees(golden, name):
    """build_sah_binned (bvh.cpp:43-147): the oracle port AND the product's host builder
    (pine_b200/csrc/bvh_build.cpp, through the C ABI on a host-only context) give the reference's
    node array node for node, so the reference-order walk and the leaf gates rest on the same tree."""
    g = golden(name)
    S = pb.Scene()
    common.build(S, name)
    P = port.PortScene(S)
    B = pb.Backend(-1)
    B.upload_scene(S)
    for tree in (P.bvh_tree(-1), B.bvh_tree(-1)):
        _same_tree(tree, g, "top_")
    if int(g["n_lbvh"][0]):
        for tree in (P.bvh_tree(0), B.bvh_tree(0)):
            _same_tree(tree, g, "mesh0_")


@needs_ref
@pytest.mark.parametrize("name", SCENES)
def test_compiled_reference_reproduces_the_fixtures(golden, name):
    from oracle import ref
    g = golden(name)
    R = ref.RefScene()
    common.build(R, name)
    _, _, spp, depth = common.SMALL_SCENES[name]
    assert np.array_equal(u32(R.primary_rays(0)), u32(g["primary0"]))
    h = R.trace_closest(g["secondary"], surf=True)
    assert np.array_equal(h["geom"], g["s_geom"]) and np.array_equal(u32(h["t"]), u32(g["s_t"]))
    assert np.array_equal(u32(h["surf"]), u32(g["s_surf"]))
    assert np.array_equal(R.trace_any(g["shadow"]), g["shadow_occ"])
    film, _ = R.render(spp, depth)
    assert np.array_equal(u32(film), u32(g["film"]))  # the reference render is deterministic


def test_embree_backed_reference_renders_the_same_film(tmp_path):
    """oracle/_ref/libpine_ref_embree.so = the same reference with its PRL-default accel (EmbreeAccel over the
    vendored Embree 4.3.1; `make -C oracle embree`). It is a TIMING baseline (bench.py cpu_baseline_embree); here
    its film is checked against the pine::BVH reference's: same integrator and sampler streams, different
    triangle kernels => equal up to rounding-induced path changes."""
    import subprocess
    import sys
    from oracle import ref
    if not (ref.available() and ref.embree_available()):
        pytest.skip("Embree-backed reference not built")
    code = ("import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r + '/tests');"
            "import common; from oracle import ref; R = ref.RefScene(); common.build(R, 'mesh_48');"
            "ref.use_embree_accel(%s); R.bounded_prepare(4, 5); t, f = R.bounded_render_rows(0, R.size[1], 1, True);"
            "np.save(%r, f)")
    films = {}
    for name, flag, env in (("bvh", "False", {}), ("embree", "True", {"PINE_REF_EMBREE": "1"})):
        out = str(tmp_path / f"{name}.npy")
        r = subprocess.run([sys.executable, "-c", code % (common.ROOT, common.ROOT, flag, out)],
                           env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        films[name] = np.load(out)
    assert common.relmse(films["embree"], films["bvh"]) <= 1e-3


@needs_ref
def test_port_plane_and_cylinder_match_the_live_reference():
    """The oracle's restatement of Plane (geometry.cpp:31-70) and Cylinder (:466-517) -- boxes, the non-strict /
    unchecked comparisons, the reference-order walk over shapes that can be hit outside their box -- against the
    live reference: camera rays, bounce rays from the hit points and shadow-length rays, (geom, t) bit for bit."""
    import pine_b200 as pb
    from oracle import port, ref
    from pine_b200 import scenes
    R, S = ref.RefScene(), pb.Scene()
    scenes.plane_yard(R, size=(96, 72))
    scenes.plane_yard(S, size=(96, 72))
    P = port.PortScene(S)
    rays = R.primary_rays(0)
    want = R.trace_closest(rays, surf=True)
    got = P.trace_closest(rays)
    assert set(np.unique(want["geom"])) >= {0, 1, 2, 3, 4}
    assert np.array_equal(got["geom"], want["geom"]) and np.array_equal(common.u32(got["t"]), common.u32(want["t"]))
    sec = common.secondary_inputs(want["surf"], want["geom"], 5)
    # spawn rays off the surfaces (o = p + n * 1e-3, the direction of secondary_inputs), long and shadow-length
    o = sec[:, 0:3] + sec[:, 3:6] * np.float32(1e-3)
    bounce = np.concatenate([o, sec[:, 6:9], np.zeros((len(sec), 1), np.float32), sec[:, 9:10]], axis=1).astype(np.float32)
    w2, g2 = R.trace_closest(bounce), P.trace_closest(bounce)
    assert np.array_equal(g2["geom"], w2["geom"]) and np.array_equal(common.u32(g2["t"]), common.u32(w2["t"]))
    shadow = common.shadow_from(bounce, 9)
    assert np.array_equal(P.trace_any(shadow)["occluded"], R.trace_any(shadow))
