"""The C-ABI shared library: loads, exports every symbol include/*.h declares, and refuses to
compute without a CUDA device (there is no CPU fallback). No GPU needed."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import pine_b200 as pb
from pine_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in ("pine_b200.h", "pine_b200_host.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(pb2h?_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    L = C.CDLL(A.LIB_PATH)
    declared = declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/ but not exported"
    assert sorted(set(A.ABI_SYMBOLS)) == declared, "pine_b200/_abi.py ABI_SYMBOLS out of date"


def test_struct_sizes_match_header():
    # sizes the C side static_asserts / the header implies
    assert C.sizeof(A.Rect) == 80 and C.sizeof(A.Obb) == 152 and C.sizeof(A.Cone) == 96
    assert C.sizeof(A.Disk) == 52 and C.sizeof(A.Sphere) == 16 and C.sizeof(A.Triangle) == 48
    assert C.sizeof(A.Material) == 76 and C.sizeof(A.Geometry) == 12 and C.sizeof(A.NodeInstr) == 16
    assert C.sizeof(A.RenderParams) == 52


def test_round_spp_and_sampler_ctor():
    L = A.lib()
    # BlueSobolSampler ctor (sampler.cpp:115-121): round up to a power of two, clamp at 256
    assert [L.pb2h_round_spp(v) for v in (1, 2, 3, 16, 17, 255, 256, 300)] == [1, 2, 4, 16, 32, 256, 256, 256]
    with pytest.warns(UserWarning):
        assert pb.BlueSampler(1000).spp() == 256
    with pytest.raises(ValueError):
        pb.PathIntegrator(pb.BlueSampler(4), 0)  # path.cpp:12-13


def _no_gpu():
    return A.lib().pb2_device_count() == 0


@pytest.mark.skipif(not _no_gpu(), reason="a CUDA device is present")
def test_compute_calls_fail_loudly_without_a_gpu():
    with pytest.raises(pb.PineB200Error) as e:
        pb.Backend(0)
    assert e.value.code == A.PB2_E_CUDA and "no CPU fallback" in str(e.value)
    # a host-only context can build trees but every compute entry point refuses
    from pine_b200 import scenes
    S = pb.Scene()
    scenes.cbox_readme(S, size=(8, 8))
    B = pb.Backend(-1)
    B.upload_scene(S)
    B._tables_loaded.add(4)
    for call in (lambda: B.render(4, 4), lambda: B.trace_closest(np.zeros((1, 8), np.float32)),
                 lambda: B.trace_any(np.zeros((1, 8), np.float32)), lambda: B.primary_rays(0)):
        with pytest.raises(pb.PineB200Error):
            call()


def test_scene_validation_errors():
    B = pb.Backend(-1)
    S = pb.Scene()
    m = S.add_material("emissive", (1, 1, 1))
    S.add_box((0, 0, 0), (1, 1, 1), np.eye(4, dtype=np.float32), m)  # emissive Box: undefined in the reference
    S.set_camera((4, 4), (0, 0, -3), (0, 0, 0), 0.3)
    with pytest.raises(pb.PineB200Error) as e:
        B.upload_scene(S)
    assert e.value.code == A.PB2_E_UNSUPPORTED
    S2 = pb.Scene()
    with pytest.raises(ValueError):
        S2.desc()  # no camera
    with pytest.raises(KeyError):
        S2.add(pb.Rect((0, 0, 0), (1, 0, 0), (0, 1, 0)), "missing-material")  # scene.cpp:52-57


def test_empty_scene_builds():
    B = pb.Backend(-1)
    S = pb.Scene()
    S.set_camera((4, 4), (0, 0, -3), (0, 0, 0), 0.3)
    B.upload_scene(S)
    assert B.counters()["wide_nodes"] == 0


def test_built_scene_is_shared_between_contexts_through_a_file(tmp_path):
    """pb2_export_host_scene / pb2_import_host_scene (one host build per node with one process per GPU): the
    importer holds the same staging bytes and scene state as the builder -- its own export is the same file."""
    from pine_b200 import scenes
    S = pb.Scene()
    scenes.mesh_field(S, n_instances=27, subdiv=2, seed=3, size=(64, 48))
    # round-2 arrays travel too: node programs, a Plane, a Cylinder
    from pine_b200 import nodes as N
    S.add_plane((0, -2, 0), (0, 1, 0), S.add_material("uber", N.lerp(N.Checkerboard(N.UV()), (1, 0, 0), (0, 0, 1)), roughness=0.3))
    S.add_cylinder((0, 0, 0), (0, 1, 0), 0.2, 0)
    builder, importer = pb.Backend(-1), pb.Backend(-1)
    builder.upload_scene(S)
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    builder.export_scene(a)
    importer.import_scene(a, S.size)
    ca, cb = builder.counters(), importer.counters()
    assert all(ca[k] == cb[k] for k in ("bvh2_nodes", "wide_nodes", "n_primitives", "scene_bytes"))
    importer.export_scene(b)
    assert open(a, "rb").read() == open(b, "rb").read()
    with open(a, "r+b") as f:  # a foreign / truncated file is refused
        f.write(b"XXXXXXXX")
    with pytest.raises(pb.PineB200Error):
        pb.Backend(-1).import_scene(a, S.size)
    with pytest.raises(pb.PineB200Error):
        pb.Backend(-1).export_scene(b)  # nothing built
