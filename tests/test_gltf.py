"""The glTF / .glb front end (pine_b200/gltf.py = scene_from_gltf, fileio.cpp:146-327), no GPU.

tests/golden/room_like.glb is written by tests/gltf_fixture.py; tests/golden/room_like.npz holds what the
unmodified reference's loader + BVH made of that file (make_golden.py gltf_kats). Here the scene the
product loader builds is traced by the oracle port on the CPU and must give the reference's hits bit for bit,
which pins vertices, indices, node transforms and the camera.
"""
import json
import os
import struct

import numpy as np
import pytest

import common
import pine_b200 as pb
from common import u32
from conftest import needs_ref
from oracle import port
from pine_b200 import _abi as A
from pine_b200 import gltf

GLB = os.path.join(common.GOLDEN, "room_like.glb")
GLB_TEXTURED = os.path.join(common.GOLDEN, "room_textured.glb")
REF_CBOX = "/root/reference/scenes/cbox.glb"


@pytest.fixture(scope="module")
def loaded():
    S = pb.Scene()
    n = gltf.scene_from_gltf(S, GLB)
    return S, n


def test_fixture_is_what_the_writer_writes():
    import gltf_fixture
    with open(GLB, "rb") as f:
        assert f.read() == gltf_fixture.room_like()
    with open(GLB_TEXTURED, "rb") as f:
        assert f.read() == gltf_fixture.room_textured()


def test_image_node_materials_load():
    # fileio.cpp:276-292: baseColorTexture -> albedo image; metallicRoughnessTexture -> roughness (unless a
    # roughnessFactor is given, applied after it in std::map order) and metallic
    S = pb.Scene()
    assert gltf.scene_from_gltf(S, GLB_TEXTURED) == 5
    assert [t.shape for t in S.textures] == [(23, 37, 4), (16, 16, 4)]  # one texture per image, RGBA
    rng = np.random.RandomState(7)
    color = rng.randint(30, 256, size=(23, 37, 3)).astype(np.uint8)
    assert np.array_equal(S.textures[0][..., :3], color) and (S.textures[0][..., 3] == 255).all()
    assert set(np.unique(S.textures[1][..., 2])) == {0, 255}
    slots = [(m.color_texture, m.roughness_texture, m.metallic_texture) for m in S.materials]
    assert slots == [(1, 0, 0), (0, 2, 2), (1, 0, 2), (1, 0, 0), (0, 0, 0)]
    assert np.float32(S.materials[2].roughness) == np.float32(0.35)
    S.set_camera((8, 8), (0, 1, 4), (0, 1, 0), 0.4)
    d, keep = S.desc()
    assert d.n_textures == 2 and d.textures[0].width == 37 and d.textures[0].height == 23
    with pytest.raises(ValueError):
        S.add_texture(np.zeros((4, 4), np.uint8))


def test_geometries_materials_and_camera(golden, loaded):
    g = golden("room_like")
    S, n = loaded
    assert n == int(g["n_geometries"][0]) == len(S.geometries) == 8
    assert S.size == tuple(g["size"]) == (96, 640)  # Film(int(640 * aspect), 640), fileio.cpp:318
    boxes = np.array([np.concatenate([v.min(0), v.max(0)]) for v, _, _, _ in S.mesh_data], np.float32)
    assert np.array_equal(u32(boxes), u32(g["geometry_aabbs"].reshape(-1, 6)))
    kinds = [S.materials[m].kind for _, _, m in S.geometries]
    uber, emissive = A.MATERIAL_KINDS["uber"], A.MATERIAL_KINDS["emissive"]
    assert kinds == [uber] * 7 + [emissive] and S.lights() == [7]
    m = {i: S.materials[mat] for i, (_, _, mat) in enumerate(S.geometries)}
    # white wall, ior default 1.45 (fileio.cpp:261); red wall with KHR_materials_ior 1
    assert np.float32(m[0].ior) == np.float32(1.45) and np.float32(m[1].ior) == np.float32(1.0)
    # brass ball; the emissiveFactor-without-strength cube stays Uber (fileio.cpp:294-299)
    assert np.float32(m[2].metallic) == np.float32(0.85) and np.float32(m[2].roughness) == np.float32(0.3)
    assert tuple(np.float32(c) for c in m[3].color) == tuple(np.float32(c) for c in (0.2, 0.3, 0.8))
    # glass cube: transmission + ior extensions
    assert np.float32(m[4].transmission) == np.float32(0.9) and np.float32(m[4].ior) == np.float32(1.5)
    # primitive without a material: Uber(1, 1, 0, 0, 1.45); material without pbr block: tinygltf defaults
    assert (m[5].roughness, m[5].metallic, m[5].transmission) == (1.0, 0.0, 0.0)
    assert (m[6].roughness, m[6].metallic) == (1.0, 1.0) and tuple(m[6].color) == (1.0, 1.0, 1.0)
    assert tuple(np.float32(c) for c in m[7].color) == tuple(np.float32(c) * np.float32(30) for c in (1.0, 0.85, 0.6))


def test_loaded_scene_traces_like_the_reference(golden, loaded):
    g = golden("room_like")
    P = port.PortScene(loaded[0])
    assert np.array_equal(u32(P.primary_rays(0)), u32(g["primary0"]))
    for tag, rays in (("p", g["primary0"]), ("f", g["fan"])):
        h = P.trace_closest(rays)
        assert np.array_equal(h["geom"], g[f"{tag}_geom"]) and np.array_equal(h["prim"], g[f"{tag}_prim"])
        assert np.array_equal(u32(h["t"]), u32(g[f"{tag}_t"]))
    assert set(np.unique(g["f_geom"])) == set(range(8))  # the fan reaches every geometry


def _rewrite(mutate):
    """room_like.glb with its JSON chunk passed through `mutate(doc)`."""
    with open(GLB, "rb") as f:
        data = f.read()
    jlen = struct.unpack_from("<I", data, 12)[0]
    doc = json.loads(data[20:20 + jlen])
    rest = data[20 + jlen:]
    mutate(doc)
    js = json.dumps(doc).encode()
    js += b" " * (-len(js) % 4)
    return struct.pack("<4sII", b"glTF", 2, 20 + len(js) + len(rest)) + struct.pack("<I4s", len(js), b"JSON") + js + rest


def test_interleaved_and_embedded_variants_load_identically(loaded):
    # byteStride views: the reference reads them as tightly packed (fileio.cpp:203-207) and so garbles
    # interleaved files; this loader honours the stride. Checked against its own packed load.
    import gltf_fixture
    w = gltf_fixture.GlbWriter()
    v, i, n = gltf_fixture.box((-1, 0, -1), (1, 2, 1))
    for inter in (True, False):
        w.node(root=True, mesh=w.mesh([w.primitive(v, i, None, normals=n, interleave=inter)]), translation=[0.5, 0, 0])
    S = pb.Scene()
    assert gltf.scene_from_gltf(S, w.tobytes()) == 2
    (va, ia, na, _), (vb, ib, nb, _) = S.mesh_data
    assert np.array_equal(va, vb) and np.array_equal(ia, ib) and np.array_equal(na, nb)
    assert np.array_equal(va, v + np.float32([0.5, 0, 0]))
    # .gltf JSON with a base64 data: URI instead of the BIN chunk
    import base64
    with open(GLB, "rb") as f:
        data = f.read()
    jlen = struct.unpack_from("<I", data, 12)[0]
    doc = json.loads(data[20:20 + jlen])
    blen = struct.unpack_from("<I", data, 20 + jlen)[0]
    doc["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(
        data[28 + jlen:28 + jlen + blen]).decode()
    S2 = pb.Scene()
    gltf.scene_from_gltf(S2, json.dumps(doc).encode())
    for (a, b, c, d), (e, f_, g_, h) in zip(S2.mesh_data, loaded[0].mesh_data):
        assert np.array_equal(a, e) and np.array_equal(b, f_)


def test_global_transform_and_load_scene(loaded):
    m = pb.mat4.mul(pb.mat4.translate((1, 2, 3)), pb.mat4.scale((2, 2, 2)))
    S = gltf.load_scene(GLB, m)
    for (v, _, _, _), (v0, _, _, _) in zip(S.mesh_data, loaded[0].mesh_data):
        np.testing.assert_allclose(v, v0 * 2 + np.float32([1, 2, 3]), rtol=1e-6, atol=1e-6)


def test_unsupported_content_is_an_error():
    def dangling_texture(doc):
        doc["materials"][0]["pbrMetallicRoughness"]["baseColorTexture"] = {"index": 0}
    with pytest.raises(gltf.GltfError, match="no source image"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(dangling_texture))

    def lines(doc):
        doc["meshes"][0]["primitives"][0]["mode"] = 1
    with pytest.raises(gltf.GltfError, match="TRIANGLES"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(lines))

    def byte_indices(doc):
        doc["accessors"][doc["meshes"][0]["primitives"][0]["indices"]]["componentType"] = 5121
    with pytest.raises(gltf.GltfError, match="UNSIGNED_SHORT"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(byte_indices))

    def camera_without_rotation(doc):
        del [n for n in doc["nodes"] if "camera" in n][0]["rotation"]
    with pytest.raises(gltf.GltfError, match="camera"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(camera_without_rotation))
    with pytest.raises(gltf.GltfError):
        gltf.scene_from_gltf(pb.Scene(), b"glTF" + struct.pack("<II", 2, 12))
    with pytest.raises(gltf.GltfError, match="version"):
        gltf.scene_from_gltf(pb.Scene(), b"glTF" + struct.pack("<II", 1, 12))


def test_malformed_buffers_are_rejected_not_read_out_of_bounds(tmp_path):
    """Untrusted input: accessors that run past their bufferView / buffer, attribute counts that differ from
    POSITION's, indices beyond the vertex count and uris that leave the file's directory are errors."""
    def too_many(doc):
        doc["accessors"][doc["meshes"][0]["primitives"][0]["attributes"]["POSITION"]]["count"] = 1 << 20
    with pytest.raises(gltf.GltfError, match="do not fit"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(too_many))

    def huge_stride(doc):
        a = doc["accessors"][doc["meshes"][0]["primitives"][0]["attributes"]["POSITION"]]
        doc["bufferViews"][a["bufferView"]]["byteStride"] = 1 << 16
    with pytest.raises(gltf.GltfError, match="do not fit"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(huge_stride))

    def short_normals(doc):
        at = doc["meshes"][0]["primitives"][0]["attributes"]
        if "NORMAL" not in at:
            at["NORMAL"] = at["POSITION"]
        doc["accessors"].append(dict(doc["accessors"][at["NORMAL"]], count=1))
        at["NORMAL"] = len(doc["accessors"]) - 1
    with pytest.raises(gltf.GltfError, match="NORMAL has 1 elements"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(short_normals))

    def few_vertices(doc):
        at = doc["meshes"][0]["primitives"][0]["attributes"]
        at.pop("NORMAL", None); at.pop("TEXCOORD_0", None)
        doc["accessors"].append(dict(doc["accessors"][at["POSITION"]], count=2))
        doc["accessors"][-1].pop("min", None); doc["accessors"][-1].pop("max", None)
        at["POSITION"] = len(doc["accessors"]) - 1
    with pytest.raises(gltf.GltfError, match="beyond POSITION"):
        gltf.scene_from_gltf(pb.Scene(), _rewrite(few_vertices))

    (tmp_path / "sub").mkdir()
    (tmp_path / "secret.bin").write_bytes(b"\0" * 64)
    doc = {"asset": {"version": "2.0"}, "buffers": [{"uri": "../secret.bin", "byteLength": 64}]}
    (tmp_path / "sub" / "x.gltf").write_text(json.dumps(doc))
    with pytest.raises(gltf.GltfError, match="outside the directory"):
        gltf.scene_from_gltf(pb.Scene(), str(tmp_path / "sub" / "x.gltf"))
    with pytest.raises(ValueError):
        pb.Scene().add_mesh(np.zeros((3, 3), np.float32), [[0, 1, 2]], 0, normals=np.zeros((2, 3), np.float32))


def test_prl_scene_load(tmp_path, loaded):
    from pine_b200 import prl
    script = tmp_path / "room.pine"
    script.write_text('scene := Scene();\nscene.load("room_like.glb");\n'
                      'PathIntegrator(BlueSampler(4), 5).render(scene);\n')
    (tmp_path / "room_like.glb").write_bytes(open(GLB, "rb").read())
    sp, = prl.load(script.read_text(), pb.Scene, search_dirs=[str(tmp_path)])
    assert len(sp.target.geometries) == 8 and sp.target.size == (96, 640)
    assert sp.config["spp"] == 4 and sp.config["depth"] == 5
    with pytest.raises(prl.PrlError, match="no such file"):
        prl.load('scene := Scene();\nscene.load("missing.glb");\n', pb.Scene)


@needs_ref
@pytest.mark.skipif(not os.path.isfile(REF_CBOX), reason="reference checkout not present")
def test_reference_cbox_glb_live():
    """scenes/cbox.glb (the reference's own asset) through both loaders."""
    from oracle import ref
    S, R = pb.Scene(), ref.RefScene()
    assert gltf.scene_from_gltf(S, REF_CBOX) == R.load_gltf(REF_CBOX) == 8
    for T in (S, R):
        T.set_camera((64, 64), (0, 1, 4), (0, 1, 0), 0.4)
    boxes = np.array([np.concatenate([v.min(0), v.max(0)]) for v, _, _, _ in S.mesh_data], np.float32)
    assert np.array_equal(u32(boxes), u32(R.geometry_aabbs().reshape(-1, 6)))
    rays = R.primary_rays(0)
    hr, hp = R.trace_closest(rays), port.PortScene(S).trace_closest(rays)
    assert (hr["geom"] >= 0).mean() > 0.5
    assert np.array_equal(hr["geom"], hp["geom"]) and np.array_equal(hr["prim"], hp["prim"])
    assert np.array_equal(u32(hr["t"]), u32(hp["t"]))
