"""Writes the synthetic .glb the glTF-loader tests run on (test infrastructure; `python tests/gltf_fixture.py`
regenerates tests/golden/room_like.glb). The file exercises everything scene_from_gltf (fileio.cpp:146-327)
reads: a node hierarchy with a matrix node, TRS nodes and children; u16 and u32 indices; primitives with and
without NORMAL / TEXCOORD_0; interleaved (byteStride) vertex data; Uber materials with metallic / roughness /
KHR_materials_transmission / KHR_materials_ior, a primitive without a material, an emitter through
KHR_materials_emissive_strength; and a perspective camera node.
"""
import json
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class GlbWriter:
    def __init__(self):
        self.bin = bytearray()
        self.doc = {"asset": {"version": "2.0", "generator": "pine_b200 tests"}, "scene": 0,
                    "scenes": [{"nodes": []}], "nodes": [], "meshes": [], "materials": [], "accessors": [],
                    "bufferViews": [], "buffers": [{}], "cameras": []}

    def view(self, raw, stride=None):
        while len(self.bin) % 4:
            self.bin.append(0)
        v = {"buffer": 0, "byteOffset": len(self.bin), "byteLength": len(raw)}
        if stride:
            v["byteStride"] = stride
        self.bin += raw
        self.doc["bufferViews"].append(v)
        return len(self.doc["bufferViews"]) - 1

    def accessor(self, arr, kind, view=None, offset=0):
        arr = np.ascontiguousarray(arr)
        ct = {np.dtype(np.uint16): 5123, np.dtype(np.uint32): 5125, np.dtype(np.float32): 5126}[arr.dtype]
        if view is None:
            view = self.view(arr.tobytes())
        a = {"bufferView": view, "componentType": ct, "count": int(arr.shape[0]) if arr.ndim > 1 else int(arr.size),
             "type": kind}
        if offset:
            a["byteOffset"] = offset
        if kind == "VEC3" and ct == 5126:
            a["min"], a["max"] = arr.min(0).tolist(), arr.max(0).tolist()
        self.doc["accessors"].append(a)
        return len(self.doc["accessors"]) - 1

    def primitive(self, verts, idx, material=None, normals=None, uvs=None, index_type=np.uint16, interleave=False):
        verts = np.asarray(verts, np.float32)
        attrs = {}
        if interleave and normals is not None:
            both = np.concatenate([verts, np.asarray(normals, np.float32)], axis=1)
            v = self.view(both.tobytes(), stride=24)
            attrs["POSITION"] = self.accessor(verts, "VEC3", view=v)
            attrs["NORMAL"] = self.accessor(np.asarray(normals, np.float32), "VEC3", view=v, offset=12)
        else:
            attrs["POSITION"] = self.accessor(verts, "VEC3")
            if normals is not None:
                attrs["NORMAL"] = self.accessor(np.asarray(normals, np.float32), "VEC3")
        if uvs is not None:
            attrs["TEXCOORD_0"] = self.accessor(np.asarray(uvs, np.float32), "VEC2")
        p = {"attributes": attrs, "indices": self.accessor(np.asarray(idx).reshape(-1).astype(index_type), "SCALAR")}
        if material is not None:
            p["material"] = material
        return p

    def mesh(self, primitives):
        self.doc["meshes"].append({"primitives": primitives})
        return len(self.doc["meshes"]) - 1

    def material(self, **kw):
        self.doc["materials"].append(kw)
        return len(self.doc["materials"]) - 1

    def node(self, root=False, **kw):
        self.doc["nodes"].append(kw)
        n = len(self.doc["nodes"]) - 1
        if root:
            self.doc["scenes"][0]["nodes"].append(n)
        return n

    def tobytes(self):
        doc = dict(self.doc)
        if not doc["cameras"]:
            del doc["cameras"]
        doc["buffers"] = [{"byteLength": len(self.bin)}]
        js = json.dumps(doc, separators=(",", ":")).encode()
        js += b" " * (-len(js) % 4)
        binc = bytes(self.bin) + b"\0" * (-len(self.bin) % 4)
        total = 12 + 8 + len(js) + 8 + len(binc)
        return (struct.pack("<4sII", b"glTF", 2, total) + struct.pack("<I4s", len(js), b"JSON") + js +
                struct.pack("<I4s", len(binc), b"BIN\0") + binc)


def png(pixels):
    """uint8 [h, w, 3|4] -> PNG bytes (8-bit, filter 0 / Sub alternating so a decoder's unfilter is exercised)."""
    import zlib
    a = np.ascontiguousarray(pixels, np.uint8)
    h, w, c = a.shape
    rows = bytearray()
    for y in range(h):
        row = a[y].reshape(-1)
        if y % 2:
            prev = np.concatenate([np.zeros(c, np.uint8), row[:-c]])
            rows += b"\x01" + ((row.astype(np.int16) - prev) & 255).astype(np.uint8).tobytes()
        else:
            rows += b"\x00" + row.tobytes()

    def chunk(kind, data):
        body = kind + data
        return struct.pack(">I", len(data)) + body + struct.pack(">I", zlib.crc32(body) & 0xffffffff)
    ihdr = struct.pack(">IIBBBBB", w, h, 8, 2 if c == 3 else 6, 0, 0, 0)
    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + chunk(b"IDAT", zlib.compress(bytes(rows), 9)) + chunk(b"IEND", b"")


def box(lo, hi):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    verts, normals, idx = [], [], []
    for axis in range(3):
        for side in (0, 1):
            u, v = (axis + 1) % 3, (axis + 2) % 3
            n = np.zeros(3, np.float32)
            n[axis] = 1 if side else -1
            base = len(verts)
            for a, b in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = np.zeros(3, np.float32)
                p[axis] = hi[axis] if side else lo[axis]
                p[u] = hi[u] if a else lo[u]
                p[v] = hi[v] if b else lo[v]
                verts.append(p)
                normals.append(n)
            idx += [[base, base + 1, base + 2], [base, base + 2, base + 3]]
    return np.array(verts, np.float32), np.array(idx, np.uint32), np.array(normals, np.float32)


def room_like():
    from pine_b200 import scenes
    w = GlbWriter()
    white = w.material(name="white", pbrMetallicRoughness={"baseColorFactor": [0.73, 0.71, 0.68, 1], "metallicFactor": 0,
                                                           "roughnessFactor": 0.9})
    red = w.material(name="red", pbrMetallicRoughness={"baseColorFactor": [0.63, 0.065, 0.05, 1], "metallicFactor": 0.0,
                                                       "roughnessFactor": 0.75},
                     extensions={"KHR_materials_ior": {"ior": 1.0}})
    brass = w.material(name="brass", pbrMetallicRoughness={"baseColorFactor": [0.9, 0.7, 0.3, 1], "metallicFactor": 0.85,
                                                           "roughnessFactor": 0.3})
    glass = w.material(name="glass", pbrMetallicRoughness={"baseColorFactor": [0.95, 0.97, 1.0, 1], "metallicFactor": 0,
                                                           "roughnessFactor": 0.05},
                       extensions={"KHR_materials_transmission": {"transmissionFactor": 0.9},
                                   "KHR_materials_ior": {"ior": 1.5}})
    default_pbr = w.material(name="defaults")  # tinygltf defaults: base colour 1, metallic 1, roughness 1
    lamp = w.material(name="lamp", emissiveFactor=[1.0, 0.85, 0.6],
                      extensions={"KHR_materials_emissive_strength": {"emissiveStrength": 30.0}})
    not_a_lamp = w.material(name="emissive without strength", emissiveFactor=[1.0, 1.0, 1.0],
                            pbrMetallicRoughness={"baseColorFactor": [0.2, 0.3, 0.8, 1], "metallicFactor": 0,
                                                  "roughnessFactor": 0.6})

    # room shell: inward-facing box, one primitive per material, u16 indices, interleaved vertex data
    v, i, n = box((-2, 0, -2), (2, 2.5, 2))
    shell = [w.primitive(v[:16], i[:8] % 16, white, normals=-n[:16]),
             w.primitive(v[16:], i[8:] - 16, red, normals=-n[16:])]
    sv, si = scenes.icosphere(2)
    sv, si = np.asarray(sv, np.float32), np.asarray(si, np.uint32)
    uv = np.stack([np.arctan2(sv[:, 2], sv[:, 0]) / (2 * np.pi) + 0.5, sv[:, 1] * 0.5 + 0.5], 1).astype(np.float32)
    ball = [w.primitive(sv, si, brass, normals=sv / np.linalg.norm(sv, axis=1, keepdims=True), uvs=uv,
                        index_type=np.uint32)]
    bv, bi, bn = box((-0.5, -0.5, -0.5), (0.5, 0.5, 0.5))
    cube = [w.primitive(bv, bi, glass, normals=bn)]
    cube_plain = [w.primitive(bv, bi, None)]           # no material, no normals
    cube_default = [w.primitive(bv, bi, default_pbr, index_type=np.uint32)]
    cube_blue = [w.primitive(bv, bi, not_a_lamp, normals=bn)]
    quad = np.array([[-0.5, 0, -0.5], [0.5, 0, -0.5], [0.5, 0, 0.5], [-0.5, 0, 0.5]], np.float32)
    light = [w.primitive(quad, [[0, 1, 2], [0, 2, 3]], lamp)]
    m_shell, m_ball, m_cube, m_plain, m_default, m_blue, m_light = (
        w.mesh(p) for p in (shell, ball, cube, cube_plain, cube_default, cube_blue, light))

    w.node(root=True, name="shell", mesh=m_shell)
    # matrix node (column-major: rotate 30 degrees about y, non-uniform scale, translate) with two children
    c, s = float(np.cos(np.pi / 6)), float(np.sin(np.pi / 6))
    child_a = w.node(name="ball", mesh=m_ball, translation=[0.3, 0.45, -0.2], scale=[0.45, 0.45, 0.45])
    grandchild = w.node(name="cube on ball", mesh=m_blue, translation=[0.0, 1.3, 0.0],
                        rotation=[0.0, 0.3826834323650898, 0.0, 0.9238795325112867], scale=[0.5, 0.25, 0.5])
    w.doc["nodes"][child_a]["children"] = [grandchild]
    child_b = w.node(name="glass cube", mesh=m_cube, translation=[-0.9, 0.5, 0.4],
                     rotation=[0.1830127, 0.1830127, 0.6830127, 0.6830127])
    w.node(root=True, name="group", children=[child_a, child_b],
           matrix=[1.2 * c, 0, -1.2 * s, 0, 0, 1.0, 0, 0, 0.8 * s, 0, 0.8 * c, 0, 0.4, 0.0, 0.3, 1])
    w.node(root=True, name="plain", mesh=m_plain, translation=[1.2, 0.3, -1.0], scale=[0.6, 0.6, 0.6])
    w.node(root=True, name="defaults", mesh=m_default, translation=[-1.3, 0.35, -1.1],
           rotation=[0.0, 0.2588190451, 0.0, 0.9659258263], scale=[0.7, 0.7, 0.7])
    w.node(root=True, name="lamp", mesh=m_light, translation=[0.0, 2.45, 0.0], scale=[1.6, 1.0, 1.6])
    w.doc["cameras"].append({"type": "perspective",
                             "perspective": {"aspectRatio": 0.15, "yfov": 0.9, "znear": 0.1}})
    w.node(root=True, name="camera", camera=0, translation=[0.1, 1.2, 1.9],
           rotation=[-0.0499791693, 0.0, 0.0, 0.9987502604])
    return w.tobytes()


def room_textured():
    """Image-node materials (fileio.cpp:285-292): a floor with a base-colour PNG (RGB, 37 x 23), a wall
    whose metallic/roughness come from an RGBA PNG (metallic texels are 0 or 255 only, so the Uber lobe
    choice stays deterministic: with_probability(0 | 1), sampler.h:317-324), a ball that uses both images
    plus an explicit roughnessFactor (which the reference lets override the texture), and a mesh without
    TEXCOORD_0 (uv falls back to the triangle's own parametrisation). No camera node: the test sets one."""
    from pine_b200 import scenes
    rng = np.random.RandomState(7)
    w = GlbWriter()
    w.doc["images"], w.doc["textures"], w.doc["samplers"] = [], [], [{}]

    def image(pixels):
        raw = png(pixels)
        w.doc["images"].append({"bufferView": w.view(raw), "mimeType": "image/png"})
        w.doc["textures"].append({"source": len(w.doc["images"]) - 1, "sampler": 0})
        return len(w.doc["textures"]) - 1
    color = rng.randint(30, 256, size=(23, 37, 3)).astype(np.uint8)
    mr = rng.randint(0, 256, size=(16, 16, 4)).astype(np.uint8)
    mr[..., 2] = np.where(rng.rand(16, 16) < 0.4, 255, 0)
    t_color, t_mr = image(color), image(mr)
    floor = w.material(name="floor", pbrMetallicRoughness={"baseColorTexture": {"index": t_color},
                                                           "baseColorFactor": [0.1, 0.1, 0.1, 1],
                                                           "metallicFactor": 0, "roughnessFactor": 0.8})
    wall = w.material(name="wall", pbrMetallicRoughness={"baseColorFactor": [0.8, 0.75, 0.7, 1], "metallicFactor": 0.5,
                                                         "metallicRoughnessTexture": {"index": t_mr}})
    ball = w.material(name="ball", pbrMetallicRoughness={"baseColorTexture": {"index": t_color},
                                                         "metallicRoughnessTexture": {"index": t_mr},
                                                         "roughnessFactor": 0.35})
    plain = w.material(name="no uvs", pbrMetallicRoughness={"baseColorTexture": {"index": t_color},
                                                            "metallicFactor": 0, "roughnessFactor": 0.9})
    lamp = w.material(name="lamp", emissiveFactor=[1.0, 0.9, 0.8],
                      extensions={"KHR_materials_emissive_strength": {"emissiveStrength": 40.0}})
    quad = np.array([[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1]], np.float32)
    quad_uv = np.array([[0, 0], [3, 0], [3, 2], [0, 2]], np.float32)  # repeats: exercises fract()
    qi = [[0, 2, 1], [0, 3, 2]]
    up = np.tile(np.float32([0, 1, 0]), (4, 1))
    w.node(root=True, name="floor", mesh=w.mesh([w.primitive(quad, qi, floor, normals=up, uvs=quad_uv)]),
           scale=[2.0, 1.0, 2.0])
    w.node(root=True, name="wall", mesh=w.mesh([w.primitive(quad, qi, wall, normals=up, uvs=quad_uv * 0.5 - 0.75)]),
           translation=[0.0, 1.5, -2.0], rotation=[0.7071067811865476, 0.0, 0.0, 0.7071067811865476],
           scale=[2.0, 1.0, 1.5])
    sv, si = scenes.icosphere(3)
    sv, si = np.asarray(sv, np.float32), np.asarray(si, np.uint32)
    uv = np.stack([np.arctan2(sv[:, 2], sv[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(sv[:, 1], -1, 1)) / np.pi], 1)
    w.node(root=True, name="ball", mesh=w.mesh([w.primitive(sv, si, ball, normals=sv, uvs=uv.astype(np.float32),
                                                            index_type=np.uint32)]),
           translation=[-0.7, 0.6, -0.3], scale=[0.6, 0.6, 0.6])
    bv, bi, bn = box((-0.4, 0.0, -0.4), (0.4, 0.8, 0.4))
    w.node(root=True, name="box without uvs", mesh=w.mesh([w.primitive(bv, bi, plain, normals=bn)]),
           translation=[0.9, 0.0, 0.4], rotation=[0.0, 0.25881904510252074, 0.0, 0.9659258262890683])
    down = np.array([[-0.5, 0, -0.5], [0.5, 0, -0.5], [0.5, 0, 0.5], [-0.5, 0, 0.5]], np.float32)
    w.node(root=True, name="lamp", mesh=w.mesh([w.primitive(down, [[0, 1, 2], [0, 2, 3]], lamp)]),
           translation=[0.0, 2.6, 0.0])
    return w.tobytes()


if __name__ == "__main__":
    for name, make in (("room_like", room_like), ("room_textured", room_textured)):
        out = os.path.join(ROOT, "tests", "golden", name + ".glb")
        with open(out, "wb") as f:
            f.write(make())
        print(out, os.path.getsize(out), "bytes")
