"""glTF 2.0 / .glb front end of the path: `scene_from_gltf` (src/pine/core/fileio.cpp:146-327) restated
over the builder interface of `pine_b200.Scene`, so `mesh (.glb)` scenes reach the CUDA integrator with
the geometry, materials and camera the reference would have built.

What the reference does, and this does identically (all arithmetic in binary32 through the C++ host
helpers, so vertices, normals and the camera are bit-identical -- tests/test_gltf.py):
  * walks `scenes[*].nodes` depth first; a node's matrix is `parent * transpose(mat4(matrix))`, then
    `* translate(T)`, `* q2m(w, x, y, z)`, `* scale(S)`, each applied when present (fileio.cpp:163-169);
  * every TRIANGLES primitive becomes one `Mesh` geometry: u16 / u32 indices (u8 fails the reference's
    CHECK, fileio.cpp:180), float POSITION / NORMAL / TEXCOORD_0, then `Mesh::apply(transform)`
    (geometry.cpp:652-658);
  * material: `UberMaterial(baseColorFactor.rgb, roughnessFactor, metallicFactor, transmission, ior)` with
    tinygltf's defaults (base colour 1, metallic 1, roughness 1), `KHR_materials_transmission`,
    `KHR_materials_ior` (default 1.45), and `EmissiveMaterial(emissiveFactor * emissiveStrength)` when
    `KHR_materials_emissive_strength` makes that non-zero (fileio.cpp:255-299). A primitive without a
    material gets Uber(1, 1, 0, 0, 1.45);
  * every node with a camera sets `ThinLenCamera(Film(int(640 * aspect), 640), T, T + R * (0,0,-1),
    yfov / 2)` from the node's OWN translation / rotation (fileio.cpp:308-321).
  * `baseColorTexture` becomes the albedo image node `NodeImage(NodeUV(), image)`, and
    `metallicRoughnessTexture` supplies roughness (component 1; unless the JSON also gives a
    roughnessFactor, which the reference applies afterwards) and metallic (component 2)
    (fileio.cpp:276-292): the 8-bit image goes to the device as a texture (pb2_texture) and is looked up
    there exactly as node.cpp:20-23 / image.cpp:7-18 do (nearest texel, wrap, pow(v / 255, 2.2)). As in
    the reference the factors are NOT multiplied in, texCoord sets other than 0, samplers and
    KHR_texture_transform are ignored. Images are decoded with Pillow: PNG is lossless, so texels equal
    the reference's stb_image decode; JPEG decoders may differ by one level in a few texels.
Not supported (error, never a silent fallback): 16-bit or float images, sparse accessors, non-triangle
modes.
"""
import base64
import json
import os
import struct

import numpy as np

from . import _abi as A
from .scene import _f, _fp, mat4

_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32,
              5126: np.float32, 5130: np.float64}
_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


class GltfError(ValueError):
    pass


def parse(data, base_dir=None):
    """-> (json dict, [buffer bytes]) of a .glb container or a .gltf JSON document."""
    if data[:4] == b"glTF":
        magic, version, length = struct.unpack_from("<4sII", data, 0)
        if version != 2:
            raise GltfError(f"glb version {version} is not supported")
        off, doc, bin_chunk = 12, None, None
        while off + 8 <= min(length, len(data)):
            clen, ctype = struct.unpack_from("<I4s", data, off)
            chunk = data[off + 8:off + 8 + clen]
            if ctype == b"JSON":
                doc = json.loads(chunk.decode("utf-8"))
            elif ctype == b"BIN\0" and bin_chunk is None:
                bin_chunk = chunk
            off += 8 + clen + (-clen % 4)
        if doc is None:
            raise GltfError("glb has no JSON chunk")
    else:
        doc, bin_chunk = json.loads(data.decode("utf-8")), None
    buffers = []
    for i, b in enumerate(doc.get("buffers", [])):
        uri = b.get("uri")
        if uri is None:
            if bin_chunk is None:
                raise GltfError(f"buffer {i} has no uri and the file has no BIN chunk")
            buffers.append(bin_chunk)
        elif uri.startswith("data:"):
            buffers.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            if base_dir is None:
                raise GltfError(f"buffer {i} refers to the external file `{uri}`")
            with open(_inside(base_dir, uri, f"buffer {i}"), "rb") as f:
                buffers.append(f.read())
    return doc, buffers


def _inside(base_dir, uri, what):
    """base_dir / uri, refusing uris that leave base_dir (absolute paths, `..`): the file is untrusted input."""
    from urllib.parse import unquote
    rel = unquote(uri)
    root = os.path.realpath(base_dir)
    path = os.path.realpath(os.path.join(root, rel))
    if os.path.isabs(rel) or (path != root and not path.startswith(root + os.sep)):
        raise GltfError(f"{what}: uri `{uri}` points outside the directory of the .gltf file")
    return path


def _accessor(doc, buffers, index):
    acc = doc["accessors"][index]
    if "sparse" in acc:
        raise GltfError("sparse accessors are not supported")
    try:
        view = doc["bufferViews"][acc["bufferView"]]
        dt, width = np.dtype(_COMPONENT[acc["componentType"]]), _WIDTH[acc["type"]]
        raw = np.frombuffer(buffers[view["buffer"]], np.uint8)
    except (KeyError, IndexError, TypeError) as e:
        raise GltfError(f"accessor {index}: bad bufferView / buffer / type reference ({e})")
    view_start, acc_off = int(view.get("byteOffset", 0)), int(acc.get("byteOffset", 0))
    start = view_start + acc_off
    stride = int(view.get("byteStride", 0)) or dt.itemsize * width
    count = int(acc["count"])
    # the file is untrusted: the last element must end inside the bufferView and the buffer
    elem = dt.itemsize * width
    view_len = int(view.get("byteLength", len(raw) - view_start))
    end = acc_off + (count - 1) * stride + elem if count > 0 else acc_off
    if (count < 0 or view_start < 0 or acc_off < 0 or stride < elem or view_start + view_len > len(raw)
            or end > view_len):
        raise GltfError(f"accessor {index}: {count} elements of {elem} bytes (stride {stride}) at offset {acc_off} "
                        f"do not fit its bufferView ({view_len} bytes) / buffer ({len(raw)} bytes)")
    if stride == dt.itemsize * width:
        out = raw[start:start + count * stride].view(dt).reshape(count, width)
    else:
        rows = np.lib.stride_tricks.as_strided(raw[start:], (count, dt.itemsize * width), (stride, 1))
        out = np.ascontiguousarray(rows).view(dt).reshape(count, width)
    return acc, out


def _col(m):  # math layout [4,4] -> the host helpers' column-major float[16]
    return _f(_f(m).T).reshape(-1)


def _op(op, args):
    return mat4._op(op, [np.float32(a) for a in args])


def _node_transform(node, transform):
    if len(node.get("matrix", [])) == 16:
        # transpose(mat4(m[0..15])): the constructor reads its arguments row by row (vecmath.h:579-581)
        m = _f(node["matrix"]).reshape(4, 4)
        transform = mat4.mul(transform, m.T.copy())
    if len(node.get("translation", [])) == 3:
        transform = mat4.mul(transform, _op(0, node["translation"]))
    if len(node.get("rotation", [])) == 4:
        x, y, z, w = node["rotation"]
        transform = mat4.mul(transform, _op(7, [w, x, y, z]))
    if len(node.get("scale", [])) == 3:
        transform = mat4.mul(transform, _op(1, node["scale"]))
    return transform


def _image(doc, buffers, index, base_dir):
    """-> uint8 [h, w, 4] of images[index] (tinygltf expands every image to 4 components)."""
    import io
    from PIL import Image
    img = doc["images"][index]
    if "bufferView" in img:
        view = doc["bufferViews"][img["bufferView"]]
        start = view.get("byteOffset", 0)
        if start < 0 or start + view["byteLength"] > len(buffers[view["buffer"]]):
            raise GltfError(f"image {index}: its bufferView does not fit the buffer")
        raw = bytes(buffers[view["buffer"]][start:start + view["byteLength"]])
    elif img.get("uri", "").startswith("data:"):
        raw = base64.b64decode(img["uri"].split(",", 1)[1])
    elif "uri" in img and base_dir is not None:
        with open(_inside(base_dir, img["uri"], f"image {index}"), "rb") as f:
            raw = f.read()
    else:
        raise GltfError(f"image {index} has no readable source")
    im = Image.open(io.BytesIO(raw))
    if im.mode in ("I;16", "I;16B", "I", "F"):
        raise GltfError(f"image {index}: only 8-bit images are supported (fileio.cpp:152)")
    return np.ascontiguousarray(np.asarray(im.convert("RGBA"), np.uint8))


def _material(doc, index, texture_of=None):
    """-> ("uber" | "emissive", kwargs for add_material). `texture_of(texture index)` -> scene texture."""
    color, rough, metal, trans, ior = (1.0, 1.0, 1.0), 1.0, 0.0, 0.0, 1.45
    emission_color, emission_strength = (1.0, 1.0, 1.0), 0.0
    if index is not None and index >= 0:
        mat = doc["materials"][index]
        ext = mat.get("extensions", {})
        if "KHR_materials_transmission" in ext:
            trans = ext["KHR_materials_transmission"].get("transmissionFactor", 0.0)
        if "KHR_materials_ior" in ext:
            ior = ext["KHR_materials_ior"].get("ior", 0.0)
        if "KHR_materials_emissive_strength" in ext:
            emission_strength = ext["KHR_materials_emissive_strength"].get("emissiveStrength", 0.0)
        pbr = mat.get("pbrMetallicRoughness", {})
        textures = {}
        for key in ("baseColorTexture", "metallicRoughnessTexture"):
            if key in pbr:
                if texture_of is None:
                    raise GltfError(f"material `{mat.get('name', index)}` uses {key} but the scene builder "
                                    "has no add_texture")
                slot = texture_of(pbr[key]["index"])
                if key == "baseColorTexture":
                    textures["color_texture"] = slot
                else:
                    # the reference applies the parsed keys in std::map order (fileio.cpp:276-292):
                    # metallicFactor < metallicRoughnessTexture < roughnessFactor, so an explicit
                    # roughnessFactor overrides the texture's roughness while metallic stays textured
                    textures["metallic_texture"] = slot
                    if "roughnessFactor" not in pbr:
                        textures["roughness_texture"] = slot
        color = tuple(pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0])[:3])
        metal = pbr.get("metallicFactor", 1.0)
        rough = pbr.get("roughnessFactor", 1.0)
        emission_color = tuple(mat.get("emissiveFactor", [0.0, 0.0, 0.0]))
    f = np.float32
    emission = tuple(f(c) * f(emission_strength) for c in emission_color)
    if any(e != 0 for e in emission):
        return "emissive", dict(color=emission)
    kw = dict(color=tuple(f(c) for c in color), roughness=float(f(rough)), metallic=float(f(metal)),
              transmission=float(f(trans)), ior=float(f(ior)))
    if index is not None and index >= 0:
        kw.update(textures)
    return "uber", kw


def scene_from_gltf(scene, source, transform=None):
    """Adds the geometries (and camera, if any) of a .glb / .gltf to `scene`; returns how many geometries
    were added. `source`: a file name or the file's bytes. `transform`: global mat4 (math layout)."""
    base_dir = None
    if isinstance(source, (str, os.PathLike)):
        base_dir = os.path.dirname(os.path.abspath(source))
        with open(source, "rb") as f:
            source = f.read()
    doc, buffers = parse(bytes(source), base_dir)
    transform = np.eye(4, dtype=np.float32) if transform is None else _f(transform).reshape(4, 4)
    added = 0
    nodes = doc.get("nodes", [])
    image_slots = {}

    def texture_of(tex_index):  # textures[i].source -> one scene texture per image
        texs = doc.get("textures", [])
        src = texs[tex_index].get("source") if 0 <= tex_index < len(texs) else None
        if src is None or not 0 <= src < len(doc.get("images", [])):
            raise GltfError(f"texture {tex_index} has no source image")
        if src not in image_slots:
            image_slots[src] = scene.add_texture(_image(doc, buffers, src, base_dir))
        return image_slots[src]

    def process(node, tr):
        nonlocal added
        tr = _node_transform(node, tr)
        if node.get("mesh", -1) >= 0:
            for prim in doc["meshes"][node["mesh"]]["primitives"]:
                if prim.get("mode", 4) != 4:
                    raise GltfError("only TRIANGLES primitives are supported (fileio.cpp:181)")
                if "indices" not in prim:
                    raise GltfError("non-indexed primitives are not supported")
                acc, idx = _accessor(doc, buffers, prim["indices"])
                if acc["componentType"] not in (5123, 5125):
                    raise GltfError("index accessors must be UNSIGNED_SHORT or UNSIGNED_INT (fileio.cpp:180)")
                idx = idx.reshape(-1)
                idx = idx[:len(idx) // 3 * 3].astype(np.uint32).reshape(-1, 3)
                verts = normals = uvs = None
                for name, a in prim.get("attributes", {}).items():
                    if name not in ("POSITION", "NORMAL", "TEXCOORD_0"):
                        continue
                    acc, arr = _accessor(doc, buffers, a)
                    want = "VEC2" if name == "TEXCOORD_0" else "VEC3"
                    if acc["type"] != want or acc["componentType"] not in (5126, 5130):
                        raise GltfError(f"{name} must be float {want}")
                    arr = np.ascontiguousarray(arr, np.float32)
                    if name == "POSITION":
                        verts = arr.copy()
                    elif name == "NORMAL":
                        normals = arr.copy()
                    else:
                        uvs = arr.copy()
                if verts is None:
                    raise GltfError("primitive without POSITION")
                for what, arr in (("NORMAL", normals), ("TEXCOORD_0", uvs)):
                    if arr is not None and len(arr) != len(verts):
                        raise GltfError(f"{what} has {len(arr)} elements but POSITION has {len(verts)}")
                if len(idx) and int(idx.max()) >= len(verts):
                    raise GltfError("an index refers to a vertex beyond POSITION's count")
                n_norm = 0 if normals is None else len(normals)
                A.lib().pb2h_mesh_apply(_fp(_col(tr)), _fp(verts), len(verts),
                                        _fp(normals) if n_norm else None, n_norm)
                kind, kw = _material(doc, prim.get("material", -1),
                                     texture_of if hasattr(scene, "add_texture") else None)
                mat = scene.add_material(kind, **kw)
                scene.add_mesh(verts, idx, mat, normals=normals, uvs=uvs)
                added += 1
        for child in node.get("children", []):
            process(nodes[child], tr)

    for sc in doc.get("scenes", []):
        for n in sc.get("nodes", []):
            process(nodes[n], transform)

    for node in nodes:
        if node.get("camera", -1) >= 0:
            cam = doc["cameras"][node["camera"]]
            T, R = node.get("translation", []), node.get("rotation", [])
            if len(T) != 3 or len(R) != 4:
                raise GltfError("camera nodes need a translation and a rotation (fileio.cpp:312-313)")
            persp = cam.get("perspective", {})
            f = np.float32
            pos = np.array([f(v) for v in T], np.float32)
            rot = _op(7, [R[3], R[0], R[1], R[2]])[:3, :3]
            # rot * vec3(0, 0, -1) = x*0 + y*0 + z*(-1) summed left to right (vecmath.h:695-697)
            fwd = (rot[:, 0] * f(0) + rot[:, 1] * f(0)) + rot[:, 2] * f(-1)
            at = pos + fwd.astype(np.float32)
            width = int(640 * float(persp.get("aspectRatio", 0.0)))
            scene.set_camera((width, 640), tuple(pos), tuple(at), float(f(float(persp.get("yfov", 0.0)) / 2)))
    return added


def load_scene(source, transform=None):
    """`load_scene(filename)` (fileio.h:76): a new Scene holding the file's geometries and camera."""
    from .scene import Scene
    s = Scene()
    scene_from_gltf(s, source, transform)
    return s
