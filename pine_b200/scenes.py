"""The benchmark scenes of BASELINE.json as builder-interface descriptions.

Each function takes any object with the builder methods of `pine_b200.Scene` /
`oracle.ref.RefScene` and feeds it plain numbers, so the reference and the CUDA backend see the
same scene. Matrices come from `pine_b200.mat4` (the C++ host mirror of vecmath.h).
"""
import numpy as np

from .scene import mat4


def _frange(a, step, b):
    """PRL `for x in a ~ step ~ b`: x = a; x <= b; x += step in f32 (jit.cpp:1534-1541)."""
    x, step, b = np.float32(a), np.float32(step), np.float32(b)
    out = []
    while x <= b:
        out.append(float(x))
        x = np.float32(x + step)
    return out


def _trs(t, ry, s):
    return mat4.mul(mat4.mul(mat4.translate(t), mat4.rotate_y(ry)), mat4.scale(s))


def cbox_readme(B, size=(640, 640), cones=True):
    """C1: the README scene (reference README.md:7-35): 5 walls, cone field, two boxes, one light.
    PathIntegrator(BlueSampler(16), 4). The PRL float loop yields 20 x 20 cones (x stops at
    0.9000002 because -1 + 20*0.1f accumulates to 1.0000001 > 1)."""
    floor = B.add_material("diffuse", (0.9, 0.9, 0.9))
    red = B.add_material("diffuse", (0.9, 0.1, 0.05))
    green = B.add_material("diffuse", (0.2, 0.9, 0.05))
    metal = B.add_material("metal", (1, 1, 1), roughness=0.0)
    B.add_material("glossy", (0.25, 0.25, 0.25), roughness=0.05)          # declared, unassigned
    B.add_material("subsurface", (1, 1, 1), roughness=0.0, sigma_s=(40.0,) * 3)  # declared, unassigned
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((0, 2, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((-1, 1, 1), (0, 0, 2), (0, 2, 0), red)
    B.add_rect((1, 1, 1), (0, 0, 2), (0, 2, 0), green)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), floor)
    if cones:
        for x in _frange(-1.0, 0.1, 1.0):
            for z in _frange(0.0, 0.1, 2.0):
                B.add_cone((x, 0, z), (0, 1, 0), 0.05, 0.05, floor)
    B.add_box((0, 0, 0), (1, 1, 1), _trs((0.0, 0.0, 0.6), 0.4, (0.6, 0.6, 0.6)), floor)
    B.add_box((0, 0, 0), (1, 1, 1), _trs((-0.6, 0.0, 1.0), -0.4, (0.6, 1.3, 0.6)), metal)
    light = B.add_material("emissive", tuple(np.float32(600) * np.float32(c) for c in (1.0, 0.64, 0.185)))
    B.add_rect((0.0, 1.9, 0.5), (0.1, 0, 0), (0, 0, 0.1), light)
    B.set_camera(size, (0, 1, -4), (0, 1, 0), 0.25)
    return {"spp": 16, "depth": 4}


def cbox_file(B, size=(1024, 1024)):
    """C1b: scenes/cbox.pine as written on disk (1024x1024, BlueSampler(256), depth 8)."""
    floor = B.add_material("diffuse", (0.9, 0.9, 0.9))
    blue = B.add_material("diffuse", (0.2, 0.5, 0.9))
    red = B.add_material("diffuse", (0.9, 0.1, 0.05))
    green = B.add_material("diffuse", (0.2, 0.9, 0.05))
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), floor, flip=True)
    B.add_rect((0, 2, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((-1, 1, 1), (0, 0, 2), (0, 2, 0), red, flip=True)
    B.add_rect((1, 1, 1), (0, 0, 2), (0, 2, 0), green)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), blue, flip=True)
    B.add_box((0, 0, 0), (1, 1, 1), _trs((0.0, 0.0, 0.6), 0.4, (0.6, 0.6, 0.6)), floor)
    B.add_box((0, 0, 0), (1, 1, 1), _trs((-0.6, 0.0, 1.0), -0.4, (0.6, 1.3, 0.6)), floor)
    light = B.add_material("emissive", tuple(np.float32(600) * np.float32(c) for c in (1.0, 0.64, 0.185)))
    B.add_rect((0.0, 1.9, 1), (0.1, 0, 0), (0, 0, 0.1), light)
    B.set_camera(size, (0, 0, 0), (0, 0, 1), 0.4)
    return {"spp": 256, "depth": 8}


def cbox_indirect(B, size=(1024, 1024)):
    """C3: scenes/cbox-indirect.pine geometry; BASELINE asks 1024x1024, 256 spp, depth 8
    (the file itself says 720x360, 64 spp, depth 5). Film(image_size) default tonemapper."""
    floor = B.add_material("diffuse", (0.95, 0.95, 0.95))
    red = B.add_material("diffuse", (0.9, 0.1, 0.05))
    green = B.add_material("diffuse", (0.2, 0.9, 0.05))
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((0, 2, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), floor)
    B.add_rect((-1, 1, 1), (0, 0, 2), (0, 2, 0), red)
    B.add_rect((1, 1, 1), (0, 0, 2), (0, 2, 0), green)
    B.add_rect((0.2, 1.5, 1), (1.6, 0, 0), (0, 0, 2), floor)
    light = B.add_material("emissive", tuple(np.float32(600) * np.float32(c) for c in (1.0, 0.64, 0.185)))
    B.add_rect((0, 1.9, 1), (0.2, 0, 0), (0, 0, 0.2), light)
    B.set_camera(size, (0, 1, -4), (0, 1, 0), 0.25)
    return {"spp": 256, "depth": 8}


# ---- procedural meshes ------------------------------------------------------------------------------
def icosphere(subdiv):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t),
         (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4),
         (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9),
         (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return np.array(v, np.float32), np.array(f, np.uint32)


def mesh_field(B, n_instances=64, subdiv=3, seed=42, size=(640, 360), spp=16, depth=6,
               with_shapes=True):
    """Deterministic flattened-instancing scene (the C4/C5 family): `n_instances`
    icospheres (20 * 4^subdiv triangles each) on a jittered grid as separate Mesh geometries
    (Pine's own BVH ignores Scene::instancings, bvh.cpp:453-495), ground Rect, Rect emitters."""
    rng = np.random.RandomState(seed)
    bv, bf = icosphere(subdiv)
    mats = [B.add_material("diffuse", (0.8, 0.8, 0.8)), B.add_material("diffuse", (0.8, 0.3, 0.2)),
            B.add_material("metal", (0.9, 0.9, 0.9), roughness=0.2),
            B.add_material("glossy", (0.3, 0.5, 0.8), roughness=0.1)]
    g = int(np.ceil(n_instances ** (1 / 3)))
    cell = 1.0
    k = 0
    for ix in range(g):
        for iy in range(g):
            for iz in range(g):
                if k >= n_instances:
                    break
                c = (np.array([ix, iy, iz], np.float32) + 0.5 + (rng.rand(3).astype(np.float32) - 0.5) * 0.5) * cell
                r = np.float32(0.25 + 0.15 * rng.rand())
                B.add_mesh((bv * r + c).astype(np.float32), bf, mats[k % len(mats)])
                k += 1
    ext = g * cell
    B.add_rect((ext / 2, -0.01, ext / 2), (4 * ext, 0, 0), (0, 0, 4 * ext), mats[0], flip=True)
    light = B.add_material("emissive", (40.0, 36.0, 30.0))
    for lx, lz in ((0.25, 0.25), (0.75, 0.25), (0.25, 0.75), (0.75, 0.75)):
        B.add_rect((ext * lx, ext + 1.0, ext * lz), (ext * 0.3, 0, 0), (0, 0, ext * 0.3), light)
    if with_shapes:
        B.add_sphere((-0.6, 0.4, ext / 2), 0.4, mats[2])
        B.add_disk((ext + 0.5, 0.3, ext / 2), (-1, 0.2, 0), 0.5, mats[1])
    B.set_camera(size, (ext / 2, ext * 0.9, -ext * 1.6), (ext / 2, ext * 0.4, ext / 2), 0.35)
    return {"spp": spp, "depth": depth}


def tessellated_box(n):
    """Unit cube [0,1]^3 with every face cut into n x n quads (2 triangles each): 12 n^2 triangles,
    vertices shared inside a face only -> (verts f32 [6 (n+1)^2, 3], faces u32 [12 n^2, 3])."""
    g = np.linspace(0.0, 1.0, n + 1, dtype=np.float32)
    u, v = np.meshgrid(g, g, indexing="ij")
    u, v = u.reshape(-1), v.reshape(-1)
    zero, one = np.zeros_like(u), np.ones_like(u)
    faces = [(zero, u, v, True), (one, u, v, False), (u, zero, v, False), (u, one, v, True),
             (u, v, zero, True), (u, v, one, False)]
    verts, tris = [], []
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    a = (i * (n + 1) + j).reshape(-1)
    b, c, d = a + (n + 1), a + (n + 2), a + 1
    for k, (x, y, z, flip) in enumerate(faces):
        verts.append(np.stack([x, y, z], axis=1))
        base = k * (n + 1) ** 2
        q = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)]) + base
        tris.append(q[:, ::-1] if flip else q)
    return np.concatenate(verts).astype(np.float32), np.concatenate(tris).astype(np.uint32)


def room_proc(B, size=(1920, 1080), spp=64, depth=6, n_spheres=60, n_boxes=64, sphere_subdiv=5, box_tess=32,
              seed=1234):
    """C4 "room-proc" (stands in for the missing scenes/room.glb): a CLOSED 8 x 3 x 6 m box room
    (six Rects), furniture = `n_spheres` icospheres (subdivision 5, 20 480 triangles each) and `n_boxes`
    tessellated boxes (12 * 32^2 = 12 288 triangles each) ~ 2.0 M triangles on a jittered floor grid, some
    raised on "table" slabs, Diffuse / Metal / Glossy mix, ONE 1 x 1 m Rect emitter under the ceiling; camera
    inside the room. Every path stays inside, so a sample traces close to the 2 D - 1 = 11 rays the depth
    allows: the indirect-heavy, memory-bound interior regime. Deterministic in `seed` (numpy RandomState; the
    same arrays feed the reference and the CUDA path)."""
    rng = np.random.RandomState(seed)
    W, H, D = 8.0, 3.0, 6.0
    wall = B.add_material("diffuse", (0.82, 0.8, 0.76))
    floor = B.add_material("glossy", (0.45, 0.32, 0.2), roughness=0.25)
    accent = B.add_material("diffuse", (0.7, 0.25, 0.2))
    mats = [B.add_material("diffuse", (0.75, 0.75, 0.7)), B.add_material("diffuse", (0.25, 0.45, 0.7)),
            B.add_material("metal", (0.9, 0.85, 0.7), roughness=0.2), B.add_material("glossy", (0.3, 0.6, 0.4), roughness=0.1),
            B.add_material("metal", (0.95, 0.95, 0.95), roughness=0.05), B.add_material("glossy", (0.8, 0.75, 0.3), roughness=0.3)]
    B.add_rect((W / 2, 0, D / 2), (W, 0, 0), (0, 0, D), floor)
    B.add_rect((W / 2, H, D / 2), (W, 0, 0), (0, 0, D), wall)
    B.add_rect((0, H / 2, D / 2), (0, 0, D), (0, H, 0), accent)
    B.add_rect((W, H / 2, D / 2), (0, 0, D), (0, H, 0), wall)
    B.add_rect((W / 2, H / 2, 0), (W, 0, 0), (0, H, 0), wall)
    B.add_rect((W / 2, H / 2, D), (W, 0, 0), (0, H, 0), wall)
    sv, sf = icosphere(sphere_subdiv)
    bv, bf = tessellated_box(box_tess)
    n_obj = n_spheres + n_boxes
    gx = int(np.ceil(np.sqrt(n_obj * W / D)))
    gz = int(np.ceil(n_obj / gx))
    cells = [(ix, iz) for iz in range(gz) for ix in range(gx)]
    order = rng.permutation(len(cells))[:n_obj]
    kinds = np.array([0] * n_spheres + [1] * n_boxes)
    rng.shuffle(kinds)
    cw, cd = (W - 0.6) / gx, (D - 1.4) / gz
    for k, ci in enumerate(order):
        ix, iz = cells[ci]
        cx = np.float32(0.3 + (ix + 0.5 + (rng.rand() - 0.5) * 0.4) * cw)
        cz = np.float32(1.1 + (iz + 0.5 + (rng.rand() - 0.5) * 0.4) * cd)
        lift = np.float32(0.0 if rng.rand() < 0.6 else 0.5 + 1.6 * rng.rand())  # some objects float / sit high
        mat = mats[int(rng.randint(len(mats)))]
        if kinds[k] == 0:
            r = np.float32(0.12 + 0.12 * rng.rand())
            c = np.array([cx, lift + r, cz], np.float32)
            B.add_mesh((sv * r + c).astype(np.float32), sf, mat)
        else:
            sx, sy, sz = (np.float32(0.18 + 0.22 * rng.rand()), np.float32(0.08 + 0.5 * rng.rand()),
                          np.float32(0.18 + 0.22 * rng.rand()))
            ang = np.float32(rng.rand() * np.pi)
            ca, sa = np.float32(np.cos(ang)), np.float32(np.sin(ang))
            p = (bv - np.array([0.5, 0.0, 0.5], np.float32)) * np.array([sx, sy, sz], np.float32)
            q = np.stack([ca * p[:, 0] + sa * p[:, 2], p[:, 1], -sa * p[:, 0] + ca * p[:, 2]], axis=1)
            B.add_mesh((q + np.array([cx, lift, cz], np.float32)).astype(np.float32), bf, mat)
    light = B.add_material("emissive", (90.0, 82.0, 70.0))
    B.add_rect((W / 2, H - 0.01, D / 2 + 0.5), (1.0, 0, 0), (0, 0, 1.0), light)
    B.set_camera(size, (W / 2, 1.6, 0.25), (W / 2, 1.0, D * 0.7), 0.55)
    return {"spp": spp, "depth": depth}


def cbox_readme_noboxes(B, size=(640, 640)):
    """The README scene without its two Box geometries: cones + rects only, i.e. a scene whose
    closest hits go through the wide tree although a Cone is unbounded below its base."""
    floor = B.add_material("diffuse", (0.9, 0.9, 0.9))
    red = B.add_material("metal", (0.9, 0.4, 0.3), roughness=0.3)
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((0, 2, 1), (2, 0, 0), (0, 0, 2), floor)
    B.add_rect((-1, 1, 1), (0, 0, 2), (0, 2, 0), red)
    B.add_rect((1, 1, 1), (0, 0, 2), (0, 2, 0), floor)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), floor)
    for x in _frange(-1.0, 0.1, 1.0):
        for z in _frange(0.0, 0.1, 2.0):
            B.add_cone((x, 0.3, z), (0, 1, 0), 0.05, 0.05, red if int(round(x * 10)) % 2 else floor)
    light = B.add_material("emissive", (600.0, 384.0, 111.0))
    B.add_rect((0.0, 1.9, 0.5), (0.1, 0, 0), (0, 0, 0.1), light)
    B.set_camera(size, (0, 1, -4), (0, 1, 0), 0.25)
    return {"spp": 16, "depth": 4}


def materials_room(B, size=(96, 96)):
    """Parity scene for the material kinds the named configs never assign: Glass, Uber and
    Subsurface (on a mesh, where the reference's BSSRDF walk is well defined) next to Diffuse,
    Metal and Glossy, inside the cbox-indirect room."""
    white = B.add_material("diffuse", (0.9, 0.9, 0.9))
    red = B.add_material("diffuse", (0.9, 0.1, 0.05))
    glass = B.add_material("glass", (0.95, 0.95, 1.0), roughness=0.0, ior=1.5)
    frosted = B.add_material("glass", (0.9, 1.0, 0.9), roughness=0.3, ior=1.4)
    uber = B.add_material("uber", (0.8, 0.6, 0.2), roughness=0.25, metallic=0.4, transmission=0.3, ior=1.45)
    skin = B.add_material("subsurface", (0.9, 0.7, 0.6), roughness=0.1, sigma_s=(20.0, 30.0, 40.0))
    metal = B.add_material("metal", (0.9, 0.9, 0.9), roughness=0.15)
    glossy = B.add_material("glossy", (0.3, 0.5, 0.8), roughness=0.1)
    B.add_rect((0, 0, 1), (2, 0, 0), (0, 0, 2), white)
    B.add_rect((0, 2, 1), (2, 0, 0), (0, 0, 2), white)
    B.add_rect((0, 1, 2), (2, 0, 0), (0, 2, 0), glossy)
    B.add_rect((-1, 1, 1), (0, 0, 2), (0, 2, 0), red)
    B.add_rect((1, 1, 1), (0, 0, 2), (0, 2, 0), metal)
    B.add_sphere((-0.5, 0.35, 0.7), 0.35, glass)
    B.add_sphere((0.5, 0.3, 0.5), 0.3, uber)
    B.add_sphere((0.0, 0.25, 1.3), 0.25, frosted)
    v, f = icosphere(2)
    B.add_mesh((v * np.float32(0.3) + np.array([0.1, 1.0, 0.9], np.float32)).astype(np.float32), f, skin)
    light = B.add_material("emissive", (60.0, 50.0, 40.0))
    B.add_rect((0, 1.95, 1), (0.6, 0, 0), (0, 0, 0.6), light)
    B.set_camera(size, (0, 1, -3.5), (0, 1, 0), 0.28)
    return {"spp": 16, "depth": 6}


def classic(B, size=(720, 360), skin=False):
    """scenes/classic.pine:1-18: a checkerboard Uber floor -- albedo and roughness are node graphs,
    lerp(Checkerboard(UV(), 0.95), a, b) -- under three unit spheres (the middle one Subsurface in the file;
    Diffuse here unless `skin`, because the reference's BSSRDF walk is ill-defined on analytic shapes)
    and a 2x2 Rect emitter. The file renders it with MicroRenderGI and keeps the PathIntegrator line commented."""
    from . import nodes as N
    check = lambda: N.Checkerboard(N.UV(), 0.95)
    floor = B.add_material("uber", N.lerp(check(), (0.5, 0.7, 1.0), (0.01, 0.02, 0.03)),
                           roughness=N.lerp(check(), 0.0, 0.4), metallic=0.0, transmission=0.0, ior=1.45)
    if skin:
        middle = B.add_material("subsurface", (0, 0, 0), roughness=0.0, sigma_s=(1.0, 0.8, 0.6))
    else:
        middle = B.add_material("diffuse", (0.8, 0.8, 0.8))
    metal = B.add_material("uber", (1.0, 1.0, 1.0), roughness=0.0, metallic=1.0, transmission=0.0, ior=1.45)
    glossy = B.add_material("uber", (0.98, 0.55, 0.02), roughness=0.0, metallic=0.0, transmission=0.0, ior=1.45)
    B.add_disk((0, 0, 0), (0, 1, 0), 100.0, floor)
    B.add_sphere((-3, 1, 0), 1.0, metal)
    B.add_sphere((0, 1, 0), 1.0, middle)
    B.add_sphere((3, 1, 0), 1.0, glossy)
    light = B.add_material("emissive", tuple(np.float32(160) * np.float32(c) for c in (1, 1, 1)))
    B.add_rect((-1, 3, -1), (1, 0, 0), (0, 0, 1), light)
    B.set_camera(size, (0, 4, -8), (0, 1, 0), 0.3)
    return {"spp": 64, "depth": 6}


def plane_yard(B, size=(128, 96)):
    """Parity scene for Plane and Cylinder (geometry.cpp:31-70, 466-517): a Plane floor and a Plane back wall (both
    reach beyond the +-100 box the reference's BVH gives them), upright, lying and tilted Cylinders -- one long enough
    to leave the radius-r box around p0 that is all the reference bounds it by --, two spheres and a Rect emitter."""
    floor = B.add_material("diffuse", (0.8, 0.8, 0.8))
    wall = B.add_material("glossy", (0.3, 0.5, 0.8), roughness=0.2)
    red = B.add_material("diffuse", (0.9, 0.2, 0.1))
    metal = B.add_material("metal", (0.9, 0.9, 0.9), roughness=0.1)
    green = B.add_material("diffuse", (0.2, 0.8, 0.2))
    B.add_plane((0, 0, 0), (0, 1, 0), floor)
    B.add_plane((0, 0, 6), (0.1, 0.05, -1), wall)
    B.add_cylinder((-2, 0, 2), (-2, 2.5, 2), 0.5, red)
    B.add_cylinder((-1, 0.4, 1), (2.5, 0.4, 3), 0.4, metal)
    B.add_cylinder((2, 0, 0.5), (3, 3, 2.5), 0.3, green)
    B.add_sphere((0.5, 0.6, 2.5), 0.6, metal)
    B.add_sphere((-0.5, 0.3, 0.5), 0.3, red)
    light = B.add_material("emissive", (40.0, 36.0, 30.0))
    B.add_rect((0, 5, 2), (2, 0, 0), (0, 0, 2), light)
    B.set_camera(size, (0, 2.5, -6), (0, 1, 2), 0.35)
    return {"spp": 16, "depth": 5}


def lights_yard(B, size=(96, 64), sky=True, delta=True, area=True):
    """Parity scene for the lights the named configs never use (light.h:21-66): an open yard (ground, one
    wall, a sphere, a cone, a box) lit by a DirectionalLight, a PointLight, a SpotLight (with a cutoff band),
    a small Emissive rect ADDED BETWEEN them (Scene::lights keeps insertion order, scene.cpp:19-31) and
    a Sky environment light (rays that leave the scene see it, light.cpp:71-84)."""
    ground = B.add_material("diffuse", (0.6, 0.6, 0.55))
    wall = B.add_material("glossy", (0.7, 0.3, 0.25), roughness=0.2)
    ball = B.add_material("metal", (0.9, 0.85, 0.6), roughness=0.25)
    cone = B.add_material("diffuse", (0.2, 0.5, 0.8))
    B.add_rect((0, 0, 0), (8, 0, 0), (0, 0, 8), ground)
    if delta:
        B.add_directional_light((1.0, 2.0, -1.5), (2.5, 2.4, 2.2))
    B.add_rect((0, 1, 2), (4, 0, 0), (0, 2, 0), wall)
    B.add_sphere((-0.9, 0.5, 0.3), 0.5, ball)
    if area:
        lamp = B.add_material("emissive", (30.0, 24.0, 16.0))
        B.add_rect((1.6, 1.8, 0.6), (0.5, 0, 0), (0, 0, 0.5), lamp)
    if delta:
        B.add_point_light((-1.5, 1.2, -0.8), (3.0, 4.0, 6.0))
    B.add_cone((0.6, 0.0, 0.2), (0, 1, 0), 0.35, 0.9, cone)
    B.add_box((0, 0, 0), (1, 1, 1), mat4.mul(mat4.translate((1.4, 0.0, -0.9)), mat4.mul(mat4.rotate_y(0.5), mat4.scale((0.5, 0.7, 0.5)))), ground)
    if delta:
        B.add_spot_light((0.0, 2.6, -1.2), (0.1, -1.0, 0.45), (40.0, 40.0, 36.0), 0.35, 0.2)
    if sky:
        B.set_sky((0.9, 1.0, 1.2))
    B.set_camera(size, (0.4, 1.5, -4.5), (0.1, 0.7, 0), 0.3)
    return {"spp": 16, "depth": 5}


SCENES = {"materials_room": materials_room, "classic": classic, "plane_yard": plane_yard, "lights_yard": lights_yard, "cbox_readme": cbox_readme, "cbox_readme_noboxes": cbox_readme_noboxes, "cbox_file": cbox_file, "cbox_indirect": cbox_indirect,
          "mesh_field": mesh_field, "room_proc": room_proc}
