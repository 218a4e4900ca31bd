"""ctypes declarations of include/pine_b200.h and include/pine_b200_host.h.

The shared library is built in-tree by `__graft_entry__.build()` (pine_b200/csrc/Makefile ->
pine_b200/lib/libpine_b200.so). There is no Python or CPU implementation behind these calls: if
the library is missing, importing a symbol raises; if no CUDA device is present, every compute
entry point returns PB2_E_CUDA and the wrappers raise `PineB200Error`.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PINE_B200_LIB") or os.path.join(HERE, "lib", "libpine_b200.so")  # env: tuning variants
INCLUDE_DIR = os.path.join(os.path.dirname(HERE), "include")

PB2_OK, PB2_E_INVALID, PB2_E_UNSUPPORTED, PB2_E_CUDA, PB2_E_STATE = 0, -1, -2, -3, -4

f3 = C.c_float * 3
f16 = C.c_float * 16


class Rect(C.Structure):
    _fields_ = [("position", f3), ("ex", f3), ("ey", f3), ("n", f3), ("lx", C.c_float),
                ("ly", C.c_float), ("rx", f3), ("ry", f3)]


class Aabb(C.Structure):
    _fields_ = [("lower", f3), ("upper", f3)]


class Obb(C.Structure):
    _fields_ = [("base", Aabb), ("m", f16), ("m_inv", f16)]


class Disk(C.Structure):
    _fields_ = [("position", f3), ("n", f3), ("u", f3), ("v", f3), ("r", C.c_float)]


class Cone(C.Structure):
    _fields_ = [("bottom", Disk), ("p", f3), ("n", f3), ("r", C.c_float), ("h", C.c_float),
                ("A", C.c_float), ("A2", C.c_float), ("S", C.c_float)]


class Plane(C.Structure):
    _fields_ = [("position", f3), ("n", f3), ("u", f3), ("v", f3)]


class Cylinder(C.Structure):
    _fields_ = [("p0", f3), ("p1", f3), ("n", f3), ("r", C.c_float), ("bottom", Disk), ("top", Disk)]


class Sphere(C.Structure):
    _fields_ = [("c", f3), ("r", C.c_float)]


class Triangle(C.Structure):
    _fields_ = [("v0", f3), ("v1", f3), ("v2", f3), ("n", f3)]


class Mesh(C.Structure):
    _fields_ = [("vertices", C.c_void_p), ("indices", C.c_void_p), ("normals", C.c_void_p),
                ("texcoords", C.c_void_p), ("n_vertices", C.c_uint32), ("n_triangles", C.c_uint32)]


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", f3), ("roughness", C.c_float), ("ior", C.c_float),
                ("metallic", C.c_float), ("transmission", C.c_float), ("sigma_s", f3),
                ("color_texture", C.c_int32), ("roughness_texture", C.c_int32),
                ("metallic_texture", C.c_int32),  # 1-based, 0 = none
                # 1-based start of the parameter's node program in SceneDesc.node_instrs, 0 = constant
                ("color_node", C.c_int32), ("roughness_node", C.c_int32), ("ior_node", C.c_int32),
                ("metallic_node", C.c_int32), ("transmission_node", C.c_int32)]


class NodeInstr(C.Structure):
    _fields_ = [("op", C.c_int32), ("a", f3)]


NODE_SLOTS = ("color", "roughness", "ior", "metallic", "transmission")


class Texture(C.Structure):
    _fields_ = [("rgba", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32)]


class Geometry(C.Structure):
    _fields_ = [("shape_kind", C.c_int32), ("shape_index", C.c_int32), ("material", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("position", f3), ("c2w", C.c_float * 9), ("fov2d", C.c_float * 2),
                ("len_radius", C.c_float), ("focus_distance", C.c_float), ("width", C.c_int32),
                ("height", C.c_int32), ("tonemapper", C.c_int32)]


def _arr(name, typ):
    return [(name, C.POINTER(typ)), ("n_" + name, C.c_uint32)]


class DeviceArray(C.Structure):
    _fields_ = [("device_ptr", C.c_void_p), ("bytes", C.c_uint64), ("part_bytes", C.c_uint64)]


class Light(C.Structure):
    _fields_ = [("kind", C.c_int32), ("geometry", C.c_int32), ("position", f3), ("direction", f3),
                ("color", f3), ("falloff_cos", C.c_float), ("cutoff_cos", C.c_float)]


LIGHT_KINDS = {"area": 0, "point": 1, "spot": 2, "directional": 3, "sky": 4}


class SceneDesc(C.Structure):
    _fields_ = (_arr("geometries", Geometry) + _arr("materials", Material) + _arr("aabbs", Aabb) +
                _arr("obbs", Obb) + _arr("spheres", Sphere) + _arr("triangles", Triangle) +
                _arr("cones", Cone) + _arr("rects", Rect) + _arr("disks", Disk) +
                _arr("meshes", Mesh) +
                [("light_geometries", C.POINTER(C.c_int32)), ("n_lights", C.c_uint32),
                 ("camera", Camera)] + _arr("textures", Texture) + _arr("light_records", Light) +
                _arr("node_instrs", NodeInstr) + _arr("planes", Plane) + _arr("cylinders", Cylinder))


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("max_path_length", C.c_int32), ("row_begin", C.c_int32),
                ("row_end", C.c_int32), ("sample_begin", C.c_int32), ("sample_end", C.c_int32),
                ("accel", C.c_int32), ("flags", C.c_int32), ("sampler", C.c_int32),
                ("sampler_seed", C.c_int32), ("row_band", C.c_int32), ("row_parts", C.c_int32),
                ("row_part", C.c_int32)]


SAMPLER_BLUE, SAMPLER_SOBOL, SAMPLER_UNIFORM, SAMPLER_HALTON = 0, 1, 2, 3


class RayRecord(C.Structure):
    _fields_ = [("capacity", C.c_uint64), ("n_closest", C.c_uint64), ("n_shadow", C.c_uint64),
                ("closest_rays", C.c_void_p), ("closest_t", C.c_void_p),
                ("closest_geom", C.c_void_p), ("closest_prim", C.c_void_p),
                ("shadow_rays", C.c_void_p), ("shadow_occluded", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("camera_samples", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("last_render_ms", C.c_double), ("last_trace_ms", C.c_double),
                ("last_shadow_ms", C.c_double), ("last_shade_ms", C.c_double),
                ("last_other_ms", C.c_double), ("last_closest_rays", C.c_uint64),
                ("last_shadow_rays", C.c_uint64), ("last_trace_launches", C.c_uint64),
                ("bvh_build_ms", C.c_double), ("upload_ms", C.c_double),
                ("scene_bytes", C.c_uint64), ("bvh2_nodes", C.c_uint32),
                ("wide_nodes", C.c_uint32), ("n_primitives", C.c_uint32),
                ("build_threads", C.c_uint32), ("last_closest_fallbacks", C.c_uint64),
                ("last_shadow_fallbacks", C.c_uint64)]


SHAPE_KINDS = {"aabb": 0, "obb": 1, "sphere": 2, "triangle": 3, "cone": 4, "rect": 5, "disk": 6,
               "mesh": 7, "plane": 8, "cylinder": 9}
MATERIAL_KINDS = {"emissive": 0, "diffuse": 1, "metal": 2, "glossy": 3, "glass": 4, "uber": 5,
                  "subsurface": 6}

# every symbol include/*.h declares (tests check the built library exports all of them)
ABI_SYMBOLS = [
    "pb2_last_error", "pb2_device_count", "pb2_create", "pb2_destroy", "pb2_upload_scene",
    "pb2_restage_scene", "pb2_restage_scene_part",
    "pb2_upload_sampler_tables", "pb2_load_sampler_tables_file", "pb2_render", "pb2_render_device", "pb2_film_finalize",
    "pb2_trace_closest", "pb2_trace_any", "pb2_trace_closest_device", "pb2_test_sampler", "pb2_test_sampler_kind",
    "pb2_test_primary_rays", "pb2_get_counters", "pb2_debug_walk_stats", "pb2_set_ray_record", "pb2_bvh_info", "pb2_bvh_nodes",
    "pb2h_make_rect", "pb2h_make_obb", "pb2h_make_disk", "pb2h_make_plane", "pb2h_make_cylinder", "pb2h_make_cone", "pb2h_make_triangle",
    "pb2h_make_camera", "pb2h_mat4", "pb2h_mat4_mul", "pb2h_mesh_apply", "pb2h_cosf", "pb2h_round_spp", "pb2h_sampler_draws",
    "pb2_create_multi", "pb2_destroy_multi", "pb2_multi_size", "pb2_multi_context", "pb2_multi_upload_sampler_tables",
    "pb2_multi_load_sampler_tables_file", "pb2_multi_upload_scene", "pb2_multi_restage_scene", "pb2_multi_render",
    "pb2_multi_get_counters", "pb2_measure_read_gbs", "pb2_film_finalize_device", "pb2_multi_film_finalize",
    "pb2_export_host_scene", "pb2_import_host_scene", "pb2_trace_closest_device_noalloc",
    "pb2_set_progress_callback", "pb2_multi_set_progress_callback", "pb2_test_eval_node",
]
SHARD_SAMPLES, SHARD_ROWS = 0, 1
PROGRESS_FN = C.CFUNCTYPE(None, C.c_float, C.c_void_p)  # pb2_progress_fn
NOALLOC_MAX_RAYS = 1 << 20  # PB2_NOALLOC_MAX_RAYS

_lib = None


class PineB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pine_b200 error {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                "pine_b200 has no Python/CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.pb2_last_error.restype = C.c_char_p
        L.pb2_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.pb2_destroy.argtypes = [C.c_void_p]
        L.pb2_destroy.restype = None
        L.pb2_upload_scene.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
        L.pb2_restage_scene.argtypes = [C.c_void_p, C.c_void_p]
        L.pb2_restage_scene_part.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                             C.c_void_p]
        L.pb2_upload_sampler_tables.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                C.c_void_p]
        L.pb2_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p]
        L.pb2_render_device.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p,
                                        C.POINTER(C.c_void_p)]
        L.pb2_film_finalize.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.pb2_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb2_trace_any.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.pb2_trace_closest_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb2_trace_closest_device_noalloc.argtypes = L.pb2_trace_closest_device.argtypes
        L.pb2_set_progress_callback.argtypes = [C.c_void_p, PROGRESS_FN, C.c_void_p]
        L.pb2_multi_set_progress_callback.argtypes = [C.c_void_p, PROGRESS_FN, C.c_void_p]
        L.pb2_test_eval_node.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
        L.pb2_test_sampler_kind.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_char_p, C.c_int, C.c_void_p]
        L.pb2_test_sampler.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_char_p, C.c_int, C.c_void_p]
        L.pb2_test_primary_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.pb2_set_ray_record.argtypes = [C.c_void_p, C.POINTER(RayRecord)]
        L.pb2_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
        L.pb2_bvh_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32),
                                   C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.pb2_bvh_nodes.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]
        L.pb2_film_finalize_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.pb2_multi_film_finalize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.pb2_export_host_scene.argtypes = [C.c_void_p, C.c_char_p]
        L.pb2_import_host_scene.argtypes = [C.c_void_p, C.c_char_p]
        L.pb2_measure_read_gbs.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        L.pb2_create_multi.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.pb2_destroy_multi.argtypes = [C.c_void_p]
        L.pb2_destroy_multi.restype = None
        L.pb2_multi_size.argtypes = [C.c_void_p]
        L.pb2_multi_context.argtypes = [C.c_void_p, C.c_int]
        L.pb2_multi_context.restype = C.c_void_p
        L.pb2_multi_upload_sampler_tables.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb2_multi_load_sampler_tables_file.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.pb2_multi_upload_scene.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
        L.pb2_multi_restage_scene.argtypes = [C.c_void_p]
        L.pb2_multi_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_int, C.c_void_p]
        L.pb2_multi_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters), C.POINTER(C.c_double),
                                             C.POINTER(C.c_double)]
        fp = C.POINTER(C.c_float)
        L.pb2h_make_rect.argtypes = [fp, fp, fp, C.c_int, C.POINTER(Rect)]
        L.pb2h_make_obb.argtypes = [fp, fp, fp, C.POINTER(Obb)]
        L.pb2h_make_disk.argtypes = [fp, fp, C.c_float, C.POINTER(Disk)]
        L.pb2h_make_cone.argtypes = [fp, fp, C.c_float, C.c_float, C.POINTER(Cone)]
        L.pb2h_make_triangle.argtypes = [fp, fp, fp, C.POINTER(Triangle)]
        L.pb2h_make_camera.argtypes = [C.c_int, C.c_int, C.c_int, fp, fp, C.c_float, C.c_float,
                                       C.c_float, C.POINTER(Camera)]
        L.pb2h_mat4.argtypes = [C.c_int, fp, fp]
        L.pb2h_mat4_mul.argtypes = [fp, fp, fp]
        L.pb2h_cosf.restype = C.c_float
        L.pb2h_cosf.argtypes = [C.c_float]
        L.pb2h_mesh_apply.argtypes = [fp, fp, C.c_size_t, fp, C.c_size_t]
        L.pb2h_sampler_draws.argtypes = [C.c_int] * 8 + [C.c_char_p, C.c_int, C.c_void_p]
        for name in ("pb2h_make_rect", "pb2h_make_obb", "pb2h_make_disk", "pb2h_make_cone",
                     "pb2h_make_triangle", "pb2h_make_camera", "pb2h_mat4", "pb2h_mat4_mul"):
            getattr(L, name).restype = None
        _lib = L
    return _lib


def check(rc):
    if rc != PB2_OK:
        raise PineB200Error(rc, lib().pb2_last_error().decode())
