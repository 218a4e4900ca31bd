// integration/cuda_path_integrator.cpp -- see cuda_path_integrator.h.
//
// flatten(): pine::Scene -> pb2_scene_desc. Shape records in include/pine_b200.h carry exactly the
// data members of the reference structs in declaration order (geometry.h:44-139, bbox.h:29-109), so
// a shape is copied bit for bit (the static_asserts below pin the layouts): both sides then
// intersect the same numbers, including the host-computed OBB::m_inv (bbox.cpp:143-147). Material
// parameters are Node graphs (node.h:46-103): a NodeConstant is copied, scene_from_gltf's image patterns become
// device textures, and any other graph is flattened to a postfix program the shade kernels evaluate
// (pb2_node_op); noise and function nodes -- and a non-constant Emissive::color -- are rejected (SEVERE, like
// the reference's own argument checks, log.h:44-49): there is no CPU fallback.
// Image-node material parameters (what scene_from_gltf builds for baseColorTexture / metallicRoughnessTexture,
// fileio.cpp:285-292) have to be recognised structurally, and Mnode<T>::value, NodeImage::{p, image} and
// NodeComponent::{x, n} are private (node.h:69-70, 102-103, 181-193, 236-243). A maintainer adds two accessors
// (INTEGRATION.md section 2); this file, which must compile against the UNMODIFIED headers, opens `private`
// around node.h only -- access specifiers do not change the layout.
#define private public
#include <pine/core/node.h>
#undef private
#include "cuda_path_integrator.h"

#include <pine/core/scene.h>
#include <pine/core/image.h>
#include <pine/core/log.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "../include/pine_b200.h"

// The Heitz et al. tables the reference links in (src/contrib/bluesobol/bluenoise_<N>spp.cpp:2-12)
// are file-static there; a maintainer exposes them with three accessor lines per file. Until then
// the shim reads the same bytes from the blob pine_b200/tools/extract_tables.cpp wrote from those sources.

namespace pine {

namespace {

static_assert(sizeof(Rect) == sizeof(pb2_rect), "pb2_rect must mirror pine::Rect");
static_assert(sizeof(AABB) == sizeof(pb2_aabb), "pb2_aabb must mirror pine::AABB");
static_assert(sizeof(OBB) == sizeof(pb2_obb), "pb2_obb must mirror pine::OBB");
static_assert(sizeof(Disk) == sizeof(pb2_disk), "pb2_disk must mirror pine::Disk");
static_assert(sizeof(Plane) == sizeof(pb2_plane), "pb2_plane must mirror pine::Plane");
static_assert(sizeof(Cylinder) == sizeof(pb2_cylinder), "pb2_cylinder must mirror pine::Cylinder");
static_assert(sizeof(Cone) == sizeof(pb2_cone), "pb2_cone must mirror pine::Cone");
static_assert(sizeof(Sphere) == sizeof(pb2_sphere), "pb2_sphere must mirror pine::Sphere");
static_assert(sizeof(Triangle) == sizeof(pb2_triangle), "pb2_triangle must mirror pine::Triangle");
static_assert(sizeof(vec3) == 12 && sizeof(vec2) == 8 && sizeof(vec3u32) == 12 && sizeof(vec4) == 16);

template <typename Dst, typename Src>
Dst bits(const Src& s) {
  Dst d;
  std::memcpy(&d, &s, sizeof(Dst));
  return d;
}

const NodeEvalCtx kProbe[3] = {
    NodeEvalCtx(vec3(0.0f), vec3(0, 0, 1), vec2(0.0f)),
    NodeEvalCtx(vec3(0.37f, -1.91f, 5.3f), vec3(0, 1, 0), vec2(0.31f, 0.77f)),
    NodeEvalCtx(vec3(-7.1f, 0.53f, 0.11f), vec3(1, 0, 0), vec2(0.93f, 0.12f)),
};
template <typename N>
auto constant_of(const N& node, const char* what) {
  auto v = node.eval(kProbe[0]);
  for (int i = 1; i < 3; i++)
    if (!(node.eval(kProbe[i]) == v))
      SEVERE("[CudaPathIntegrator]material parameter `", what,
             "` is not a constant node; textured/procedural parameters are not supported by the "
             "CUDA path (and there is no CPU fallback)");
  return v;
}
void put3(float* dst, vec3 v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }

struct Flat;
// A parameter that is neither a constant nor one of scene_from_gltf's image patterns: the graph is flattened to
// a postfix program (include/pine_b200.h pb2_node_op) appended to Flat::node_instrs -> 1-based start index.
int32_t node_program(Flat& f, const Node3f& node, const char* what);
int32_t node_program(Flat& f, const Nodef& node, const char* what);
template <typename N>
bool is_constant_node(const N& node) {
  return node.value->template is<NodeConstant<decltype(node.eval(kProbe[0]))>>();
}
// NodeImage(NodeUV(), image) (node.cpp:20-23) over an 8-bit RGBA image -> 1-based pb2_texture index; 0 if
// `node` is not an image node. Any other non-constant node is rejected by constant_of afterwards.
int32_t image_texture(Flat& f, const Node3f& node);
int32_t component_texture(Flat& f, const Nodef& node, int component, const char* what);

// constant -> *dst, anything else -> a program in *node (Emissive::color keeps the constant-only rule)
void param3(Flat& f, const Node3f& node, float* dst, int32_t* node_slot, const char* what) {
  if (is_constant_node(node)) put3(dst, node.eval(kProbe[0]));
  else *node_slot = node_program(f, node, what);
}
void paramf(Flat& f, const Nodef& node, float* dst, int32_t* node_slot, const char* what) {
  if (is_constant_node(node)) *dst = node.eval(kProbe[0]);
  else *node_slot = node_program(f, node, what);
}

pb2_material flatten_material(Flat& f, const Material& m) {
  pb2_material o{};
  o.ior = 1.4f;
  if (m.is<EmissiveMaterial>()) {
    o.kind = PB2_MAT_EMISSIVE;
    put3(o.color, constant_of(m.as<EmissiveMaterial>().color, "Emissive.color"));
  } else if (m.is<DiffuseMaterial>()) {
    o.kind = PB2_MAT_DIFFUSE;
    param3(f, m.as<DiffuseMaterial>().albedo, o.color, &o.color_node, "Diffuse.albedo");
  } else if (m.is<MetalMaterial>()) {
    auto& x = m.as<MetalMaterial>();
    o.kind = PB2_MAT_METAL;
    param3(f, x.albedo, o.color, &o.color_node, "Metal.albedo");
    paramf(f, x.roughness, &o.roughness, &o.roughness_node, "Metal.roughness");
  } else if (m.is<GlossyMaterial>()) {
    auto& x = m.as<GlossyMaterial>();
    o.kind = PB2_MAT_GLOSSY;
    param3(f, x.albedo, o.color, &o.color_node, "Glossy.albedo");
    paramf(f, x.roughness, &o.roughness, &o.roughness_node, "Glossy.roughness");
    paramf(f, x.ior, &o.ior, &o.ior_node, "Glossy.ior");
  } else if (m.is<GlassMaterial>()) {
    auto& x = m.as<GlassMaterial>();
    o.kind = PB2_MAT_GLASS;
    param3(f, x.albedo, o.color, &o.color_node, "Glass.albedo");
    paramf(f, x.roughness, &o.roughness, &o.roughness_node, "Glass.roughness");
    paramf(f, x.ior, &o.ior, &o.ior_node, "Glass.ior");
  } else if (m.is<UberMaterial>()) {
    auto& x = m.as<UberMaterial>();
    o.kind = PB2_MAT_UBER;
    // image nodes become device textures; the constants they replace are unused (render.cu material_bxdf)
    o.color_texture = image_texture(f, x.albedo);
    o.roughness_texture = component_texture(f, x.roughness, 1, "Uber.roughness");
    o.metallic_texture = component_texture(f, x.metallic, 2, "Uber.metallic");
    if (!o.color_texture) param3(f, x.albedo, o.color, &o.color_node, "Uber.albedo");
    if (!o.roughness_texture) paramf(f, x.roughness, &o.roughness, &o.roughness_node, "Uber.roughness");
    if (!o.metallic_texture) paramf(f, x.metallic, &o.metallic, &o.metallic_node, "Uber.metallic");
    paramf(f, x.transmission, &o.transmission, &o.transmission_node, "Uber.transmission");
    o.ior = x.ior;
  } else {
    auto& x = m.as<SubsurfaceMaterial>();
    o.kind = PB2_MAT_SUBSURFACE;
    param3(f, x.albedo, o.color, &o.color_node, "Subsurface.albedo");
    paramf(f, x.roughness, &o.roughness, &o.roughness_node, "Subsurface.roughness");
    put3(o.sigma_s, x.sigma_s);
    o.ior = x.ior;
  }
  return o;
}

struct Flat {
  std::vector<pb2_geometry> geometries;
  std::vector<pb2_material> materials;
  std::vector<const Material*> material_keys;
  std::vector<pb2_aabb> aabbs;
  std::vector<pb2_obb> obbs;
  std::vector<pb2_sphere> spheres;
  std::vector<pb2_triangle> triangles;
  std::vector<pb2_cone> cones;
  std::vector<pb2_rect> rects;
  std::vector<pb2_disk> disks;
  std::vector<pb2_plane> planes;
  std::vector<pb2_cylinder> cylinders;
  std::vector<pb2_mesh> meshes;
  std::vector<int32_t> lights;
  std::vector<pb2_light> light_records;
  std::vector<pb2_texture> textures;
  std::vector<const Image*> texture_keys;
  std::vector<pb2_node_instr> node_instrs;
  pb2_scene_desc desc{};
};

// ---- Node graph -> postfix program (node.h:105-250) ----
void emit(Flat& f, const Node3f& node, const char* what);
void emit(Flat& f, const Nodef& node, const char* what);
void ins(Flat& f, int op, float a0 = 0, float a1 = 0, float a2 = 0) {
  f.node_instrs.push_back(pb2_node_instr{op, {a0, a1, a2}});
}
template <typename T, typename M>
bool emit_arith(Flat& f, const M& node, const char* what) {
  auto& v = *node.value;
#define PB2_BIN(ch, code) \
  if (v.template is<NodeBinary<T, ch>>()) { auto& x = v.template as<NodeBinary<T, ch>>(); emit(f, x.a, what); emit(f, x.b, what); ins(f, code); return true; }
#define PB2_UN(ch, code) \
  if (v.template is<NodeUnary<T, ch>>()) { emit(f, v.template as<NodeUnary<T, ch>>().x, what); ins(f, code); return true; }
  PB2_BIN('+', PB2_NODE_ADD) PB2_BIN('-', PB2_NODE_SUB) PB2_BIN('*', PB2_NODE_MUL) PB2_BIN('/', PB2_NODE_DIV)
  PB2_BIN('^', PB2_NODE_POW)
  PB2_UN('-', PB2_NODE_NEG) PB2_UN('a', PB2_NODE_ABS) PB2_UN('s', PB2_NODE_SQR) PB2_UN('r', PB2_NODE_SQRT)
  PB2_UN('f', PB2_NODE_FRACT)
#undef PB2_BIN
#undef PB2_UN
  return false;
}
void emit(Flat& f, const Node3f& node, const char* what) {
  auto& v = *node.value;
  if (v.is<NodeConstant<vec3>>()) { const vec3 c = node.eval(kProbe[0]); ins(f, PB2_NODE_CONST, c.x, c.y, c.z); }
  else if (v.is<NodePosition>()) ins(f, PB2_NODE_POSITION);
  else if (v.is<NodeNormal>()) ins(f, PB2_NODE_NORMAL);
  else if (v.is<NodeUV>()) ins(f, PB2_NODE_UV);
  else if (emit_arith<vec3>(f, node, what)) {}
  else if (v.is<NodeToVec3>()) {
    auto& x = v.as<NodeToVec3>();
    emit(f, x.x, what);  // the one-argument form is a broadcast: the identity on the device stack
    if (x.y) { emit(f, *x.y, what); emit(f, *x.z, what); ins(f, PB2_NODE_VEC3); }
  } else if (v.is<NodeImage>()) {
    auto& x = v.as<NodeImage>();
    emit(f, x.p, what);
    Node3f over_uv{NodeImage(Node3f(NodeUV()), x.image)};
    ins(f, PB2_NODE_IMAGE, float(image_texture(f, over_uv)));
  } else
    SEVERE("[CudaPathIntegrator]material parameter `", what, "` uses a noise or function node, which has no CUDA "
           "form (and there is no CPU fallback)");
}
void emit(Flat& f, const Nodef& node, const char* what) {
  auto& v = *node.value;
  if (v.is<NodeConstant<float>>()) { const float c = node.eval(kProbe[0]); ins(f, PB2_NODE_CONST, c, c, c); }
  else if (emit_arith<float>(f, node, what)) {}
  else if (v.is<NodeComponent>()) {
    auto& x = v.as<NodeComponent>();
    emit(f, x.x, what);
    ins(f, PB2_NODE_COMPONENT, float(x.n));
  } else if (v.is<NodeCheckerboard>()) {
    auto& x = v.as<NodeCheckerboard>();
    emit(f, x.p, what);
    ins(f, PB2_NODE_CHECKERBOARD, x.ratio);
  } else if (v.is<NodeImagef>()) {
    auto& x = v.as<NodeImagef>();
    emit(f, x.p, what);
    Node3f over_uv{NodeImage(Node3f(NodeUV()), x.image)};
    ins(f, PB2_NODE_IMAGEF, float(image_texture(f, over_uv)));
  } else
    SEVERE("[CudaPathIntegrator]material parameter `", what, "` uses a noise or function node, which has no CUDA "
           "form (and there is no CPU fallback)");
}
int32_t node_program(Flat& f, const Node3f& node, const char* what) {
  const int32_t start = int32_t(f.node_instrs.size()) + 1;
  emit(f, node, what);
  ins(f, PB2_NODE_END);
  return start;
}
int32_t node_program(Flat& f, const Nodef& node, const char* what) {
  const int32_t start = int32_t(f.node_instrs.size()) + 1;
  emit(f, node, what);
  ins(f, PB2_NODE_END);
  return start;
}

int32_t image_texture(Flat& f, const Node3f& node) {
  if (!node.value->is<NodeImage>()) return 0;
  const NodeImage& ni = node.value->as<NodeImage>();
  if (!ni.p.value->is<NodeUV>())
    SEVERE("[CudaPathIntegrator]image nodes are supported with UV coordinates only (NodeImage(NodeUV(), image))");
  const Image* img = ni.image.get();
  if (!img || !img->is<Array2d<vec4u8>>())
    SEVERE("[CudaPathIntegrator]image nodes need an 8-bit RGBA image (what scene_from_gltf loads)");
  for (size_t i = 0; i < f.texture_keys.size(); i++)
    if (f.texture_keys[i] == img) return int32_t(i) + 1;
  const auto& px = img->as<Array2d<vec4u8>>();
  pb2_texture t{};
  t.rgba = reinterpret_cast<const uint8_t*>(px.data());
  t.width = px.size().x;
  t.height = px.size().y;
  f.texture_keys.push_back(img);
  f.textures.push_back(t);
  return int32_t(f.textures.size());
}
int32_t component_texture(Flat& f, const Nodef& node, int component, const char* what) {
  if (!node.value->is<NodeComponent>()) return 0;
  const NodeComponent& nc = node.value->as<NodeComponent>();
  const int32_t t = image_texture(f, nc.x);
  if (!t) return 0;  // a component of something else: constant_of decides
  if (nc.n != component)
    SEVERE("[CudaPathIntegrator]`", what, "` reads component ", nc.n, " of an image; the CUDA path binds "
           "roughness to G and metallic to B as scene_from_gltf does (fileio.cpp:288-291)");
  return t;
}

template <typename T, typename S>
int32_t push(std::vector<T>& v, const S& s) {
  v.push_back(bits<T>(s));
  return int32_t(v.size()) - 1;
}

void flatten(Scene& scene, Flat& f) {
  if (scene.mediums.size())
    SEVERE("[CudaPathIntegrator]participating media are not supported by the CUDA path");
  if (scene.env_light && !scene.env_light->is<Sky>())
    SEVERE("[CudaPathIntegrator]only the Sky environment light is supported by the CUDA path");
  if (scene.instancings.size())  // pine::BVH ignores them too (bvh.cpp:453-495)
    SEVERE("[CudaPathIntegrator]Scene::instancings are not supported: add flattened Mesh geometries");
  for (size_t gi = 0; gi < scene.geometries.size(); gi++) {
    const Geometry& g = *scene.geometries[gi];
    pb2_geometry rec{};
    const Material* key = g.material.get();
    size_t mi = 0;
    while (mi < f.material_keys.size() && f.material_keys[mi] != key) mi++;
    if (mi == f.material_keys.size()) {
      f.material_keys.push_back(key);
      f.materials.push_back(flatten_material(f, *key));
    }
    rec.material = int32_t(mi);
    const Shape& s = g.shape;
    if (s.is<Rect>()) { rec.shape_kind = PB2_SHAPE_RECT; rec.shape_index = push(f.rects, s.as<Rect>()); }
    else if (s.is<OBB>()) { rec.shape_kind = PB2_SHAPE_OBB; rec.shape_index = push(f.obbs, s.as<OBB>()); }
    else if (s.is<AABB>()) { rec.shape_kind = PB2_SHAPE_AABB; rec.shape_index = push(f.aabbs, s.as<AABB>()); }
    else if (s.is<Cone>()) { rec.shape_kind = PB2_SHAPE_CONE; rec.shape_index = push(f.cones, s.as<Cone>()); }
    else if (s.is<Sphere>()) { rec.shape_kind = PB2_SHAPE_SPHERE; rec.shape_index = push(f.spheres, s.as<Sphere>()); }
    else if (s.is<Disk>()) { rec.shape_kind = PB2_SHAPE_DISK; rec.shape_index = push(f.disks, s.as<Disk>()); }
    else if (s.is<Triangle>()) { rec.shape_kind = PB2_SHAPE_TRIANGLE; rec.shape_index = push(f.triangles, s.as<Triangle>()); }
    else if (s.is<Plane>()) { rec.shape_kind = PB2_SHAPE_PLANE; rec.shape_index = push(f.planes, s.as<Plane>()); }
    else if (s.is<Cylinder>()) { rec.shape_kind = PB2_SHAPE_CYLINDER; rec.shape_index = push(f.cylinders, s.as<Cylinder>()); }
    else if (s.is<Mesh>()) {
      const Mesh& m = s.as<Mesh>();
      pb2_mesh pm{};
      pm.vertices = &m.vertices[0][0];
      pm.indices = &m.indices[0][0];
      pm.normals = m.normals.size() ? &m.normals[0][0] : nullptr;
      pm.texcoords = m.texcoords.size() ? &m.texcoords[0][0] : nullptr;
      pm.n_vertices = uint32_t(m.vertices.size());
      pm.n_triangles = uint32_t(m.indices.size());
      rec.shape_kind = PB2_SHAPE_MESH;
      f.meshes.push_back(pm);
      rec.shape_index = int32_t(f.meshes.size()) - 1;
    } else {
      SEVERE("[CudaPathIntegrator]geometry ", gi, " has a shape kind the CUDA path does not implement "
             "(Line/SDF/CSG); there is no CPU fallback");
    }
    f.geometries.push_back(rec);
  }
  // UniformLightSampler::lights (lightsampler.cpp:6-10): Scene::lights in order (AreaLights were added per
  // Emissive geometry, scene.cpp:19-20; the others by Scene::add_light), then the environment light
  bool only_area = !scene.env_light;
  for (const Light& l : scene.lights) {
    pb2_light r{};
    r.geometry = -1;
    if (l.is<AreaLight>()) {
      const Geometry* lg = l.as<AreaLight>().geometry.get();
      for (size_t i = 0; i < scene.geometries.size(); i++)
        if (scene.geometries[i].get() == lg) r.geometry = int32_t(i);
      if (r.geometry < 0) SEVERE("[CudaPathIntegrator]AreaLight over a geometry that is not in Scene::geometries");
      r.kind = PB2_LIGHT_AREA;
      f.lights.push_back(r.geometry);
    } else if (l.is<PointLight>()) {
      r.kind = PB2_LIGHT_POINT;
      put3(r.position, l.as<PointLight>().position);
      put3(r.color, l.as<PointLight>().color);
    } else if (l.is<SpotLight>()) {
      auto& x = l.as<SpotLight>();
      r.kind = PB2_LIGHT_SPOT;
      put3(r.position, x.position); put3(r.direction, x.direction); put3(r.color, x.color);
      r.falloff_cos = x.falloff_cos; r.cutoff_cos = x.cutoff_cos;
    } else if (l.is<DirectionalLight>()) {
      r.kind = PB2_LIGHT_DIRECTIONAL;
      put3(r.direction, l.as<DirectionalLight>().direction);
      put3(r.color, l.as<DirectionalLight>().color);
    } else {
      SEVERE("[CudaPathIntegrator]Atmosphere / ImageSky lights are not supported by the CUDA path");
    }
    only_area = only_area && r.kind == PB2_LIGHT_AREA;
    f.light_records.push_back(r);
  }
  if (scene.env_light) {
    // Sky's only data member is `vec3 sun_color` (light.h:57-66, private): read it by value
    static_assert(sizeof(Sky) == 3 * sizeof(float), "Sky is expected to hold sun_color only");
    pb2_light r{};
    r.kind = PB2_LIGHT_SKY;
    r.geometry = -1;
    std::memcpy(r.color, &scene.env_light->as<Sky>(), sizeof(r.color));
    f.light_records.push_back(r);
  }
  if (!scene.camera.is<ThinLenCamera>()) SEVERE("[CudaPathIntegrator]unsupported camera");
  auto& cam = scene.camera.as<ThinLenCamera>();
  pb2_camera& c = f.desc.camera;
  put3(c.position, cam.position);
  std::memcpy(c.c2w, &cam.c2w, sizeof(float) * 9);  // column-major mat3 (vecmath.h Matrix3 {x,y,z})
  c.fov2d[0] = cam.fov2d.x; c.fov2d[1] = cam.fov2d.y;
  c.len_radius = cam.len_radius;
  c.focus_distance = cam.focus_distance;
  c.width = cam.film_.width();
  c.height = cam.film_.height();
  c.tonemapper = 0;  // only used by pb2_film_finalize; Film::tone_mapper is private and the shim
                     // leaves Film::finalize to the host as the reference does

  pb2_scene_desc& d = f.desc;
  d.geometries = f.geometries.data(); d.n_geometries = uint32_t(f.geometries.size());
  d.materials = f.materials.data();   d.n_materials = uint32_t(f.materials.size());
  d.aabbs = f.aabbs.data();           d.n_aabbs = uint32_t(f.aabbs.size());
  d.obbs = f.obbs.data();             d.n_obbs = uint32_t(f.obbs.size());
  d.spheres = f.spheres.data();       d.n_spheres = uint32_t(f.spheres.size());
  d.triangles = f.triangles.data();   d.n_triangles = uint32_t(f.triangles.size());
  d.cones = f.cones.data();           d.n_cones = uint32_t(f.cones.size());
  d.rects = f.rects.data();           d.n_rects = uint32_t(f.rects.size());
  d.disks = f.disks.data();           d.n_disks = uint32_t(f.disks.size());
  d.meshes = f.meshes.data();         d.n_meshes = uint32_t(f.meshes.size());
  d.light_geometries = f.lights.data(); d.n_lights = uint32_t(f.lights.size());
  d.textures = f.textures.data();     d.n_textures = uint32_t(f.textures.size());
  d.node_instrs = f.node_instrs.data(); d.n_node_instrs = uint32_t(f.node_instrs.size());
  d.planes = f.planes.data();         d.n_planes = uint32_t(f.planes.size());
  d.cylinders = f.cylinders.data();   d.n_cylinders = uint32_t(f.cylinders.size());
  if (!only_area) { d.light_records = f.light_records.data(); d.n_light_records = uint32_t(f.light_records.size()); }
}

void check(int rc, const char* what) {
  if (rc != PB2_OK) SEVERE("[CudaPathIntegrator]", what, " failed: ", pb2_last_error());
}

}  // namespace

CudaPathIntegrator::CudaPathIntegrator(Accel, Sampler sampler, LightSampler, int max_path_length)
    : sampler_(MOVE(sampler)), max_path_length_(max_path_length) {
  if (max_path_length <= 0)  // path.cpp:12-13
    SEVERE("`PathIntegrator` expect `max_path_length` to be positive, get", max_path_length);
  if (!sampler_.is<BlueSobolSampler>() && !sampler_.is<SobolSampler>() && !sampler_.is<UniformSampler>() &&
      !sampler_.is<HaltonSampler>())
    SEVERE("[CudaPathIntegrator]only BlueSampler, SobolSampler, HaltonSampler and UniformSampler run on the CUDA path");
}
CudaPathIntegrator::CudaPathIntegrator(Sampler sampler, int max_path_length)
    : CudaPathIntegrator(Accel(), MOVE(sampler), LightSampler(UniformLightSampler()), max_path_length) {}
CudaPathIntegrator::CudaPathIntegrator(CudaPathIntegrator&& o) noexcept
    : sampler_(MOVE(o.sampler_)), max_path_length_(o.max_path_length_), devices_(MOVE(o.devices_)),
      shard_mode_(o.shard_mode_), n_devices_(o.n_devices_), multi_(o.multi_) {
  o.multi_ = nullptr;
}
CudaPathIntegrator::~CudaPathIntegrator() {
  if (multi_) pb2_destroy_multi(multi_);
}

void CudaPathIntegrator::render(Scene& scene) {
  // RTIntegrator::render prologue (integrator.cpp:26-35): spp from the sampler, accel build, light list
  const int spp = sampler_.spp();
  if (!multi_) {
    check(pb2_create_multi(devices_.size() ? &devices_[0] : nullptr, int(devices_.size()), &multi_), "pb2_create_multi");
    n_devices_ = pb2_multi_size(multi_);
  }
  Flat flat;
  flatten(scene, flat);
  check(pb2_multi_upload_scene(multi_, &flat.desc), "pb2_multi_upload_scene");
  const int kind = sampler_.is<SobolSampler>()     ? PB2_SAMPLER_SOBOL
                   : sampler_.is<UniformSampler>() ? PB2_SAMPLER_UNIFORM
                   : sampler_.is<HaltonSampler>()  ? PB2_SAMPLER_HALTON
                                                   : PB2_SAMPLER_BLUE;
  if (kind == PB2_SAMPLER_BLUE) {
    const char* tables = std::getenv("PINE_B200_TABLES");
    check(pb2_multi_load_sampler_tables_file(multi_, tables ? tables : "", spp), "pb2_multi_load_sampler_tables_file");
  }
  set_progress(0);
  // path.cpp:39 publishes progress from the pixel loop; here every completed device wave does (driver threads;
  // set_progress is an atomic store, integrator.cpp:11-16)
  check(pb2_multi_set_progress_callback(multi_, [](float f, void*) { set_progress(f); }, nullptr),
        "pb2_multi_set_progress_callback");
  pb2_render_params p{};
  p.spp = spp;
  p.max_path_length = max_path_length_;
  p.row_begin = 0; p.row_end = flat.desc.camera.height;
  p.sample_begin = 0; p.sample_end = spp;
  p.sampler = kind;  // UniformSampler keeps its seed private (sampler.h:36-38): the CUDA stream uses seed 0
  auto& film = scene.camera.film();
  // Film::pixels is Array2d<vec4>, row-major p.x + p.y*w (array.h:51-55): the ABI's film layout
  check(pb2_multi_render(multi_, &p, shard_mode_, &film.data()[0][0]), "pb2_multi_render");
  set_progress(1);
  pb2_counters c{};
  pb2_multi_get_counters(multi_, &c, nullptr, nullptr);
  last_closest_ = c.last_closest_rays;
  last_shadow_ = c.last_shadow_rays;
  launches_ = c.kernel_launches;
  last_ms_ = c.last_render_ms;
}

void CudaPathIntegrator::finalize_rgba8(const Film& film, unsigned char* rgba8, bool flip_y) {
  if (!multi_) SEVERE("[CudaPathIntegrator]finalize_rgba8 needs a render() first");
  // Film keeps its ToneMapper private; a copy carries it. Shrink the copy to one probe pixel, let the reference
  // finalize it, and see which operator it applied.
  Film probe = film;
  probe.pixels = Array2d<vec4>(vec2i(1, 1));
  probe[vec2i(0, 0)] = vec4(0.5f, 1.0f, 4.0f, 1.0f);
  probe.finalize();
  const vec3 u2 = Uncharted2ToneMapper()(vec3(0.5f, 1.0f, 4.0f));
  const vec4 got = probe[vec2i(0, 0)];
  const int tonemapper = (got.x == u2.x && got.y == u2.y && got.z == u2.z) ? 0 : 1;
  check(pb2_multi_film_finalize(multi_, tonemapper, flip_y ? 1 : 0, rgba8, nullptr), "pb2_multi_film_finalize");
}

}  // namespace pine
