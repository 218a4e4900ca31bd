// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// C-ABI harness around the UNMODIFIED reference (wicstas/pine) so that pytest (ctypes) and
// bench.py's cpu_baseline / --impl reference legs can drive Pine's own C++ API:
//   * build a pine::Scene from the same plain-number description our own host layer takes,
//   * run PathIntegrator(BVH(), <sampler>, UniformLightSampler(), depth).render(scene)
//     (src/pine/impl/integrator/path.cpp:26-41) and hand back Film::pixels pre-tonemap,
//   * trace a fixed ray set through pine::BVH and report (geom, prim, t) per ray -- the
//     recording lambda is the one of BVH::intersect (src/pine/impl/accel/bvh.cpp:518-536),
//   * dump the reference's binary SAH trees node for node (to pin our own builder),
//   * known-answer generators for the sampler / RNG / camera / shape constructors.
//
// `private` is opened ONLY around bvh.h so BVH::tbvh/lbvh/indices and BVHImpl::nodes are visible;
// bvh.cpp is #included (the reference's own test does the same with a .cpp:
// test/path_opengl_test.cpp:1) because BVHImpl::Intersect/hit are templates defined there.
#define private public
#include <pine/impl/accel/bvh.h>
#undef private
// RTIntegrator::samplers/spp/accel are protected; BoundedPath below derives from PathIntegrator and
// needs PathIntegrator::Vertex, which is only complete inside path.cpp -- so path.cpp is #included
// here too (unmodified) instead of being linked as its own object.
#include <pine/core/scene.h>
#include <pine/core/parallel.h>
#include <pine/core/integrator.h>
#include <pine/core/lightsampler.h>
#include <pine/core/sampler.h>
#include <pine/core/fileio.h>
#include <pine/core/profiler.h>
#include <pine/core/rng.h>
#include <pine/impl/integrator/path.h>
#include <pine/impl/accel/embree.h>

#include <pine/impl/accel/bvh.cpp>
#include <pine/impl/integrator/path.cpp>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>

extern float bluenoise_1spp(int, int, int, int);
extern float bluenoise_2spp(int, int, int, int);
extern float bluenoise_4spp(int, int, int, int);
extern float bluenoise_8spp(int, int, int, int);
extern float bluenoise_16spp(int, int, int, int);
extern float bluenoise_32spp(int, int, int, int);
extern float bluenoise_64spp(int, int, int, int);
extern float bluenoise_128spp(int, int, int, int);
extern float bluenoise_256spp(int, int, int, int);

using namespace pine;

namespace {
// PathIntegrator whose prologue (RTIntegrator::render: sampler clones, accel.build,
// light_sampler.build -- integrator.cpp:26-35) and pixel loop (path.cpp:31-40) can be run
// separately, so that bench.py can time a BOUNDED SAMPLE of a big render (a subset of the film's
// rows) without paying the reference's single-threaded BVH build once per timed step. The loop body
// is the reference's, verbatim in meaning; radiance() is the reference's own code.
struct BoundedPath : PathIntegrator {
  using PathIntegrator::PathIntegrator;
  void prepare(Scene& scene) { RTIntegrator::render(scene); }
  // rows row_begin, row_begin + row_step, ... < row_end of the film
  void render_rows(Scene& scene, int row_begin, int row_end, int row_step) {
    auto& film = scene.camera.film();
    const int n_rows = row_end > row_begin ? (row_end - row_begin + row_step - 1) / row_step : 0;
    parallel_for(vec2i(film.size().x, n_rows), [&](vec2i q) {
      vec2i p(q.x, row_begin + q.y * row_step);
      Sampler& sampler = samplers[threadIdx].start_pixel(p, 0);
      auto L = vec3(0.0f);
      for (int si = 0; si < spp; si++, sampler.start_next_sample()) {
        auto ray = scene.camera.gen_ray((p + sampler.rand2f()) / film.size(), sampler.rand2f());
        L += radiance(scene, ray, sampler, Vertex::first_vertex()).Lo;
      }
      film[p] = vec4(L / spp, 1.0f);
    });
  }
  // Samples [s0, s1) of every pixel, the sampler started at sample index s0 (Sampler::start_pixel(p, sample_index),
  // sampler.h:286-290: RNG(hash(p, s0)), BlueSobol index = s0): what one rank of a sample-sharded render owns
  // (SURVEY 8(e)). film[p] = (sum over the range / spp, (s1 - s0) / spp), so that the ranks' films ADD to a full one.
  void render_sample_range(Scene& scene, int s0, int s1) {
    auto& film = scene.camera.film();
    parallel_for(film.size(), [&](vec2i p) {
      Sampler& sampler = samplers[threadIdx].start_pixel(p, s0);
      auto L = vec3(0.0f);
      for (int si = s0; si < s1; si++, sampler.start_next_sample()) {
        auto ray = scene.camera.gen_ray((p + sampler.rand2f()) / film.size(), sampler.rand2f());
        L += radiance(scene, ray, sampler, Vertex::first_vertex()).Lo;
      }
      film[p] = vec4(L / spp, float(s1 - s0) / float(spp));
    });
  }
  // The same loop over the whole film, also returning per pixel and channel the sum and the sum of squares
  // of the spp per-sample radiances (what the 3-sigma test of SURVEY 8(c) T4 takes its sigma from).
  void render_moments(Scene& scene, double* sum, double* sumsq) {
    auto& film = scene.camera.film();
    const int w = film.size().x;
    parallel_for(film.size(), [&](vec2i p) {
      Sampler& sampler = samplers[threadIdx].start_pixel(p, 0);
      auto L = vec3(0.0f);
      double* s1 = sum + 3 * (size_t(p.x) + size_t(p.y) * w);
      double* s2 = sumsq + 3 * (size_t(p.x) + size_t(p.y) * w);
      for (int k = 0; k < 3; k++) s1[k] = s2[k] = 0;
      for (int si = 0; si < spp; si++, sampler.start_next_sample()) {
        auto ray = scene.camera.gen_ray((p + sampler.rand2f()) / film.size(), sampler.rand2f());
        auto Lo = radiance(scene, ray, sampler, Vertex::first_vertex()).Lo;
        L += Lo;
        for (int k = 0; k < 3; k++) { s1[k] += double(Lo[k]); s2[k] += double(Lo[k]) * double(Lo[k]); }
      }
      film[p] = vec4(L / spp, 1.0f);
    });
  }
};
}  // namespace
struct RefScene {
  Scene scene;
  std::vector<psl::shared_ptr<Material>> materials;
  BVH bvh;
  bool bvh_built = false;
  BoundedPath* bounded = nullptr;
  ~RefScene() { delete bounded; }
};
namespace {
vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }
mat4 m4_rowmajor(const float* m) {
  return mat4(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], m[12],
              m[13], m[14], m[15]);
}
void store_m4_rowmajor(const mat4& m, float* out) {
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) out[r * 4 + c] = m[c][r];
}
}  // namespace

extern "C" {

void* ref_scene_new() { return new RefScene(); }
void ref_scene_free(void* h) { delete (RefScene*)h; }
// the live pine::Scene of a handle, for oracle/shim_harness.cpp (the drop-in test lives in its own library)
void* ref_scene_ptr(void* h) { return &((RefScene*)h)->scene; }

// kind: 0 Emissive(color) 1 Diffuse(albedo) 2 Metal(albedo, rough) 3 Glossy(albedo, rough, ior)
//       4 Glass(albedo, rough, ior) 5 Uber(albedo, rough, metallic, transmission, ior)
//       6 Subsurface(albedo, rough, sigma_s[3])
// p = {c0,c1,c2, roughness, ior, metallic, transmission, s0,s1,s2}
int ref_add_material(void* h, int kind, const float* p) {
  auto& S = *(RefScene*)h;
  auto c = v3(p);
  psl::shared_ptr<Material> m;
  switch (kind) {
    case 0: m = psl::make_shared<Material>(EmissiveMaterial(c)); break;
    case 1: m = psl::make_shared<Material>(DiffuseMaterial(c)); break;
    case 2: m = psl::make_shared<Material>(MetalMaterial(c, p[3])); break;
    case 3: m = psl::make_shared<Material>(GlossyMaterial(c, p[3], p[4])); break;
    case 4: m = psl::make_shared<Material>(GlassMaterial(c, p[3], p[4])); break;
    case 5: m = psl::make_shared<Material>(UberMaterial(c, p[3], p[5], p[6], p[4])); break;
    case 6: m = psl::make_shared<Material>(SubsurfaceMaterial(c, p[3], v3(p + 7))); break;
    default: return -1;
  }
  S.materials.push_back(m);
  return int(S.materials.size()) - 1;
}

}  // extern "C"
namespace {
// A typed postfix program of pine_b200/nodes.py (op | 0x100 = the instruction yields a Nodef) rebuilt as the
// reference's OWN node objects (node.h:105-250), so that the reference evaluates the graph with its own code.
struct NodeValue {
  bool is_float;
  psl::optional<Nodef> f;
  psl::optional<Node3f> v;
};
Node3f to3(const NodeValue& x) { return x.is_float ? Node3f(NodeToVec3(*x.f)) : *x.v; }
template <char op>
NodeValue binary(const NodeValue& a, const NodeValue& b, bool is_float) {
  if (is_float) return NodeValue{true, Nodef(NodeBinary<float, op>(*a.f, *b.f)), psl::nullopt};
  return NodeValue{false, psl::nullopt, Node3f(NodeBinary<vec3, op>(to3(a), to3(b)))};
}
template <char op>
NodeValue unary(const NodeValue& x, bool is_float) {
  if (is_float) return NodeValue{true, Nodef(NodeUnary<float, op>(*x.f)), psl::nullopt};
  return NodeValue{false, psl::nullopt, Node3f(NodeUnary<vec3, op>(*x.v))};
}
// returns false on a malformed program or an op the harness does not build (image nodes: the glTF tests cover them)
bool build_node(const int* ops, const float* args, int n, int start, NodeValue& out) {
  std::vector<NodeValue> st;
  for (int pc = start - 1; pc < n; pc++) {
    const int op = ops[pc] & 0xff;
    const bool is_float = (ops[pc] & 0x100) != 0;
    const float* a = args + 3 * pc;
    auto pop = [&]() { NodeValue x = st.back(); st.pop_back(); return x; };
    if (op == 0) {
      if (st.size() != 1) return false;
      out = st.back();
      return true;
    }
    const size_t need = (op >= 5 && op <= 9) ? 2 : (op == 16 ? 3 : (op >= 10 ? 1 : 0));
    if (st.size() < need) return false;
    switch (op) {
      case 1:
        if (is_float) st.push_back(NodeValue{true, Nodef(a[0]), psl::nullopt});
        else st.push_back(NodeValue{false, psl::nullopt, Node3f(vec3(a[0], a[1], a[2]))});
        break;
      case 2: st.push_back(NodeValue{false, psl::nullopt, Node3f(NodePosition())}); break;
      case 3: st.push_back(NodeValue{false, psl::nullopt, Node3f(NodeNormal())}); break;
      case 4: st.push_back(NodeValue{false, psl::nullopt, Node3f(NodeUV())}); break;
      case 5: { auto b = pop(), x = pop(); st.push_back(binary<'+'>(x, b, is_float)); break; }
      case 6: { auto b = pop(), x = pop(); st.push_back(binary<'-'>(x, b, is_float)); break; }
      case 7: { auto b = pop(), x = pop(); st.push_back(binary<'*'>(x, b, is_float)); break; }
      case 8: { auto b = pop(), x = pop(); st.push_back(binary<'/'>(x, b, is_float)); break; }
      case 9: { auto b = pop(), x = pop(); st.push_back(binary<'^'>(x, b, is_float)); break; }
      case 10: { auto x = pop(); st.push_back(unary<'-'>(x, is_float)); break; }
      case 11: { auto x = pop(); st.push_back(unary<'a'>(x, is_float)); break; }
      case 12: { auto x = pop(); st.push_back(unary<'s'>(x, is_float)); break; }
      case 13: { auto x = pop(); st.push_back(unary<'r'>(x, is_float)); break; }
      case 14: { auto x = pop(); st.push_back(unary<'f'>(x, is_float)); break; }
      case 15: { auto x = pop(); st.push_back(NodeValue{true, Nodef(NodeComponent(to3(x), int(a[0]))), psl::nullopt}); break; }
      case 16: {
        auto z = pop(), y = pop(), x = pop();
        if (!x.is_float || !y.is_float || !z.is_float) return false;
        st.push_back(NodeValue{false, psl::nullopt, Node3f(NodeToVec3(*x.f, *y.f, *z.f))});
        break;
      }
      case 17: { auto x = pop(); st.push_back(NodeValue{true, Nodef(NodeCheckerboard(to3(x), a[0])), psl::nullopt}); break; }
      default: return false;
    }
  }
  return false;
}
}  // namespace
extern "C" {

// ref_add_material with graph parameters. starts[5]: 1-based program start per slot {color, roughness, ior,
// metallic, transmission}, 0 = the constant of `p`. -1: unknown kind, -2: a program the harness cannot build.
int ref_add_material_nodes(void* h, int kind, const float* p, const int* ops, const float* args, int n,
                           const int* starts) {
  auto& S = *(RefScene*)h;
  Node3f color(v3(p));
  Nodef rough(p[3]), ior(p[4]), metallic(p[5]), transmission(p[6]);
  for (int slot = 0; slot < 5; slot++) {
    if (!starts[slot]) continue;
    NodeValue x;
    if (!build_node(ops, args, n, starts[slot], x)) return -2;
    if (slot == 0) color = to3(x);
    else {
      if (!x.is_float) return -2;
      (slot == 1 ? rough : slot == 2 ? ior : slot == 3 ? metallic : transmission) = *x.f;
    }
  }
  psl::shared_ptr<Material> m;
  switch (kind) {
    case 1: m = psl::make_shared<Material>(DiffuseMaterial(color)); break;
    case 2: m = psl::make_shared<Material>(MetalMaterial(color, rough)); break;
    case 3: m = psl::make_shared<Material>(GlossyMaterial(color, rough, ior)); break;
    case 4: m = psl::make_shared<Material>(GlassMaterial(color, rough, ior)); break;
    case 5: m = psl::make_shared<Material>(UberMaterial(color, rough, metallic, transmission, p[4])); break;
    case 6: m = psl::make_shared<Material>(SubsurfaceMaterial(color, rough, v3(p + 7))); break;
    default: return -1;
  }
  S.materials.push_back(m);
  return int(S.materials.size()) - 1;
}
// Mnode<T>::operator()(NodeEvalCtx{p, n, uv}) of a program at n_pts points (ctx8: p, n, uv; out3: vec3, a Nodef's
// value in all three components): the known answers for pb2_test_eval_node.
int ref_eval_node(const int* ops, const float* args, int n, int start, int n_pts, const float* ctx8, float* out3) {
  NodeValue x;
  if (!build_node(ops, args, n, start, x)) return -2;
  for (int i = 0; i < n_pts; i++) {
    const float* c = ctx8 + 8 * i;
    const NodeEvalCtx ec(v3(c), v3(c + 3), vec2(c[6], c[7]));
    const vec3 v = x.is_float ? vec3((*x.f)(ec)) : (*x.v)(ec);
    out3[3 * i] = v.x; out3[3 * i + 1] = v.y; out3[3 * i + 2] = v.z;
  }
  return 0;
}

static int add_shape(RefScene& S, Shape shape, int mat) {
  S.scene.add_geometry(MOVE(shape), S.materials[mat]);
  S.bvh_built = false;
  return int(S.scene.geometries.size()) - 1;
}
int ref_add_rect(void* h, const float* pos, const float* ex, const float* ey, int flip, int mat) {
  return add_shape(*(RefScene*)h, Rect(v3(pos), v3(ex), v3(ey), flip != 0), mat);
}
int ref_add_box(void* h, const float* lo, const float* hi, const float* m_rowmajor, int mat) {
  return add_shape(*(RefScene*)h, OBB(AABB(v3(lo), v3(hi)), m4_rowmajor(m_rowmajor)), mat);
}
int ref_add_aabb(void* h, const float* lo, const float* hi, int mat) {
  return add_shape(*(RefScene*)h, AABB(v3(lo), v3(hi)), mat);
}
int ref_add_cone(void* h, const float* p, const float* n, float r, float hgt, int mat) {
  return add_shape(*(RefScene*)h, Cone(v3(p), v3(n), r, hgt), mat);
}
int ref_add_sphere(void* h, const float* c, float r, int mat) {
  return add_shape(*(RefScene*)h, Sphere(v3(c), r), mat);
}
int ref_add_disk(void* h, const float* p, const float* n, float r, int mat) {
  return add_shape(*(RefScene*)h, Disk(v3(p), v3(n), r), mat);
}
int ref_add_plane(void* h, const float* p, const float* n, int mat) {
  return add_shape(*(RefScene*)h, Plane(v3(p), v3(n)), mat);
}
int ref_add_cylinder(void* h, const float* p0, const float* p1, float r, int mat) {
  return add_shape(*(RefScene*)h, Cylinder(v3(p0), v3(p1), r), mat);
}
int ref_add_triangle(void* h, const float* a, const float* b, const float* c, int mat) {
  return add_shape(*(RefScene*)h, Triangle(v3(a), v3(b), v3(c)), mat);
}
int ref_add_mesh(void* h, const float* verts, int nv, const uint32_t* idx, int nt,
                 const float* normals, const float* uvs, int mat) {
  psl::vector<vec3> V(nv);
  for (int i = 0; i < nv; i++) V[i] = v3(verts + 3 * i);
  psl::vector<vec3u32> I(nt);
  for (int i = 0; i < nt; i++) I[i] = vec3u32(idx[3 * i], idx[3 * i + 1], idx[3 * i + 2]);
  psl::vector<vec3> N;
  if (normals) {
    N.resize(nv);
    for (int i = 0; i < nv; i++) N[i] = v3(normals + 3 * i);
  }
  psl::vector<vec2> T;
  if (uvs) {
    T.resize(nv);
    for (int i = 0; i < nv; i++) T[i] = vec2(uvs[2 * i], uvs[2 * i + 1]);
  }
  return add_shape(*(RefScene*)h, Mesh(MOVE(V), MOVE(I), MOVE(T), MOVE(N)), mat);
}
// tonemapper: 0 Uncharted2, 1 ACES
void ref_set_camera(void* h, int w, int hgt, int tonemapper, const float* from, const float* to,
                    float fov, float len_radius, float focus) {
  auto& S = *(RefScene*)h;
  auto tm = tonemapper ? ToneMapper(ACESToneMapper()) : ToneMapper(Uncharted2ToneMapper());
  S.scene.set_camera(ThinLenCamera(Film(vec2i(w, hgt), tm), v3(from), v3(to), fov, len_radius,
                                   focus));
}
// Scene::add_light / set_env_light (scene.cpp:29-31,44-46) for the non-area lights (light.h:21-66)
void ref_add_point_light(void* h, const float* position, const float* color) {
  ((RefScene*)h)->scene.add_light(PointLight(v3(position), v3(color)));
}
void ref_add_directional_light(void* h, const float* direction, const float* color) {
  ((RefScene*)h)->scene.add_light(DirectionalLight(v3(direction), v3(color)));
}
void ref_add_spot_light(void* h, const float* position, const float* direction, const float* color,
                        float falloff_radian, float cutoff_additional_radian) {
  ((RefScene*)h)->scene.add_light(
      SpotLight(v3(position), v3(direction), v3(color), falloff_radian, cutoff_additional_radian));
}
void ref_set_sky(void* h, const float* sun_color) { ((RefScene*)h)->scene.set_env_light(Sky(v3(sun_color))); }
// Film size of the scene's camera (a glTF camera node sets it inside scene_from_gltf, fileio.cpp:308-321)
void ref_camera_size(void* h, int* wh) {
  auto size = ((RefScene*)h)->scene.camera.film().size();
  wh[0] = size.x;
  wh[1] = size.y;
}
int ref_num_geometries(void* h) { return int(((RefScene*)h)->scene.geometries.size()); }
int ref_num_lights(void* h) { return int(((RefScene*)h)->scene.lights.size()); }
int ref_num_threads() { return n_threads(); }

// ---- render ----------------------------------------------------------------------------------
// sampler_kind: 0 BlueSobolSampler (PRL "BlueSampler"), 1 SobolSampler, 2 UniformSampler.
// Returns wall seconds around render() (includes BVH::build, as the reference's own flow does,
// integrator.cpp:26-35). film_rgba receives Film::pixels (w*h*4 floats, row-major, pre-tonemap).
double ref_render(void* h, int sampler_kind, int spp, int depth, float* film_rgba) {
  auto& S = *(RefScene*)h;
  auto sampler = sampler_kind == 0   ? Sampler(BlueSobolSampler(spp))
                 : sampler_kind == 1 ? Sampler(SobolSampler(spp))
                                     : sampler_kind == 3 ? Sampler(HaltonSampler(spp)) : Sampler(UniformSampler(spp));
  auto integrator = PathIntegrator(Accel(BVH()), MOVE(sampler), LightSampler(UniformLightSampler()),
                                   depth);
  auto t0 = std::chrono::steady_clock::now();
  integrator.render(S.scene);
  auto t1 = std::chrono::steady_clock::now();
  if (film_rgba) {
    auto& film = S.scene.camera.film();
    std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
// Bounded-sample variant for timing (see BoundedPath): prepare once (returns the seconds the
// prologue took, dominated by BVH::build), then render row subsets. BlueSobolSampler only.
// 1 when this library carries the reference's EmbreeAccel (libpine_ref_embree.so, vendored Embree 4.3.1), else 0
int ref_has_embree() {
#ifdef PINE_REF_WITH_EMBREE
  return 1;
#else
  return 0;
#endif
}
static int g_bounded_accel = 0;  // 0: pine::BVH (the bit-exact oracle), 1: EmbreeAccel (the PRL default,
                                 // program_context.cpp:79-81) -- timing baseline only
void ref_set_bounded_accel(int accel) { g_bounded_accel = accel; }
double ref_bounded_prepare(void* h, int spp, int depth) {
  auto& S = *(RefScene*)h;
  delete S.bounded;
  S.bounded = new BoundedPath(g_bounded_accel == 1 ? Accel(EmbreeAccel()) : Accel(BVH()), Sampler(BlueSobolSampler(spp)),
                              LightSampler(UniformLightSampler()), depth);
  auto t0 = std::chrono::steady_clock::now();
  S.bounded->prepare(S.scene);
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
double ref_bounded_render_rows(void* h, int row_begin, int row_end, int row_step, float* film_rgba) {
  auto& S = *(RefScene*)h;
  if (!S.bounded) return -1.0;
  auto t0 = std::chrono::steady_clock::now();
  S.bounded->render_rows(S.scene, row_begin, row_end, row_step);
  auto t1 = std::chrono::steady_clock::now();
  if (film_rgba) {
    auto& film = S.scene.camera.film();
    std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
// One rank's share of a sample-sharded render (see BoundedPath::render_sample_range). BlueSobolSampler.
double ref_render_sample_range(void* h, int spp, int depth, int s0, int s1, float* film_rgba) {
  auto& S = *(RefScene*)h;
  BoundedPath path(Accel(BVH()), Sampler(BlueSobolSampler(spp)), LightSampler(UniformLightSampler()), depth);
  auto t0 = std::chrono::steady_clock::now();
  path.prepare(S.scene);
  path.render_sample_range(S.scene, s0, s1);
  auto t1 = std::chrono::steady_clock::now();
  auto& film = S.scene.camera.film();
  std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  return std::chrono::duration<double>(t1 - t0).count();
}
// Whole-film render that also reports the per-pixel moments of the per-sample radiance (sum, sumsq:
// w*h*3 doubles each). BlueSobolSampler; film_rgba as ref_render.
double ref_render_moments(void* h, int spp, int depth, float* film_rgba, double* sum, double* sumsq) {
  auto& S = *(RefScene*)h;
  BoundedPath path(Accel(BVH()), Sampler(BlueSobolSampler(spp)), LightSampler(UniformLightSampler()), depth);
  auto t0 = std::chrono::steady_clock::now();
  path.prepare(S.scene);
  path.render_moments(S.scene, sum, sumsq);
  auto t1 = std::chrono::steady_clock::now();
  if (film_rgba) {
    auto& film = S.scene.camera.film();
    std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
// Film::finalize (film.cpp:21-27) + to_uint8 (fileio.cpp:42-54) on a caller-supplied float film;
// used to pin our tone-map/quantise epilogue. out = w*h*4 u8, y-flipped like save_film_as_image.
void ref_film_finalize(void* h, const float* film_rgba, float* out_tonemapped) {
  auto& S = *(RefScene*)h;
  Film film = S.scene.camera.film();
  std::memcpy(film.data(), film_rgba, sizeof(float) * 4 * size_t(film.width()) * film.height());
  film.finalize();
  std::memcpy(out_tonemapped, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
}

// ---- accel: build / trace / dump ----------------------------------------------------------------
double ref_accel_build(void* h) {
  auto& S = *(RefScene*)h;
  auto t0 = std::chrono::steady_clock::now();
  S.bvh = BVH();
  S.bvh.build(&S.scene);
  S.bvh_built = true;
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
// rays: n x {ox,oy,oz, dx,dy,dz, tmin, tmax}. geom/prim = -1 on miss; t = ray.tmax after traversal.
// Optional surf (n x 8: p, n, uv) = compute_surface_info outputs (bvh.cpp:537-545).
// Optional n_prim_tests (per ray) = number of primitive intersect() calls the traversal made.
void ref_trace_closest(void* h, const float* rays, int64_t n, int32_t* geom, int32_t* prim,
                       float* t, float* surf, int32_t* n_prim_tests) {
  auto& S = *(RefScene*)h;
  if (!S.bvh_built) ref_accel_build(h);
  const BVH& bvh = S.bvh;
  const Scene* scene = &S.scene;
  parallel_for(n, [&](int64_t i) {
    const float* r = rays + 8 * i;
    Ray ray(vec3(r[0], r[1], r[2]), vec3(r[3], r[4], r[5]), r[6], r[7]);
    SurfaceInteraction it;
    uint32_t geom_index = 0, prim_index = 0;
    int tests = 0;
    bool hit = false;
    if (scene->geometries.size())
      hit = bvh.tbvh.Intersect(ray, [&](Ray& ray, int i_lbvh) {
        auto& geometry = scene->geometries[bvh.indices[i_lbvh]];
        if (i_lbvh < int(bvh.lbvh.size())) {
          auto& mesh = geometry->shape.as<Mesh>();
          auto hit = bvh.lbvh[i_lbvh].Intersect(ray, [&](Ray& ray, int index) {
            tests++;
            auto hit = mesh.intersect(ray, index);
            if (hit) prim_index = index;
            return hit;
          });
          if (hit) geom_index = bvh.indices[i_lbvh];
          return hit;
        } else {
          tests++;
          auto hit = geometry->intersect(ray, it);
          if (hit) { geom_index = bvh.indices[i_lbvh]; prim_index = 0; }
          return hit;
        }
      });
    geom[i] = hit ? int32_t(geom_index) : -1;
    prim[i] = hit ? int32_t(prim_index) : -1;
    t[i] = ray.tmax;
    if (n_prim_tests) n_prim_tests[i] = tests;
    if (surf) {
      float* s = surf + 8 * i;
      for (int k = 0; k < 8; k++) s[k] = 0;
      if (hit) {
        auto& shape = scene->geometries[geom_index]->shape;
        if (shape.is<Mesh>())
          shape.as<Mesh>().compute_surface_info(ray(), it, prim_index);
        else
          shape.compute_surface_info(ray(), it);
        s[0] = it.p.x; s[1] = it.p.y; s[2] = it.p.z;
        s[3] = it.n.x; s[4] = it.n.y; s[5] = it.n.z;
        s[6] = it.uv.x; s[7] = it.uv.y;
      }
    }
  });
}
void ref_trace_any(void* h, const float* rays, int64_t n, uint8_t* occluded) {
  auto& S = *(RefScene*)h;
  if (!S.bvh_built) ref_accel_build(h);
  parallel_for(n, [&](int64_t i) {
    const float* r = rays + 8 * i;
    occluded[i] = S.bvh.hit(Ray(vec3(r[0], r[1], r[2]), vec3(r[3], r[4], r[5]), r[6], r[7]));
  });
}
// which = -1: top-level tree; which = k >= 0: per-mesh tree k. Returns node count (0 if absent).
static const BVHImpl* pick(RefScene& S, int which) {
  if (which < 0) return &S.bvh.tbvh;
  if (which >= int(S.bvh.lbvh.size())) return nullptr;
  return &S.bvh.lbvh[which];
}
int ref_bvh_num_lbvh(void* h) { return int(((RefScene*)h)->bvh.lbvh.size()); }
int ref_bvh_num_indices(void* h) { return int(((RefScene*)h)->bvh.indices.size()); }
void ref_bvh_indices(void* h, int32_t* out) {
  auto& v = ((RefScene*)h)->bvh.indices;
  for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
}
int ref_bvh_info(void* h, int which, int32_t* root, int32_t* n_leaf_prims) {
  auto* b = pick(*(RefScene*)h, which);
  if (!b) return 0;
  *root = b->rootIndex;
  int np = 0;
  for (auto& nd : b->nodes) np += int(nd.primitiveIndices.size());
  *n_leaf_prims = np;
  return int(b->nodes.size());
}
// aabbs: n x 12 (aabbs[0].lower, upper, aabbs[1].lower, upper); children: n x 2;
// leaf_range: n x 2 (offset, count) into leaf_prims.
void ref_bvh_nodes(void* h, int which, float* aabbs, int32_t* children, int32_t* leaf_range,
                   int32_t* leaf_prims) {
  auto* b = pick(*(RefScene*)h, which);
  if (!b) return;
  int off = 0;
  for (size_t i = 0; i < b->nodes.size(); i++) {
    auto& nd = b->nodes[i];
    for (int c = 0; c < 2; c++)
      for (int k = 0; k < 3; k++) {
        aabbs[i * 12 + c * 6 + k] = nd.aabbs[c].lower[k];
        aabbs[i * 12 + c * 6 + 3 + k] = nd.aabbs[c].upper[k];
      }
    children[i * 2] = nd.children[0];
    children[i * 2 + 1] = nd.children[1];
    leaf_range[i * 2] = off;
    leaf_range[i * 2 + 1] = int(nd.primitiveIndices.size());
    for (int p : nd.primitiveIndices) leaf_prims[off++] = p;
  }
}

// ---- known-answer generators ---------------------------------------------------------------------
float ref_bluenoise(int spp, int px, int py, int index, int dim) {
  switch (spp) {
    case 1: return bluenoise_1spp(px, py, index, dim);
    case 2: return bluenoise_2spp(px, py, index, dim);
    case 4: return bluenoise_4spp(px, py, index, dim);
    case 8: return bluenoise_8spp(px, py, index, dim);
    case 16: return bluenoise_16spp(px, py, index, dim);
    case 32: return bluenoise_32spp(px, py, index, dim);
    case 64: return bluenoise_64spp(px, py, index, dim);
    case 128: return bluenoise_128spp(px, py, index, dim);
    case 256: return bluenoise_256spp(px, py, index, dim);
  }
  return -1.0f;
}
// Sampler wrapper stream (sampler.h:286-309): start_pixel(p, sample_index) then `n_rand` randf()
// draws into out_rand, and per sample (n_samples of them) a draw pattern string of '1' (get1d) and
// '2' (get2d) into out_draws (sum of widths floats per sample), calling start_next_sample between.
int ref_sampler_spp(int spp_requested) { return BlueSobolSampler(spp_requested).spp(); }
void ref_sampler_stream(int spp, int px, int py, int sample_index, int n_rand, float* out_rand,
                        const char* pattern, int n_samples, float* out_draws) {
  Sampler sampler = BlueSobolSampler(spp);
  sampler.init(vec2i(1024, 1024));
  sampler.start_pixel(vec2i(px, py), sample_index);
  for (int i = 0; i < n_rand; i++) out_rand[i] = sampler.randf();
  for (int s = 0; s < n_samples; s++, sampler.start_next_sample())
    for (const char* c = pattern; *c; c++) {
      if (*c == '1') {
        *out_draws++ = sampler.get1d();
      } else {
        auto u = sampler.get2d();
        *out_draws++ = u.x;
        *out_draws++ = u.y;
      }
    }
}
// The same stream for the other Sampler alternatives (sampler_kind as in ref_render): init(film
// size) matters for SobolSampler (nbase4_digits, sampler.cpp:82-85).
void ref_sampler_stream_kind(int sampler_kind, int spp, int width, int height, int px, int py,
                             int sample_index, const char* pattern, int n_samples, float* out_draws) {
  Sampler sampler = sampler_kind == 0   ? Sampler(BlueSobolSampler(spp))
                    : sampler_kind == 1 ? Sampler(SobolSampler(spp))
                                        : sampler_kind == 3 ? Sampler(HaltonSampler(spp)) : Sampler(UniformSampler(spp));
  sampler.init(vec2i(width, height));
  sampler.start_pixel(vec2i(px, py), sample_index);
  for (int s = 0; s < n_samples; s++, sampler.start_next_sample())
    for (const char* c = pattern; *c; c++) {
      if (*c == '1') {
        *out_draws++ = sampler.get1d();
      } else {
        auto u = sampler.get2d();
        *out_draws++ = u.x;
        *out_draws++ = u.y;
      }
    }
}
uint64_t ref_hash_pixel(int px, int py, int sample_index) { return hash(vec2i(px, py), sample_index); }
void ref_rng_u64(uint64_t seed, int n, uint64_t* out) {
  RNG rng(seed);
  for (int i = 0; i < n; i++) out[i] = rng.next64u();
}
// Order of the two draws inside LightSampler::sample (lightsampler.h:27) as compiled here:
// returns 1 if get2d() is evaluated before get1d() (g++ right-to-left), else 0.
int ref_nee_draw_order_2d_first() {
  Sampler sampler = BlueSobolSampler(16);
  sampler.start_pixel(vec2i(3, 5), 0);
  Sampler probe = sampler;
  float a0 = probe.get1d();
  // replicate the call expression shape: f(get1d(), get2d())
  struct L { static void f(float u1, vec2 u2, float* o) { o[0] = u1; o[1] = u2.x; o[2] = u2.y; } };
  float o[3];
  L::f(sampler.get1d(), sampler.get2d(), o);
  return o[0] == a0 ? 0 : 1;
}
// Camera::gen_ray (camera.cpp:22-33) for n (p_film, u2) pairs -> rays n x 8.
void ref_gen_rays(void* h, const float* pfilm_u2, int64_t n, float* rays) {
  auto& S = *(RefScene*)h;
  for (int64_t i = 0; i < n; i++) {
    const float* a = pfilm_u2 + 4 * i;
    Ray r = S.scene.camera.gen_ray(vec2(a[0], a[1]), vec2(a[2], a[3]));
    float* o = rays + 8 * i;
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
    o[6] = r.tmin; o[7] = r.tmax;
  }
}
// The camera rays PathIntegrator::render generates (path.cpp:32-35) for sample `si` of every pixel
// of a w x h film, assuming no other rand*() draws in between (true for non-Uber/BSSRDF scenes).
void ref_primary_rays(void* h, int si, float* rays) {
  auto& S = *(RefScene*)h;
  auto size = S.scene.camera.film().size();
  parallel_for(size, [&](vec2i p) {
    Sampler sampler = BlueSobolSampler(1);
    sampler.start_pixel(p, 0);
    Ray r;
    for (int s = 0; s <= si; s++)
      r = S.scene.camera.gen_ray((p + sampler.rand2f()) / size, sampler.rand2f());
    float* o = rays + 8 * (size_t(p.x) + size_t(p.y) * size.x);
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
    o[6] = r.tmin; o[7] = r.tmax;
  });
}
// spawn_ray (interaction.cpp:6-13, ray.h:25-37): in n x {p3, n3, wo3, distance} -> rays n x 8
void ref_spawn_rays(const float* in, int64_t n, float* rays) {
  for (int64_t i = 0; i < n; i++) {
    const float* a = in + 10 * i;
    SurfaceInteraction it;
    it.p = v3(a); it.n = v3(a + 3);
    Ray r = it.spawn_ray(v3(a + 6), a[9]);
    float* o = rays + 8 * i;
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
    o[6] = r.tmin; o[7] = r.tmax;
  }
}
// Matrix helpers as PRL exposes them (vecmath.h:1088-1180); row-major 16 floats in/out.
// op: 0 translate(a) 1 scale(a) 2 rotate_x(a[0]) 3 rotate_y(a[0]) 4 rotate_z(a[0]) 5 inverse(m)
//     6 look_at(a[0..2], a[3..5])
void ref_mat4(int op, const float* a, float* out) {
  mat4 m;
  switch (op) {
    case 0: m = translate(v3(a)); break;
    case 1: m = scale(v3(a)); break;
    case 2: m = rotate_x(a[0]); break;
    case 3: m = rotate_y(a[0]); break;
    case 4: m = rotate_z(a[0]); break;
    case 5: m = inverse(m4_rowmajor(a)); break;
    case 6: m = look_at(v3(a), v3(a + 3)); break;
  }
  store_m4_rowmajor(m, out);
}
void ref_mat4_mul(const float* a, const float* b, float* out) {
  store_m4_rowmajor(m4_rowmajor(a) * m4_rowmajor(b), out);
}
// Geometry AABBs as BVH::build sees them (bvh.cpp:485) : out n_geom x 6 (mesh geometries: union).
void ref_geometry_aabbs(void* h, float* out) {
  auto& S = *(RefScene*)h;
  for (size_t i = 0; i < S.scene.geometries.size(); i++) {
    AABB b = S.scene.geometries[i]->get_aabb();
    for (int k = 0; k < 3; k++) { out[i * 6 + k] = b.lower[k]; out[i * 6 + 3 + k] = b.upper[k]; }
  }
}
// Shape::sample / Shape::pdf of geometry g (geometry.h:327-347): in n x {p3, u2, u1} ->
// out n x {valid, p3, n3, w3, distance, pdf}
void ref_shape_sample(void* h, int g, const float* in, int64_t n, float* out) {
  auto& S = *(RefScene*)h;
  for (int64_t i = 0; i < n; i++) {
    const float* a = in + 6 * i;
    float* o = out + 12 * i;
    for (int k = 0; k < 12; k++) o[k] = 0;
    if (auto ss = S.scene.geometries[g]->sample(v3(a), vec2(a[3], a[4]), a[5])) {
      o[0] = 1;
      o[1] = ss->p.x; o[2] = ss->p.y; o[3] = ss->p.z;
      o[4] = ss->n.x; o[5] = ss->n.y; o[6] = ss->n.z;
      o[7] = ss->w.x; o[8] = ss->w.y; o[9] = ss->w.z;
      o[10] = ss->distance; o[11] = ss->pdf;
    }
  }
}
// scene_from_gltf (fileio.cpp:146-327) into this scene; returns number of geometries added.
int ref_load_gltf(void* h, const char* filename) {
  auto& S = *(RefScene*)h;
  size_t before = S.scene.geometries.size();
  scene_from_gltf(S.scene, psl::string(filename));
  S.bvh_built = false;
  return int(S.scene.geometries.size() - before);
}
}  // extern "C"
