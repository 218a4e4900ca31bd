// scene_upload.cu -- pb2_upload_scene: validate the flattened scene, build the BVHs on the host
// (reference-identical binary trees, then the wide collapse) and lay everything out in HBM.
// Replaces accel.build(&scene) + light_sampler.build(&scene) (integrator.cpp:32-33).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "bvh_layout.h"
#include "context.h"

using namespace pb2;

namespace {

// Runs fn(0..n-1) on up to `threads` host threads (dynamic distribution).
template <class F>
void parallel_jobs(size_t n, int threads, F&& fn) {
  if (n == 0) return;
  std::atomic<size_t> next{0};
  auto work = [&] { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads && size_t(t) < n; t++) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
}

// Height of a binary tree in internal nodes along the longest root-to-leaf path: the deepest the reference-order
// walk's stack can get (it holds at most one deferred sibling per internal node on the current path). The
// builder numbers nodes in post-order, children before their parent.
int tree_height(const Tree2& t) {
  std::vector<int> h(t.nodes.size(), 0);
  int best = 0;
  for (size_t i = 0; i < t.nodes.size(); i++) {
    const Node2& nd = t.nodes[i];
    if (nd.leaf_count) continue;
    h[i] = 1 + std::max(h[nd.child[0]], h[nd.child[1]]);
    best = std::max(best, h[i]);
  }
  return best;
}
constexpr int kRefOrderStack = 64;  // traverse.cuh kStack

bool valid_material(const pb2_material& m) {
  return m.kind >= PB2_MAT_EMISSIVE && m.kind <= PB2_MAT_SUBSURFACE;
}

}  // namespace

static_assert(sizeof(pb2_material) == 76 && sizeof(pb2_node_instr) == 16, "pine_b200/_abi.py mirrors these layouts");
int restage_unpinned(pb2_context* ctx);  // first staging of a fresh scene: from pageable memory (below)

extern "C" int pb2_upload_scene(pb2_context* ctx, const pb2_scene_desc* sd) {
  if (!ctx || !sd) { set_error("null argument"); return PB2_E_INVALID; }
  if (ctx->device >= 0) PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  ctx->host_scene_ready = false;
  ctx->host_view = nullptr;
  auto t0 = std::chrono::steady_clock::now();
  const bool verbose = std::getenv("PB2_VERBOSE") != nullptr;
  auto stage = [&](const char* name) {
    if (verbose)
      fprintf(stderr, "[pb2_upload_scene] %-28s +%.1f ms\n", name,
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  };
  ctx->scene_ready = false;
  if (sd->camera.width <= 0 || sd->camera.height <= 0) {
    set_error("camera film size must be positive");
    return PB2_E_INVALID;
  }
  const uint32_t counts[10] = {sd->n_aabbs, sd->n_obbs, sd->n_spheres, sd->n_triangles, sd->n_cones,
                               sd->n_rects, sd->n_disks, sd->n_meshes, sd->n_planes, sd->n_cylinders};
  for (uint32_t i = 0; i < sd->n_materials; i++) {
    const pb2_material& m = sd->materials[i];
    if (!valid_material(m)) { set_error("unknown material kind"); return PB2_E_UNSUPPORTED; }
    const int32_t tex[3] = {m.color_texture, m.roughness_texture, m.metallic_texture};
    bool tex_ok = true;
    for (int32_t t : tex) tex_ok = tex_ok && t >= 0 && uint32_t(t) <= sd->n_textures;
    if (!tex_ok) {
      set_error("material " + std::to_string(i) + ": texture index out of range");
      return PB2_E_INVALID;
    }
    if ((tex[0] || tex[1] || tex[2]) && m.kind != PB2_MAT_UBER) {
      set_error("material " + std::to_string(i) + ": image nodes are supported on Uber materials only");
      return PB2_E_UNSUPPORTED;
    }
  }
  // node programs: every *_node names the start of a well-formed postfix program
  ctx->has_node_programs = false;
  for (uint32_t i = 0; i < sd->n_materials; i++) {
    const pb2_material& m = sd->materials[i];
    const int32_t prog[5] = {m.color_node, m.roughness_node, m.ior_node, m.metallic_node, m.transmission_node};
    for (int k = 0; k < 5; k++) {
      if (!prog[k]) continue;
      const std::string where = "material " + std::to_string(i) + " node program " + std::to_string(k) + ": ";
      if (m.kind == PB2_MAT_EMISSIVE) { set_error(where + "Emissive::color must be a constant on the CUDA path"); return PB2_E_UNSUPPORTED; }
      if (prog[k] < 0 || uint32_t(prog[k]) > sd->n_node_instrs || !sd->node_instrs) { set_error(where + "start out of range"); return PB2_E_INVALID; }
      int depth = 0;
      bool ended = false;
      for (uint32_t pc = uint32_t(prog[k]) - 1; pc < sd->n_node_instrs && !ended; pc++) {
        const pb2_node_instr& in = sd->node_instrs[pc];
        int pops = 0, pushes = 1;
        switch (in.op) {
          case PB2_NODE_END: ended = true; pushes = 0; break;
          case PB2_NODE_CONST: case PB2_NODE_POSITION: case PB2_NODE_NORMAL: case PB2_NODE_UV: break;
          case PB2_NODE_ADD: case PB2_NODE_SUB: case PB2_NODE_MUL: case PB2_NODE_DIV: case PB2_NODE_POW: pops = 2; break;
          case PB2_NODE_NEG: case PB2_NODE_ABS: case PB2_NODE_SQR: case PB2_NODE_SQRT: case PB2_NODE_FRACT:
          case PB2_NODE_CHECKERBOARD: pops = 1; break;
          case PB2_NODE_COMPONENT:
            pops = 1;
            if (!(in.a[0] == 0.0f || in.a[0] == 1.0f || in.a[0] == 2.0f)) { set_error(where + "component index must be 0, 1 or 2"); return PB2_E_INVALID; }
            break;
          case PB2_NODE_VEC3: pops = 3; break;
          case PB2_NODE_IMAGE: case PB2_NODE_IMAGEF:
            pops = 1;
            if (!(in.a[0] >= 1.0f && in.a[0] <= float(sd->n_textures) && in.a[0] == float(int(in.a[0])))) { set_error(where + "texture index out of range"); return PB2_E_INVALID; }
            break;
          default: set_error(where + "unknown node op " + std::to_string(in.op) + " (noise and function nodes have no device form)"); return PB2_E_UNSUPPORTED;
        }
        if (ended) break;
        if (depth < pops) { set_error(where + "stack underflow"); return PB2_E_INVALID; }
        depth += pushes - pops;
        if (depth > PB2_NODE_MAX_STACK) { set_error(where + "needs more than PB2_NODE_MAX_STACK values"); return PB2_E_UNSUPPORTED; }
      }
      if (!ended || depth != 1) { set_error(where + "does not end with exactly one value before PB2_NODE_END"); return PB2_E_INVALID; }
      ctx->has_node_programs = true;
    }
  }
  for (uint32_t i = 0; i < sd->n_textures; i++)
    if (!sd->textures[i].rgba || sd->textures[i].width <= 0 || sd->textures[i].height <= 0) {
      set_error("texture " + std::to_string(i) + ": empty image");
      return PB2_E_INVALID;
    }

  std::vector<DGeom> geoms(sd->n_geometries);
  ctx->needs_rng_in_shading = false;
  ctx->material_kinds = 0;
  for (uint32_t i = 0; i < sd->n_geometries; i++) {
    const pb2_geometry& g = sd->geometries[i];
    if (g.shape_kind < 0 || g.shape_kind > PB2_SHAPE_CYLINDER) {
      set_error("geometry " + std::to_string(i) + ": shape kind not supported by the CUDA path");
      return PB2_E_UNSUPPORTED;
    }
    if (g.shape_index < 0 || uint32_t(g.shape_index) >= counts[g.shape_kind] || g.material < 0 ||
        uint32_t(g.material) >= sd->n_materials) {
      set_error("geometry " + std::to_string(i) + ": index out of range");
      return PB2_E_INVALID;
    }
    const pb2_material& m = sd->materials[g.material];
    if (m.kind == PB2_MAT_EMISSIVE && (g.shape_kind == PB2_SHAPE_AABB || g.shape_kind == PB2_SHAPE_OBB)) {
      // AABB::pdf / OBB::pdf are PINE_UNREACHABLE in the reference (bbox.h:87,103)
      set_error("emissive Box geometry is undefined in the reference; not supported");
      return PB2_E_UNSUPPORTED;
    }
    if (m.kind == PB2_MAT_EMISSIVE && (g.shape_kind == PB2_SHAPE_PLANE || g.shape_kind == PB2_SHAPE_CYLINDER)) {
      // Cylinder::sample / pdf / area are PINE_UNREACHABLE (geometry.h:148-150); a Plane light (area = float_max,
      // hemisphere sampling, geometry.cpp:57-70) has no device sampler
      set_error("emissive Plane / Cylinder geometry is not supported");
      return PB2_E_UNSUPPORTED;
    }
    if (m.kind == PB2_MAT_SUBSURFACE && (g.shape_kind == PB2_SHAPE_PLANE || g.shape_kind == PB2_SHAPE_CYLINDER)) {
      set_error("Subsurface on a Plane / Cylinder is not supported (the reference's walk is ill-defined there)");
      return PB2_E_UNSUPPORTED;
    }
    if (m.kind == PB2_MAT_UBER || m.kind == PB2_MAT_SUBSURFACE) ctx->needs_rng_in_shading = true;
    ctx->material_kinds |= 1u << m.kind;
    geoms[i] = DGeom{g.shape_kind, g.shape_index, g.material, -1};
  }
  // UniformLightSampler::lights: light_records when given, else one AreaLight per light_geometries entry
  std::vector<pb2_light> light_recs;
  if (sd->light_records) {
    light_recs.assign(sd->light_records, sd->light_records + sd->n_light_records);
  } else {
    for (uint32_t l = 0; l < sd->n_lights; l++) {
      pb2_light r{};
      r.kind = PB2_LIGHT_AREA;
      r.geometry = sd->light_geometries[l];
      light_recs.push_back(r);
    }
  }
  std::vector<int32_t> lights(light_recs.size(), -1);
  bool only_area = true;
  int32_t env_kind = 0;
  float env_color[3] = {0, 0, 0};
  for (size_t l = 0; l < light_recs.size(); l++) {
    const pb2_light& r = light_recs[l];
    if (r.kind < PB2_LIGHT_AREA || r.kind > PB2_LIGHT_SKY) { set_error("unknown light kind"); return PB2_E_UNSUPPORTED; }
    if (r.kind != PB2_LIGHT_AREA) {
      only_area = false;
      if (r.kind == PB2_LIGHT_SKY) {
        if (l + 1 != light_recs.size()) { set_error("the environment light must be the last light record"); return PB2_E_INVALID; }
        env_kind = PB2_LIGHT_SKY;
        for (int k = 0; k < 3; k++) env_color[k] = r.color[k];
      }
      continue;
    }
    if (r.geometry < 0 || uint32_t(r.geometry) >= sd->n_geometries) { set_error("light index out of range"); return PB2_E_INVALID; }
    if (sd->materials[sd->geometries[r.geometry].material].kind != PB2_MAT_EMISSIVE) {
      set_error("light geometry is not emissive");
      return PB2_E_INVALID;
    }
    lights[l] = r.geometry;
    geoms[r.geometry].light = int32_t(l);
  }
  ctx->only_area_lights = only_area;

  // ---- host copies (kept, pinned, for re-staging) ----
  HostScene& H = ctx->host;
  H.unpin();
  H.unmap();
  H = HostScene();
  H.geoms = std::move(geoms);
  H.lights = std::move(lights);
  H.light_recs = std::move(light_recs);
  H.materials.assign(sd->materials, sd->materials + sd->n_materials);
  for (uint32_t i = 0; i < sd->n_textures; i++) {
    const pb2_texture& t = sd->textures[i];
    const size_t n = size_t(t.width) * size_t(t.height);
    if (H.texels.size() + n > 0xffffffffull) { set_error("more than 2^32 texels"); return PB2_E_UNSUPPORTED; }
    H.textures.push_back(DTexture{uint32_t(H.texels.size()), t.width, t.height});
    const size_t at = H.texels.size();
    H.texels.resize(at + n);
    memcpy(H.texels.data() + at, t.rgba, n * 4);
  }
  if (sd->node_instrs) H.node_instrs.assign(sd->node_instrs, sd->node_instrs + sd->n_node_instrs);
  H.srgb_lut.resize(256);
  // inverse_gamma_correction(value / 255.0f) = pow(v, 2.2f) (color.h:13-18) with the host's powf, as the
  // reference evaluates it per lookup
  for (int v = 0; v < 256; v++) H.srgb_lut[v] = powf(float(v) / 255.0f, 2.2f);
  H.aabbs.assign(sd->aabbs, sd->aabbs + sd->n_aabbs);
  H.obbs.assign(sd->obbs, sd->obbs + sd->n_obbs);
  H.spheres.assign(sd->spheres, sd->spheres + sd->n_spheres);
  H.triangles.assign(sd->triangles, sd->triangles + sd->n_triangles);
  H.cones.assign(sd->cones, sd->cones + sd->n_cones);
  H.rects.assign(sd->rects, sd->rects + sd->n_rects);
  H.disks.assign(sd->disks, sd->disks + sd->n_disks);
  if (sd->n_planes) H.planes.assign(sd->planes, sd->planes + sd->n_planes);
  if (sd->n_cylinders) H.cylinders.assign(sd->cylinders, sd->cylinders + sd->n_cylinders);

  // ---- mesh pools ----
  H.meshes.resize(sd->n_meshes);
  bool any_normals = false, any_uvs = false;
  size_t n_verts = 0, n_tris = 0;
  for (uint32_t k = 0; k < sd->n_meshes; k++) {
    any_normals |= sd->meshes[k].normals != nullptr;
    any_uvs |= sd->meshes[k].texcoords != nullptr;
    n_verts += sd->meshes[k].n_vertices;
    n_tris += sd->meshes[k].n_triangles;
  }
  H.verts.resize(n_verts * 3);
  H.tris.resize(n_tris * 3);
  if (any_normals) H.normals.resize(n_verts * 3);  // meshes without the attribute are zero-filled in their job below
  if (any_uvs) H.uvs.resize(n_verts * 2);
  {
    size_t vb = 0, tb = 0;
    for (uint32_t k = 0; k < sd->n_meshes; k++) {
      const pb2_mesh& m = sd->meshes[k];
      DMesh& dm = H.meshes[k];
      std::memset(&dm, 0, sizeof(dm));
      dm.tree = DTree{-1, 0, 0, 0};
      dm.vert_base = uint32_t(vb);
      dm.tri_base = uint32_t(tb);
      dm.n_tris = m.n_triangles;
      dm.has_normals = m.normals != nullptr;
      dm.has_uvs = m.texcoords != nullptr;
      vb += m.n_vertices;
      tb += m.n_triangles;
    }
  }
  const int hw = int(std::thread::hardware_concurrency());
  const char* env_threads = std::getenv("PB2_BUILD_THREADS");
  const int n_threads = env_threads ? std::max(1, std::atoi(env_threads)) : std::max(1, std::min(hw > 0 ? hw : 1, 64));
  ctx->counters.build_threads = uint32_t(n_threads);
  std::atomic<bool> bad_index{false};
  parallel_jobs(sd->n_meshes, n_threads, [&](size_t k) {
    const pb2_mesh& m = sd->meshes[k];
    const DMesh& dm = H.meshes[k];
    for (size_t t = 0; t < size_t(m.n_triangles) * 3; t++)
      if (m.indices[t] >= m.n_vertices) { bad_index = true; return; }
    std::memcpy(H.verts.data() + size_t(dm.vert_base) * 3, m.vertices, size_t(m.n_vertices) * 12);
    std::memcpy(H.tris.data() + size_t(dm.tri_base) * 3, m.indices, size_t(m.n_triangles) * 12);
    if (m.normals) std::memcpy(H.normals.data() + size_t(dm.vert_base) * 3, m.normals, size_t(m.n_vertices) * 12);
    else if (any_normals) std::memset(H.normals.data() + size_t(dm.vert_base) * 3, 0, size_t(m.n_vertices) * 12);
    if (m.texcoords) std::memcpy(H.uvs.data() + size_t(dm.vert_base) * 2, m.texcoords, size_t(m.n_vertices) * 8);
    else if (any_uvs) std::memset(H.uvs.data() + size_t(dm.vert_base) * 2, 0, size_t(m.n_vertices) * 8);
  });
  if (bad_index) { set_error("mesh index out of range"); return PB2_E_INVALID; }

  stage("validate + mesh pools");
  // ---- per-mesh trees (bvh.cpp:460-476): meshes with >=1 triangle, in geometry order ----
  ctx->mesh_trees.clear();
  std::vector<int> mesh_tree_of(sd->n_meshes, -1);
  std::vector<BuildPrim> top;
  struct MeshJob { uint32_t geom; int mesh; };
  std::vector<MeshJob> jobs;
  for (uint32_t i = 0; i < sd->n_geometries; i++) {
    const pb2_geometry& g = sd->geometries[i];
    if (g.shape_kind != PB2_SHAPE_MESH) continue;
    if (sd->meshes[g.shape_index].n_triangles == 0) continue;
    if (mesh_tree_of[g.shape_index] != -1) { set_error("a mesh may back only one geometry"); return PB2_E_INVALID; }
    mesh_tree_of[g.shape_index] = int(jobs.size());
    jobs.push_back(MeshJob{i, g.shape_index});
  }
  ctx->mesh_trees.resize(jobs.size());
  std::atomic<int> max_height{0};
  // Small meshes: one job per mesh. Big meshes additionally split inside build_sah_binary.
  set_build_threads(jobs.size() >= size_t(n_threads) ? 1 : n_threads);
  parallel_jobs(jobs.size(), n_threads, [&](size_t j) {
    const pb2_mesh& m = sd->meshes[jobs[j].mesh];
    std::vector<BuildPrim> prims(m.n_triangles);
    for (uint32_t t = 0; t < m.n_triangles; t++) {
      const uint32_t* f = m.indices + 3 * size_t(t);
      prims[t].box = aabb_of_tri(ld3(m.vertices + 3 * size_t(f[0])), ld3(m.vertices + 3 * size_t(f[1])),
                                 ld3(m.vertices + 3 * size_t(f[2])));
      prims[t].index = int(t);
    }
    build_sah_binary(std::move(prims), ctx->mesh_trees[j]);
    int h = tree_height(ctx->mesh_trees[j]), seen = max_height.load();
    while (h > seen && !max_height.compare_exchange_weak(seen, h)) {}
  });
  for (size_t j = 0; j < jobs.size(); j++) {
    BuildPrim tp;
    tp.box = ctx->mesh_trees[j].bounds();
    tp.index = int(top.size());
    top.push_back(tp);
    H.top_prims.push_back(DTopPrim{int32_t(jobs[j].geom), PB2_SHAPE_MESH, jobs[j].mesh, 0});
  }
  // ---- analytic shapes (bvh.cpp:485-493) ----
  for (uint32_t i = 0; i < sd->n_geometries; i++) {
    const pb2_geometry& g = sd->geometries[i];
    if (g.shape_kind == PB2_SHAPE_MESH) continue;
    BuildPrim tp;
    switch (g.shape_kind) {
      case PB2_SHAPE_AABB: tp.box = aabb_of(sd->aabbs[g.shape_index]); break;
      case PB2_SHAPE_OBB: tp.box = aabb_of(sd->obbs[g.shape_index]); break;
      case PB2_SHAPE_SPHERE: tp.box = aabb_of(sd->spheres[g.shape_index]); break;
      case PB2_SHAPE_TRIANGLE: tp.box = aabb_of(sd->triangles[g.shape_index]); break;
      case PB2_SHAPE_CONE: tp.box = aabb_of(sd->cones[g.shape_index]); break;
      case PB2_SHAPE_RECT: tp.box = aabb_of(sd->rects[g.shape_index]); break;
      case PB2_SHAPE_DISK: tp.box = aabb_of(sd->disks[g.shape_index]); break;
      case PB2_SHAPE_PLANE: tp.box = aabb_of(sd->planes[g.shape_index]); break;
      case PB2_SHAPE_CYLINDER: tp.box = aabb_of(sd->cylinders[g.shape_index]); break;
    }
    tp.index = int(top.size());
    top.push_back(tp);
    H.top_prims.push_back(DTopPrim{int32_t(i), g.shape_kind, g.shape_index, 0});
  }
  stage("mesh trees");
  set_build_threads(n_threads);
  std::vector<Box3> top_boxes(top.size());
  for (size_t i = 0; i < top.size(); i++) top_boxes[i] = top[i].box;
  build_sah_binary(top, ctx->top_tree);
  // The reference-order walk (verification path and certified fallback) keeps a 64-entry stack per tree level
  // (the reference itself has 32 and undefined behaviour beyond, bvh.cpp:363): a deeper tree -- degenerate
  // input such as thousands of coincident centroids -- is refused instead of being walked with dropped pushes.
  if (std::max(max_height.load(), tree_height(ctx->top_tree)) > kRefOrderStack) {
    set_error("a BVH is deeper than " + std::to_string(kRefOrderStack) +
              " levels (degenerate geometry): the reference-order traversal stack would overflow");
    return PB2_E_UNSUPPORTED;
  }

  // ---- device layouts ----
  LayoutOut& lay = H.lay;
  DTree top_tree;
  {
    std::vector<const Tree2*> trees{&ctx->top_tree};
    std::vector<uint32_t> mesh_of_entry;
    for (uint32_t k = 0; k < sd->n_meshes; k++)
      if (mesh_tree_of[k] >= 0) { trees.push_back(&ctx->mesh_trees[mesh_tree_of[k]]); mesh_of_entry.push_back(k); }
    const std::vector<DTree> roots = append_binary_many(trees, lay, n_threads);
    top_tree = roots[0];
    for (size_t i = 0; i < mesh_of_entry.size(); i++) H.meshes[mesh_of_entry[i]].tree = roots[i + 1];
  }
  stage("top tree + binary layout");
  WideInput wi;
  wi.top = &ctx->top_tree; wi.top_prims = &H.top_prims; wi.top_boxes = &top_boxes;
  wi.mesh_trees = &ctx->mesh_trees; wi.mesh_tree_of = &mesh_tree_of; wi.meshes = sd->meshes;
  build_wide(wi, lay, n_threads);
  stage("wide tree");
  // more AABB/OBB shapes than the per-ray pre-test handles: every closest hit in reference order
  const bool order_dependent = lay.box_shapes.size() > size_t(kMaxBoxShapes);

  ctx->counters.bvh2_nodes = uint32_t(lay.nodes2.size());
  ctx->counters.wide_nodes = uint32_t(lay.nodes8.size());
  ctx->counters.n_primitives = uint32_t(lay.leaf_prims.size());
  ctx->counters.bvh_build_ms =
      std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  DScene& s = ctx->scene;
  s.top = top_tree;
  s.n_lights = int32_t(H.lights.size()); s.env_kind = env_kind;
  for (int k = 0; k < 3; k++) s.env_color[k] = env_color[k];
  s.n_geoms = int32_t(sd->n_geometries);
  s.wide_ready = lay.nodes8.empty() ? 0 : 1;
  s.n_nodes8 = int32_t(lay.nodes8.size());
  for (int k = 0; k < 3; k++) s.node_origin_absmax[k] = 0.0f;
  for (const DNode8& nd : lay.nodes8)
    for (int k = 0; k < 3; k++) s.node_origin_absmax[k] = std::max(s.node_origin_absmax[k], std::fabs(nd.origin[k]));
  s.n_box_shapes = int32_t(lay.box_shapes.size());
  s.order_dependent = order_dependent ? 1 : 0;
  s.shape_mask = 0;
  for (uint32_t i = 0; i < sd->n_geometries; i++)
    if (sd->geometries[i].shape_kind != PB2_SHAPE_MESH) s.shape_mask |= 1 << sd->geometries[i].shape_kind;
  s.has_rare_shapes = (s.shape_mask & ((1 << PB2_SHAPE_PLANE) | (1 << PB2_SHAPE_CYLINDER))) ? 1 : 0;
  s.has_meshes = jobs.empty() ? 0 : 1;
  if (s.has_meshes) s.shape_mask |= 1 << PB2_SHAPE_MESH;
  s.camera = sd->camera;
  ctx->n_pixels = sd->camera.width * sd->camera.height;
  ctx->host_scene_ready = true;

  // ---- staging list + page-lock ----
  uint64_t bytes = 0;
  auto item = [&](DBuf pb2_context::* dst, const void* src, size_t n) {
    H.items.push_back(HostScene::Item{dst, src, n});
    bytes += n;
  };
#define ITEM(buf, vec) item(&pb2_context::buf, (vec).data(), (vec).size() * sizeof((vec)[0]))
  ITEM(nodes2, lay.nodes2); ITEM(nodes8, lay.nodes8); ITEM(leaves, lay.leaves);
  ITEM(wide_prims, lay.wide_prims); ITEM(box_shapes, lay.box_shapes); ITEM(leaf_prims, lay.leaf_prims);
  ITEM(top_prims, H.top_prims); ITEM(meshes, H.meshes); ITEM(geoms, H.geoms); ITEM(lights, H.lights); ITEM(light_recs, H.light_recs);
  ITEM(verts, H.verts); ITEM(tris, H.tris); ITEM(normals, H.normals); ITEM(uvs, H.uvs);
  ITEM(texels, H.texels); ITEM(textures, H.textures); ITEM(srgb_lut, H.srgb_lut); ITEM(node_instrs, H.node_instrs);
  ITEM(materials, H.materials); ITEM(aabbs, H.aabbs); ITEM(obbs, H.obbs); ITEM(spheres, H.spheres);
  ITEM(triangles, H.triangles); ITEM(cones, H.cones); ITEM(rects, H.rects); ITEM(disks, H.disks);
  ITEM(planes, H.planes); ITEM(cylinders, H.cylinders);
#undef ITEM
  // Page-locking 1.4 GB costs ~0.4 s, more than copying it once from pageable memory: the first staging (this
  // call) copies from pageable memory and the arrays are page-locked by the first RE-staging (HostScene::pin),
  // which is the caller that profits from pinned copies.
  ctx->counters.scene_bytes = bytes;
  if (ctx->device < 0) return PB2_OK;  // host-only context: builder inspection / export, no device work
  if (ctx->defer_staging) return PB2_OK;  // pb2_multi_upload_scene stages the replicas itself (sharded + all-gather)
  int rc = restage_unpinned(ctx);
  if (rc) return rc;
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
#ifdef PB2_COUNT
  PB2_CUDA_CHECK(ctx->walk_stats.reserve(80));
  PB2_CUDA_CHECK(cudaMemset(ctx->walk_stats.ptr, 0, 80));
  ctx->scene.walk_stats = ctx->walk_stats.as<unsigned long long>();
#endif
  ctx->scene_ready = true;
  return PB2_OK;
}

namespace {
// Copies the staging list (whole, or part `part` of `n_parts` of every array >= kShardMinBytes) and rebinds the
// device pointers of ctx->scene.
constexpr size_t kShardMinBytes = size_t(1) << 20;
int restage(pb2_context* ctx, void* cuda_stream, int part, int n_parts, pb2_device_array* arrays, int max_arrays,
            int* n_arrays, bool pin = true) {
  if (!ctx || !ctx->host_scene_ready || ctx->device < 0) { set_error("no host scene / no device"); return PB2_E_STATE; }
  if (n_parts < 1 || part < 0 || part >= n_parts) { set_error("part out of range"); return PB2_E_INVALID; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  auto t0 = std::chrono::steady_clock::now();
  int n_out = 0;
  const HostScene& host = ctx->host_view ? *ctx->host_view : ctx->host;
  // page-lock on the first RE-staging only (the first staging of a fresh scene copies from pageable memory)
  if (pin && ctx->scene_ready) const_cast<HostScene&>(host).pin();
  if (!host.pinned && n_parts == 1) {
    // First staging of a fresh scene, from PAGEABLE memory: the driver stages such copies through its own
    // bounce buffers with one host thread's memcpy (4.5 GB/s measured, 0.32 s for C5's 1.43 GB), so the
    // arrays are cut into chunks and copied by several host threads on their own streams. Synchronous: every
    // copy has landed when this returns.
    struct Chunk { char* dst; const char* src; size_t bytes; };
    std::vector<Chunk> chunks;
    constexpr size_t kChunk = size_t(32) << 20;
    for (const HostScene::Item& it : host.items) {
      DBuf& dst = ctx->*(it.dst);
      PB2_CUDA_CHECK(dst.reserve(it.bytes));
      for (size_t off = 0; off < it.bytes; off += kChunk)
        chunks.push_back(Chunk{static_cast<char*>(dst.ptr) + off, static_cast<const char*>(it.src) + off, std::min(kChunk, it.bytes - off)});
    }
    PB2_CUDA_CHECK(cudaStreamSynchronize(st));
    const int copy_threads = int(std::min<size_t>(8, std::max<size_t>(1, chunks.size() / 2)));
    std::atomic<size_t> next{0};
    std::atomic<int> failed{0};
    auto work = [&] {
      cudaStream_t cs = nullptr;
      if (cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { failed = 1; return; }
      for (size_t i; (i = next.fetch_add(1)) < chunks.size();)
        if (cudaMemcpyAsync(chunks[i].dst, chunks[i].src, chunks[i].bytes, cudaMemcpyHostToDevice, cs) != cudaSuccess) failed = 1;
      if (cudaStreamSynchronize(cs) != cudaSuccess) failed = 1;
      cudaStreamDestroy(cs);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < copy_threads; t++) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
    if (failed) { cudaGetLastError(); set_error("host->device staging failed"); return PB2_E_CUDA; }
  } else
  for (const HostScene::Item& it : host.items) {
    DBuf& dst = ctx->*(it.dst);
    if (n_parts == 1 || it.bytes < kShardMinBytes) {
      PB2_CUDA_CHECK(dst.reserve(it.bytes));
      if (it.bytes) PB2_CUDA_CHECK(cudaMemcpyAsync(dst.ptr, it.src, it.bytes, cudaMemcpyHostToDevice, st));
      continue;
    }
    const size_t part_bytes = ((it.bytes + size_t(n_parts) - 1) / size_t(n_parts) + 255) / 256 * 256;
    PB2_CUDA_CHECK(dst.reserve(part_bytes * size_t(n_parts)));
    const size_t off = part_bytes * size_t(part);
    if (off < it.bytes) {
      const size_t len = std::min(part_bytes, it.bytes - off);
      PB2_CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(dst.ptr) + off, static_cast<const char*>(it.src) + off, len,
                                     cudaMemcpyHostToDevice, st));
    }
    if (arrays) {
      if (n_out >= max_arrays) { set_error("pb2_restage_scene_part: max_arrays too small"); return PB2_E_INVALID; }
      arrays[n_out] = pb2_device_array{dst.ptr, uint64_t(part_bytes) * uint64_t(n_parts), uint64_t(part_bytes)};
    }
    n_out++;
  }
  if (n_arrays) *n_arrays = n_out;
  if (ctx->tables_spp) {  // the active sampler tables travel with the scene
    int k = 0;
    while ((1 << k) != ctx->tables_spp) k++;
    const SamplerTables& T = ctx->tables[k];
    PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->tab_sobol.ptr, T.sobol.data(), T.sobol.size(), cudaMemcpyHostToDevice, st));
    PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->tab_scramble.ptr, T.scramble.data(), T.scramble.size(), cudaMemcpyHostToDevice, st));
    PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->tab_rank.ptr, T.rank.data(), T.rank.size(), cudaMemcpyHostToDevice, st));
  }
  DScene& s = ctx->scene;
  s.nodes2 = ctx->nodes2.as<DNode2>(); s.nodes8 = ctx->nodes8.as<DNode8>();
  s.leaves = ctx->leaves.as<DLeaf>(); s.wide_prims = ctx->wide_prims.as<DPrim>(); s.box_shapes = ctx->box_shapes.as<DBoxShape>(); s.leaf_prims = ctx->leaf_prims.as<int32_t>();
  s.top_prims = ctx->top_prims.as<DTopPrim>(); s.meshes = ctx->meshes.as<DMesh>();
  s.geoms = ctx->geoms.as<DGeom>(); s.materials = ctx->materials.as<pb2_material>();
  s.lights = ctx->lights.as<int32_t>();
  s.light_recs = ctx->light_recs.as<pb2_light>();
  s.aabbs = ctx->aabbs.as<pb2_aabb>(); s.obbs = ctx->obbs.as<pb2_obb>();
  s.spheres = ctx->spheres.as<pb2_sphere>(); s.triangles = ctx->triangles.as<pb2_triangle>();
  s.cones = ctx->cones.as<pb2_cone>(); s.rects = ctx->rects.as<pb2_rect>();
  s.disks = ctx->disks.as<pb2_disk>();
  s.planes = ctx->planes.as<pb2_plane>(); s.cylinders = ctx->cylinders.as<pb2_cylinder>();
  s.verts = ctx->verts.as<float>(); s.tris = ctx->tris.as<uint32_t>();
  s.normals = ctx->normals.as<float>(); s.uvs = ctx->uvs.as<float>();
  s.texels = ctx->texels.as<uint32_t>(); s.textures = ctx->textures.as<DTexture>();
  s.srgb_lut = ctx->srgb_lut.as<float>(); s.node_instrs = ctx->node_instrs.as<pb2_node_instr>();
  ctx->counters.upload_ms =
      std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return PB2_OK;
}
}  // namespace

int restage_unpinned(pb2_context* ctx) { return restage(ctx, nullptr, 0, 1, nullptr, 0, nullptr, false); }
extern "C" int pb2_restage_scene(pb2_context* ctx, void* cuda_stream) {
  return restage(ctx, cuda_stream, 0, 1, nullptr, 0, nullptr);
}
extern "C" int pb2_restage_scene_part(pb2_context* ctx, void* cuda_stream, int part, int n_parts,
                                      pb2_device_array* arrays, int max_arrays, int* n_arrays) {
  return restage(ctx, cuda_stream, part, n_parts, arrays, max_arrays, n_arrays);
}

// ---- sharing one host build between the processes of a node ---------------------------------------------------
// With one process per GPU (torchrun) every rank would build the same tree on the same host. Instead ONE rank
// builds (pb2_upload_scene) and writes its staging arrays plus the scalar scene state to a file -- /dev/shm: page
// cache, no disk -- and the others map that file and stage from it; everything downstream (pb2_restage_scene,
// pb2_restage_scene_part, the renders) is unchanged because a HostScene is just a list of (member, pointer, bytes).
namespace {
struct SharedHeader {
  char magic[8];  // "PB2SCN06"
  uint32_t sizeof_dscene, n_items;
  DScene scene;   // device pointers inside are meaningless to the importer and rebound by the staging
  int32_t only_area_lights, needs_rng_in_shading, n_pixels, has_node_programs;
  uint32_t material_kinds;
  pb2_counters counters;
};
// the staging members in the order pb2_upload_scene lists them (ITEM(...) above)
DBuf pb2_context::* const kStageMembers[] = {
    &pb2_context::nodes2, &pb2_context::nodes8, &pb2_context::leaves, &pb2_context::wide_prims, &pb2_context::box_shapes,
    &pb2_context::leaf_prims, &pb2_context::top_prims, &pb2_context::meshes, &pb2_context::geoms, &pb2_context::lights,
    &pb2_context::light_recs, &pb2_context::verts, &pb2_context::tris, &pb2_context::normals, &pb2_context::uvs,
    &pb2_context::texels, &pb2_context::textures, &pb2_context::srgb_lut, &pb2_context::node_instrs,
    &pb2_context::materials, &pb2_context::aabbs,
    &pb2_context::obbs, &pb2_context::spheres, &pb2_context::triangles, &pb2_context::cones, &pb2_context::rects,
    &pb2_context::disks, &pb2_context::planes, &pb2_context::cylinders};
constexpr uint32_t kNumStageMembers = sizeof(kStageMembers) / sizeof(kStageMembers[0]);
constexpr size_t kShareAlign = 4096;
size_t align_up(size_t v) { return (v + kShareAlign - 1) / kShareAlign * kShareAlign; }
}  // namespace

void pb2::HostScene::unmap() {
  if (map_base) munmap(map_base, map_size);
  map_base = nullptr;
  map_size = 0;
}

extern "C" int pb2_export_host_scene(pb2_context* ctx, const char* path) {
  if (!ctx || !path) { set_error("null argument"); return PB2_E_INVALID; }
  if (!ctx->host_scene_ready || ctx->host_view) { set_error("pb2_export_host_scene needs a scene this context built"); return PB2_E_STATE; }
  const HostScene& H = ctx->host;
  if (H.items.size() != kNumStageMembers) { set_error("staging list out of date"); return PB2_E_STATE; }
  for (uint32_t i = 0; i < kNumStageMembers; i++)
    if (H.items[i].dst != kStageMembers[i]) { set_error("staging list order changed"); return PB2_E_STATE; }
  SharedHeader hd{};
  std::memcpy(hd.magic, "PB2SCN06", 8);
  hd.sizeof_dscene = uint32_t(sizeof(DScene));
  hd.n_items = kNumStageMembers;
  hd.scene = ctx->scene;
  hd.only_area_lights = ctx->only_area_lights; hd.needs_rng_in_shading = ctx->needs_rng_in_shading;
  hd.has_node_programs = ctx->has_node_programs;
  hd.n_pixels = ctx->n_pixels; hd.material_kinds = ctx->material_kinds;
  hd.counters = ctx->counters;
  std::vector<uint64_t> table(2 * size_t(kNumStageMembers));
  size_t off = align_up(sizeof(SharedHeader) + table.size() * 8);
  for (uint32_t i = 0; i < kNumStageMembers; i++) {
    table[2 * i] = H.items[i].bytes;
    table[2 * i + 1] = off;
    off += align_up(H.items[i].bytes);
  }
  const std::string tmp = std::string(path) + ".tmp";
  int fd = open(tmp.c_str(), O_CREAT | O_TRUNC | O_RDWR, 0600);
  if (fd < 0) { set_error("cannot create " + tmp); return PB2_E_INVALID; }
  bool ok = ftruncate(fd, off_t(off)) == 0;
  void* base = ok ? mmap(nullptr, off, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0) : MAP_FAILED;
  close(fd);
  if (base == MAP_FAILED) { unlink(tmp.c_str()); set_error("cannot map " + tmp); return PB2_E_INVALID; }
  std::memcpy(base, &hd, sizeof(hd));
  std::memcpy(static_cast<char*>(base) + sizeof(hd), table.data(), table.size() * 8);
  // the copies are the cost of this call (1.4 GB on C5): spread them over the builder's threads
  const int threads = int(std::max<uint32_t>(1, ctx->counters.build_threads));
  parallel_jobs(kNumStageMembers, threads, [&](size_t i) {
    if (H.items[i].bytes) std::memcpy(static_cast<char*>(base) + table[2 * i + 1], H.items[i].src, H.items[i].bytes);
  });
  munmap(base, off);
  if (rename(tmp.c_str(), path) != 0) { unlink(tmp.c_str()); set_error(std::string("cannot rename to ") + path); return PB2_E_INVALID; }
  return PB2_OK;
}

extern "C" int pb2_import_host_scene(pb2_context* ctx, const char* path) {
  if (!ctx || !path) { set_error("null argument"); return PB2_E_INVALID; }
  if (ctx->device >= 0) PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  int fd = open(path, O_RDONLY);
  if (fd < 0) { set_error(std::string("cannot open ") + path); return PB2_E_INVALID; }
  struct stat st{};
  if (fstat(fd, &st) != 0 || size_t(st.st_size) < sizeof(SharedHeader)) { close(fd); set_error("shared scene file too small"); return PB2_E_INVALID; }
  // private mapping: pages stay shared with the page cache until written (never), and can be page-locked
  void* base = mmap(nullptr, size_t(st.st_size), PROT_READ | PROT_WRITE, MAP_PRIVATE, fd, 0);
  close(fd);
  if (base == MAP_FAILED) { set_error("cannot map the shared scene file"); return PB2_E_INVALID; }
  SharedHeader hd;
  std::memcpy(&hd, base, sizeof(hd));
  const uint64_t* table = reinterpret_cast<const uint64_t*>(static_cast<const char*>(base) + sizeof(hd));
  bool ok = std::memcmp(hd.magic, "PB2SCN06", 8) == 0 && hd.sizeof_dscene == sizeof(DScene) && hd.n_items == kNumStageMembers &&
            size_t(st.st_size) >= sizeof(hd) + size_t(hd.n_items) * 16;
  for (uint32_t i = 0; ok && i < hd.n_items; i++) ok = table[2 * i + 1] + table[2 * i] <= uint64_t(st.st_size);
  if (!ok) { munmap(base, size_t(st.st_size)); set_error("not a scene file of this library version"); return PB2_E_INVALID; }
  ctx->scene_ready = false;
  ctx->host_scene_ready = false;
  ctx->host_view = nullptr;
  HostScene& H = ctx->host;
  H.unpin();
  H.unmap();
  H = HostScene();
  H.map_base = base;
  H.map_size = size_t(st.st_size);
  for (uint32_t i = 0; i < hd.n_items; i++)
    H.items.push_back(HostScene::Item{kStageMembers[i], static_cast<const char*>(base) + table[2 * i + 1], size_t(table[2 * i])});
  ctx->top_tree = Tree2();  // the tree dumps (pb2_bvh_nodes) stay with the process that built
  ctx->mesh_trees.clear();
  ctx->scene = hd.scene;
  ctx->only_area_lights = hd.only_area_lights != 0;
  ctx->needs_rng_in_shading = hd.needs_rng_in_shading != 0;
  ctx->has_node_programs = hd.has_node_programs != 0;
  ctx->n_pixels = hd.n_pixels;
  ctx->material_kinds = hd.material_kinds;
  const pb2_counters keep = ctx->counters;
  ctx->counters = hd.counters;
  ctx->counters.closest_rays = keep.closest_rays; ctx->counters.shadow_rays = keep.shadow_rays;
  ctx->counters.camera_samples = keep.camera_samples; ctx->counters.kernel_launches = keep.kernel_launches;
  ctx->host_scene_ready = true;
  if (ctx->device < 0) return PB2_OK;
  int rc = restage_unpinned(ctx);
  if (rc) return rc;
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  ctx->scene_ready = true;
  return PB2_OK;
}

extern "C" int pb2_bvh_info(pb2_context* ctx, int which, int32_t* n_nodes, int32_t* root,
                            int32_t* n_leaf_prims) {
  if (!ctx || !ctx->host_scene_ready) { set_error("no scene"); return PB2_E_STATE; }
  const Tree2* t = which < 0 ? &ctx->top_tree
                   : size_t(which) < ctx->mesh_trees.size() ? &ctx->mesh_trees[which] : nullptr;
  if (!t) { *n_nodes = 0; return PB2_OK; }
  *n_nodes = int32_t(t->nodes.size());
  *root = t->root;
  *n_leaf_prims = int32_t(t->leaf_prims.size());
  return PB2_OK;
}
extern "C" int pb2_bvh_nodes(pb2_context* ctx, int which, float* aabbs, int32_t* children,
                             int32_t* leaf_range, int32_t* leaf_prims) {
  if (!ctx || !ctx->host_scene_ready) { set_error("no scene"); return PB2_E_STATE; }
  const Tree2* t = which < 0 ? &ctx->top_tree
                   : size_t(which) < ctx->mesh_trees.size() ? &ctx->mesh_trees[which] : nullptr;
  if (!t) return PB2_E_INVALID;
  for (size_t i = 0; i < t->nodes.size(); i++) {
    const Node2& nd = t->nodes[i];
    for (int c = 0; c < 2; c++)
      for (int k = 0; k < 3; k++) {
        aabbs[i * 12 + c * 6 + k] = nd.box[c].lower[k];
        aabbs[i * 12 + c * 6 + 3 + k] = nd.box[c].upper[k];
      }
    children[i * 2] = nd.child[0];
    children[i * 2 + 1] = nd.child[1];
    leaf_range[i * 2] = nd.leaf_offset;
    leaf_range[i * 2 + 1] = nd.leaf_count;
  }
  for (size_t i = 0; i < t->leaf_prims.size(); i++) leaf_prims[i] = t->leaf_prims[i];
  return PB2_OK;
}
