// scene_upload.cu -- pb2_upload_scene: validate the flattened scene, build the BVHs on the host
// (reference-identical binary trees, then the wide collapse) and lay everything out in HBM.
// Replaces accel.build(&scene) + light_sampler.build(&scene) (integrator.cpp:32-33).
#include <chrono>
#include <cstring>

#include "bvh_layout.h"
#include "context.h"

using namespace pb2;

namespace {

bool valid_material(const pb2_material& m) {
  return m.kind >= PB2_MAT_EMISSIVE && m.kind <= PB2_MAT_SUBSURFACE;
}

}  // namespace

extern "C" int pb2_upload_scene(pb2_context* ctx, const pb2_scene_desc* sd) {
  if (!ctx || !sd) { set_error("null argument"); return PB2_E_INVALID; }
  if (ctx->device >= 0) PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  ctx->host_scene_ready = false;
  auto t0 = std::chrono::steady_clock::now();
  ctx->scene_ready = false;
  if (sd->camera.width <= 0 || sd->camera.height <= 0) {
    set_error("camera film size must be positive");
    return PB2_E_INVALID;
  }
  const uint32_t counts[8] = {sd->n_aabbs, sd->n_obbs, sd->n_spheres, sd->n_triangles,
                              sd->n_cones, sd->n_rects, sd->n_disks, sd->n_meshes};
  for (uint32_t i = 0; i < sd->n_materials; i++)
    if (!valid_material(sd->materials[i])) { set_error("unknown material kind"); return PB2_E_UNSUPPORTED; }

  std::vector<DGeom> geoms(sd->n_geometries);
  ctx->needs_rng_in_shading = false;
  for (uint32_t i = 0; i < sd->n_geometries; i++) {
    const pb2_geometry& g = sd->geometries[i];
    if (g.shape_kind < 0 || g.shape_kind > PB2_SHAPE_MESH) {
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
    if (m.kind == PB2_MAT_UBER || m.kind == PB2_MAT_SUBSURFACE) ctx->needs_rng_in_shading = true;
    geoms[i] = DGeom{g.shape_kind, g.shape_index, g.material, -1};
  }
  std::vector<int32_t> lights(sd->light_geometries, sd->light_geometries + sd->n_lights);
  for (uint32_t l = 0; l < sd->n_lights; l++) {
    if (lights[l] < 0 || uint32_t(lights[l]) >= sd->n_geometries) { set_error("light index out of range"); return PB2_E_INVALID; }
    if (sd->materials[sd->geometries[lights[l]].material].kind != PB2_MAT_EMISSIVE) {
      set_error("light geometry is not emissive");
      return PB2_E_INVALID;
    }
    geoms[lights[l]].light = int32_t(l);
  }

  // ---- per-mesh trees (bvh.cpp:460-476): meshes with >=1 triangle, in geometry order ----
  std::vector<DMesh> meshes(sd->n_meshes);
  std::vector<float> verts, normals, uvs;
  std::vector<uint32_t> tris;
  bool any_normals = false, any_uvs = false;
  for (uint32_t k = 0; k < sd->n_meshes; k++) {
    any_normals |= sd->meshes[k].normals != nullptr;
    any_uvs |= sd->meshes[k].texcoords != nullptr;
  }
  ctx->mesh_trees.clear();
  std::vector<int> mesh_tree_of(sd->n_meshes, -1);
  std::vector<BuildPrim> top;
  std::vector<DTopPrim> top_prims;
  for (uint32_t k = 0; k < sd->n_meshes; k++) {
    const pb2_mesh& m = sd->meshes[k];
    DMesh& dm = meshes[k];
    std::memset(&dm, 0, sizeof(dm));
    dm.tree = DTree{-1, 0, 0, -1};
    dm.vert_base = uint32_t(verts.size() / 3);
    dm.tri_base = uint32_t(tris.size() / 3);
    dm.n_tris = m.n_triangles;
    dm.has_normals = m.normals != nullptr;
    dm.has_uvs = m.texcoords != nullptr;
    for (uint32_t t = 0; t < m.n_triangles * 3; t++)
      if (m.indices[t] >= m.n_vertices) { set_error("mesh index out of range"); return PB2_E_INVALID; }
    verts.insert(verts.end(), m.vertices, m.vertices + size_t(m.n_vertices) * 3);
    tris.insert(tris.end(), m.indices, m.indices + size_t(m.n_triangles) * 3);
    if (any_normals) {
      if (m.normals) normals.insert(normals.end(), m.normals, m.normals + size_t(m.n_vertices) * 3);
      else normals.resize(normals.size() + size_t(m.n_vertices) * 3, 0.0f);
    }
    if (any_uvs) {
      if (m.texcoords) uvs.insert(uvs.end(), m.texcoords, m.texcoords + size_t(m.n_vertices) * 2);
      else uvs.resize(uvs.size() + size_t(m.n_vertices) * 2, 0.0f);
    }
  }
  for (uint32_t i = 0; i < sd->n_geometries; i++) {
    const pb2_geometry& g = sd->geometries[i];
    if (g.shape_kind != PB2_SHAPE_MESH) continue;
    const pb2_mesh& m = sd->meshes[g.shape_index];
    if (m.n_triangles == 0) continue;
    if (mesh_tree_of[g.shape_index] != -1) { set_error("a mesh may back only one geometry"); return PB2_E_INVALID; }
    std::vector<BuildPrim> prims(m.n_triangles);
    for (uint32_t t = 0; t < m.n_triangles; t++) {
      const uint32_t* f = m.indices + 3 * size_t(t);
      prims[t].box = aabb_of_tri(ld3(m.vertices + 3 * size_t(f[0])), ld3(m.vertices + 3 * size_t(f[1])),
                                 ld3(m.vertices + 3 * size_t(f[2])));
      prims[t].index = int(t);
    }
    mesh_tree_of[g.shape_index] = int(ctx->mesh_trees.size());
    ctx->mesh_trees.emplace_back();
    build_sah_binary(std::move(prims), ctx->mesh_trees.back());
    BuildPrim tp;
    tp.box = ctx->mesh_trees.back().bounds();
    tp.index = int(top.size());
    top.push_back(tp);
    top_prims.push_back(DTopPrim{int32_t(i), PB2_SHAPE_MESH, g.shape_index, 0});
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
    }
    tp.index = int(top.size());
    top.push_back(tp);
    top_prims.push_back(DTopPrim{int32_t(i), g.shape_kind, g.shape_index, 0});
  }
  build_sah_binary(top, ctx->top_tree);

  // ---- device layouts ----
  LayoutOut lay;
  DTree top_tree = append_tree(ctx->top_tree, lay);
  std::vector<int> geom_of_mesh(sd->n_meshes, -1);
  for (uint32_t i = 0; i < sd->n_geometries; i++)
    if (sd->geometries[i].shape_kind == PB2_SHAPE_MESH) geom_of_mesh[sd->geometries[i].shape_index] = int(i);
  for (uint32_t k = 0; k < sd->n_meshes; k++)
    if (mesh_tree_of[k] >= 0)
      meshes[k].tree = append_tree(ctx->mesh_trees[mesh_tree_of[k]], lay, &sd->meshes[k], geom_of_mesh[k]);
  ctx->counters.bvh2_nodes = uint32_t(lay.nodes2.size());
  ctx->counters.wide_nodes = uint32_t(lay.nodes8.size());
  ctx->counters.n_primitives = uint32_t(lay.leaf_prims.size());
  ctx->counters.bvh_build_ms =
      std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  ctx->host_scene_ready = true;
  if (ctx->device < 0) return PB2_OK;  // host-only context: builder inspection, no device work

#define UP(buf, vec) PB2_CUDA_CHECK(ctx->buf.upload(vec))
  UP(nodes2, lay.nodes2); UP(nodes8, lay.nodes8); UP(leaves, lay.leaves); UP(leaf_tris, lay.leaf_tris); UP(leaf_prims, lay.leaf_prims);
  UP(top_prims, top_prims); UP(meshes, meshes); UP(geoms, geoms); UP(lights, lights);
  UP(verts, verts); UP(tris, tris); UP(normals, normals); UP(uvs, uvs);
#undef UP
#define UPA(buf, ptr, n, T) \
  PB2_CUDA_CHECK(ctx->buf.upload(std::vector<T>(ptr, ptr + n)))
  UPA(materials, sd->materials, sd->n_materials, pb2_material);
  UPA(aabbs, sd->aabbs, sd->n_aabbs, pb2_aabb);
  UPA(obbs, sd->obbs, sd->n_obbs, pb2_obb);
  UPA(spheres, sd->spheres, sd->n_spheres, pb2_sphere);
  UPA(triangles, sd->triangles, sd->n_triangles, pb2_triangle);
  UPA(cones, sd->cones, sd->n_cones, pb2_cone);
  UPA(rects, sd->rects, sd->n_rects, pb2_rect);
  UPA(disks, sd->disks, sd->n_disks, pb2_disk);
#undef UPA

  DScene& s = ctx->scene;
  s.nodes2 = ctx->nodes2.as<DNode2>(); s.nodes8 = ctx->nodes8.as<DNode8>();
  s.leaves = ctx->leaves.as<DLeaf>(); s.leaf_tris = ctx->leaf_tris.as<DTri>(); s.leaf_prims = ctx->leaf_prims.as<int32_t>();
  s.top = top_tree;
  s.top_prims = ctx->top_prims.as<DTopPrim>(); s.meshes = ctx->meshes.as<DMesh>();
  s.geoms = ctx->geoms.as<DGeom>(); s.materials = ctx->materials.as<pb2_material>();
  s.lights = ctx->lights.as<int32_t>();
  s.n_lights = int32_t(sd->n_lights); s.n_geoms = int32_t(sd->n_geometries);
  s.aabbs = ctx->aabbs.as<pb2_aabb>(); s.obbs = ctx->obbs.as<pb2_obb>();
  s.spheres = ctx->spheres.as<pb2_sphere>(); s.triangles = ctx->triangles.as<pb2_triangle>();
  s.cones = ctx->cones.as<pb2_cone>(); s.rects = ctx->rects.as<pb2_rect>();
  s.disks = ctx->disks.as<pb2_disk>();
  s.verts = ctx->verts.as<float>(); s.tris = ctx->tris.as<uint32_t>();
  s.normals = ctx->normals.as<float>(); s.uvs = ctx->uvs.as<float>();
  s.camera = sd->camera;
  ctx->n_pixels = sd->camera.width * sd->camera.height;

  ctx->counters.bvh_build_ms =
      std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
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
