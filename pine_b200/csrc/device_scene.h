// device_scene.h -- layout of the scene in HBM as the kernels see it.
//
// Everything is flat arrays of 16-byte-aligned records. The reference-order binary BVH keeps the
// information content of one reference node (two child boxes + two child links, bvh.h:21-23) in
// one 64-byte record = four 128-bit loads. All trees (top level + one per mesh) share one node
// array; `DTree` names a root.
#pragma once
#include <cstdint>

#include "../../include/pine_b200.h"
#include "pmath.h"

namespace pb2 {

struct alignas(16) DNode2 {
  float box[2][6];  // child c: lower.xyz, upper.xyz (the `p` array of AABB::hit, bbox.h:60)
  int32_t link[2];  // internal child: node index; leaf child: offset into DScene::leaf_prims
  int32_t count[2]; // 0 = internal child, >0 = leaf child with that many primitives
};
static_assert(sizeof(DNode2) == 64, "DNode2 must be 64 bytes");

struct DTree {
  int32_t root;        // node index, or -1 for an empty tree
  int32_t root_count;  // >0: the root itself is a leaf (single-leaf tree, bvh.cpp:398-400)
  int32_t root_link;   // leaf offset when root_count > 0
  int32_t wide_root;   // root of the wide tree for the same primitives (-1 if none)
};

struct DTopPrim {      // one primitive of the top-level tree (bvh.cpp:478-493)
  int32_t geom;        // index into Scene::geometries == BVH::indices[i]
  int32_t kind;        // pb2_shape_kind
  int32_t index;       // index into the per-kind array, or mesh id
  int32_t pad;
};

struct DMesh {
  DTree tree;
  uint32_t vert_base, tri_base;  // into DScene::verts (x3 floats) / tris (x3 u32, mesh-local ids)
  uint32_t n_tris;
  uint32_t has_normals, has_uvs;
  uint32_t pad[3];
};

struct DGeom {
  int32_t kind, index, material, light;  // light = index in the light list or -1
};

// ---- wide BVH (production traversal) -------------------------------------------------------------
// 8-wide node, child boxes quantised to 8 bits per plane relative to the node's own box, rounded
// OUTWARD (conservative). 96 bytes = six 128-bit loads. Leaves reference a DLeaf: the reference
// leaf's exact fp32 box (the gate the reference applies, bvh.cpp:409-418) + its primitive range.
struct alignas(16) DNode8 {
  float origin[3];      // lower corner of the node box
  uint8_t exp[3];       // per-axis scale = 2^(exp-127): child plane = origin + q * scale
  uint8_t n_children;
  uint8_t qlo[3][8];    // quantised lower planes, axis-major
  uint8_t qhi[3][8];    // quantised upper planes
  int32_t child[8];     // >=0: DNode8 index; <0: ~leaf_index into DScene::leaves
};
static_assert(sizeof(DNode8) == 96, "DNode8 must be 96 bytes");

struct alignas(16) DLeaf {
  float box[6];         // exact reference leaf box
  int32_t link;         // top-level leaf: offset into DScene::leaf_prims (ids of DTopPrim);
                        // mesh leaf: offset into DScene::leaf_tris
  int32_t count;        // bit 31 set = mesh leaf; low bits = number of primitives
};
static_assert(sizeof(DLeaf) == 32, "DLeaf must be 32 bytes");

// Triangle gathered in leaf order for the wide traversal: 48 bytes = three 128-bit loads.
struct alignas(16) DTri {
  float v0[3], v1[3], v2[3];
  int32_t prim, geom, pad;
};
static_assert(sizeof(DTri) == 48, "DTri must be 48 bytes");

struct DScene {
  const DNode2* nodes2;
  const DNode8* nodes8;
  const DLeaf* leaves;
  const DTri* leaf_tris;
  const int32_t* leaf_prims;
  DTree top;
  const DTopPrim* top_prims;
  const DMesh* meshes;
  const DGeom* geoms;
  const pb2_material* materials;
  const int32_t* lights;  // geometry index per light
  int32_t n_lights, n_geoms;
  // per-kind shape arrays (members of the reference structs, include/pine_b200.h)
  const pb2_aabb* aabbs;
  const pb2_obb* obbs;
  const pb2_sphere* spheres;
  const pb2_triangle* triangles;
  const pb2_cone* cones;
  const pb2_rect* rects;
  const pb2_disk* disks;
  // mesh pools
  const float* verts;
  const uint32_t* tris;
  const float* normals;
  const float* uvs;
  // sampler tables for the active spp variant
  const uint8_t* sobol;
  const uint8_t* scramble;
  const uint8_t* rank;
  pb2_camera camera;
};

struct Ray {
  V3 o, d;
  float tmin, tmax;
};

struct Hit {
  int32_t geom, prim;
};

}  // namespace pb2
