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
  int32_t pad;
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
// ONE 8-wide tree over every primitive of the scene (all mesh triangles and all analytic shapes),
// obtained by collapsing the reference-identical binary trees (top level with each mesh's tree
// hanging below it). 80-byte nodes = five 128-bit loads, in the spirit of Ylitie et al. 2017:
// child boxes are 8-bit offsets from the node's lower corner in units of a per-axis power of two,
// rounded OUTWARD; the internal children of a node are consecutive in memory (child_base + rank)
// and so are the primitive records of its leaf children (prim_base + offset); child slots are
// assigned so that `slot ^ octant` enumerates them roughly front to back for every ray octant.
// Nodes are laid out breadth first, so the top levels are one contiguous prefix (the L2 window).
struct alignas(16) DNode8 {
  float origin[3];      // lower corner of the node box
  uint8_t exp[3];       // per-axis scale = 2^(exp-127): plane = origin + q * scale (exact in fp32)
  uint8_t imask;        // bit i set: child slot i is an internal node
  uint32_t child_base;  // index of the first internal child
  uint32_t prim_base;   // index of the first DPrim of this node's leaf children; bit 31 = kNoCullBit
  uint8_t meta[8];      // internal: 001sssss (s = 24 + slot); leaf: unary count (1..3) << 5 | offset;
                        // empty slot: 0
  uint8_t qlo[3][8];    // quantised lower planes, axis-major
  uint8_t qhi[3][8];    // quantised upper planes
};
static_assert(sizeof(DNode8) == 80, "DNode8 must be 80 bytes");

// Exact fp32 box of a reference leaf: the gate the reference applies before it tests the leaf's
// primitives (bvh.cpp:409-418). Read once per ray, for the winner only (traverse.cuh).
struct alignas(16) DLeaf {
  float box[6];
  int32_t pad[2];
};
static_assert(sizeof(DLeaf) == 32, "DLeaf must be 32 bytes");

// Primitive record of the wide tree, 48 bytes = three 128-bit loads. Triangle: the three vertices
// gathered from the mesh; analytic shape: prim = -1 - shape_kind and v0[0] carries the index into
// the per-kind array.
struct alignas(16) DPrim {
  float v0[3], v1[3], v2[3];
  int32_t prim;  // triangle index inside its mesh, or -1 - pb2_shape_kind
  int32_t geom;  // Scene::geometries index
  int32_t leaf;  // DLeaf that gates this primitive in the reference, or -1 for "no gate"
};
static_assert(sizeof(DPrim) == 48, "DPrim must be 48 bytes");

// AABB/OBB shapes, tested one by one by the wide any-hit walk: a scaled Box compares a LOCAL-space
// distance with the world-space tmax (bbox.cpp:149-161), so it can report a hit although the ray
// segment [tmin,tmax] never reaches its world-space box -- box culling cannot be used for it.
struct DBoxShape {
  int32_t kind, index, gate, geom;
};

struct DTexture {
  uint32_t offset;  // first texel in DScene::texels
  int32_t width, height;
};

struct DScene {
  const DNode2* nodes2;
  const DNode8* nodes8;
  const DLeaf* leaves;
  const DPrim* wide_prims;
  const int32_t* leaf_prims;
  const DBoxShape* box_shapes;
  int32_t n_nodes8;             // nodes in nodes8
  float node_origin_absmax[3];  // max |DNode8::origin| per axis: bounds |origin * inv| per ray (traverse.cuh WideRay::err)
  int32_t n_box_shapes;     // > kMaxBoxShapes: any-hit walks the reference-order tree instead
  int32_t wide_ready;       // 1: nodes8/wide_prims hold a tree (root = node 0)
  int32_t order_dependent;  // 1: more than kMaxBoxShapes AABB/OBB shapes, whose closest-hit result
                            // depends on the visiting order (bbox.cpp:149-161) => every closest
                            // hit walks the reference-order tree (traverse.cuh (iii))
  DTree top;
  const DTopPrim* top_prims;
  const DMesh* meshes;
  const DGeom* geoms;
  const pb2_material* materials;
  const int32_t* lights;  // geometry index per light (-1: not an area light, see light_recs)
  const pb2_light* light_recs;  // UniformLightSampler::lights, same order
  int32_t n_lights, n_geoms;
  int32_t env_kind;       // 0, or PB2_LIGHT_SKY: Scene::env_light
  float env_color[3];     // Sky::sun_color
  // per-kind shape arrays (members of the reference structs, include/pine_b200.h)
  const pb2_aabb* aabbs;
  const pb2_obb* obbs;
  const pb2_sphere* spheres;
  const pb2_triangle* triangles;
  const pb2_cone* cones;
  const pb2_rect* rects;
  const pb2_disk* disks;
  const pb2_plane* planes;
  const pb2_cylinder* cylinders;
  // mesh pools
  const float* verts;
  const uint32_t* tris;
  const float* normals;
  const float* uvs;
  // image nodes of Uber materials (pb2_texture): texel pool, per-texture {offset, w, h}, and the host-made
  // table byte -> pow(byte / 255, 2.2) (Image::operator[], image.cpp:7-18)
  const uint32_t* texels;  // r | g << 8 | b << 16 | a << 24
  const DTexture* textures;
  const float* srgb_lut;
  // postfix programs of the materials' *_node parameters (pb2_node_instr = one float4: op bits, a[0..2])
  const pb2_node_instr* node_instrs;
  // sampler tables for the active spp variant
  const uint8_t* sobol;
  const uint8_t* scramble;
  const uint8_t* rank;
  // sampler of the current render (PB2_SAMPLER_*); SobolSampler::log2_spp / nbase4_digits (sampler.h:128-130,
  // sampler.cpp:82-85)
  int sampler_kind, sampler_seed, log2_spp, nbase4_digits;
  // HaltonSampler tables (sampling.cuh HaltonTables) and multInverse[0..1] (sampler.cpp:61-62)
  const int32_t* halton_primes;
  const int32_t* halton_sums;
  const uint16_t* halton_perms;
  int halton_inv0, halton_inv1;
  int shape_mask;       // OR of 1 << pb2_shape_kind over the geometries (Mesh: non-empty ones): picks the traversal
                        // kernels' instantiation (shapes.cuh kShapes*)
  int has_rare_shapes;  // a Plane or a Cylinder somewhere (the reference-order kernels' two-way switch)
  int has_meshes;       // some geometry is a non-empty Mesh
  pb2_camera camera;
  unsigned long long* walk_stats;  // PB2_COUNT builds only: {closest nodes, closest prims, any nodes, any prims,
                                   // closest rays, any rays}; NULL otherwise
};

constexpr int kMaxBoxShapes = 16;

struct Ray {
  V3 o, d;
  float tmin, tmax;
};

struct Hit {
  int32_t geom, prim;
};

}  // namespace pb2
