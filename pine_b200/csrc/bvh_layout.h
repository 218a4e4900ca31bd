// bvh_layout.h -- turn the host's reference-identical binary trees into the two device layouts of
// device_scene.h: the binary nodes walked in reference order (verification / fallback / scenes
// whose closest hit is order dependent) and the single 8-wide compressed tree used in production.
#pragma once
#include <memory>
#include <utility>
#include <vector>

#include "bvh_build.h"
#include "device_scene.h"

namespace pb2 {

// std::vector that leaves trivially-constructible elements UNINITIALISED on resize(): the layout arrays are
// hundreds of MB that the builder's threads overwrite completely, so zero-filling them first on one thread
// (and faulting their pages in there) was a quarter of the layout time.
template <class T>
struct default_init_allocator : std::allocator<T> {
  template <class U> struct rebind { using other = default_init_allocator<U>; };
  default_init_allocator() = default;
  template <class U> default_init_allocator(const default_init_allocator<U>&) {}
  template <class U> void construct(U* p) { ::new (static_cast<void*>(p)) U; }
  template <class U, class... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};
template <class T> using raw_vector = std::vector<T, default_init_allocator<T>>;

struct LayoutOut {
  raw_vector<DNode2> nodes2;       // every binary tree (top level + one per mesh)
  raw_vector<int32_t> leaf_prims;  // primitive ids of the binary leaves
  raw_vector<DNode8> nodes8;       // the wide tree, breadth first, root = node 0
  raw_vector<DLeaf> leaves;        // reference leaf boxes (gates)
  raw_vector<DPrim> wide_prims;    // primitive records of the wide tree
  std::vector<DBoxShape> box_shapes;  // AABB/OBB shapes with their gates (device_scene.h)
};

// Appends binary tree `t` (node indices and leaf offsets rebased); returns its root descriptor.
DTree append_binary(const Tree2& t, LayoutOut& out);
// The same for a list of trees, in order, on up to `threads` host threads (sizes by prefix sum, then every
// tree fills its own range): identical arrays to appending them one after the other.
std::vector<DTree> append_binary_many(const std::vector<const Tree2*>& trees, LayoutOut& out, int threads);

struct WideInput {
  const Tree2* top = nullptr;                     // top-level tree over `top_prims`
  const std::vector<DTopPrim>* top_prims = nullptr;
  const std::vector<Box3>* top_boxes = nullptr;   // the box BVH::build gave each top primitive
  const std::vector<Tree2>* mesh_trees = nullptr; // per-mesh trees ...
  const std::vector<int>* mesh_tree_of = nullptr; // ... indexed through mesh id -> tree (or -1)
  const pb2_mesh* meshes = nullptr;               // vertex / index data of every mesh
};
// Collapses top tree + mesh trees into out.nodes8 / out.leaves / out.wide_prims using up to
// `threads` host threads (the result does not depend on the thread count).
void build_wide(const WideInput& in, LayoutOut& out, int threads);

}  // namespace pb2
