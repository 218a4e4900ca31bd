// bvh_build.h -- host BVH construction for the CUDA backend.
//
// Stage 1 (build_sah_binary): restates the reference's top-down 16-bucket binned-SAH builder
// (src/pine/impl/accel/bvh.cpp:30-147, BVHImpl::build / build_sah_binned) so that the tree --
// split decisions, child boxes, post-order node numbering, leaf grouping and the order of the
// primitive indices inside each leaf -- is identical node for node. Traversal results can only be
// bit-identical to the reference if the same leaves gate the same primitives (bvh.cpp:123-124).
// Stage 2 (flatten to device layouts) lives in bvh_layout.h.
#pragma once
#include <cstdint>
#include <vector>

#include "host_shapes.h"

namespace pb2 {

struct BuildPrim {
  Box3 box;
  int index = 0;
};

struct Node2 {  // one reference BVHImpl::Node (bvh.h:14-30), without the optimiser's bookkeeping
  Box3 box[2];
  int child[2] = {-1, -1};
  int leaf_offset = 0, leaf_count = 0;  // into Tree2::leaf_prims; leaf_count > 0 <=> leaf
};

struct Tree2 {
  std::vector<Node2> nodes;
  std::vector<int> leaf_prims;
  int root = -1;
  Box3 bounds() const { return box_union(nodes[root].box[0], nodes[root].box[1]); }
};

void build_sah_binary(std::vector<BuildPrim> prims, Tree2& out);
// Extra threads build_sah_binary may use for big trees (process-wide budget; the result does not
// depend on it). pb2_upload_scene sets it to the number of idle hardware threads.
void set_build_threads(int n);

}  // namespace pb2
