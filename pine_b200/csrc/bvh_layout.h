// bvh_layout.h -- turn host Tree2 trees into the two device layouts of device_scene.h.
#pragma once
#include <vector>

#include "bvh_build.h"
#include "device_scene.h"

namespace pb2 {

struct LayoutOut {
  std::vector<DNode2> nodes2;
  std::vector<DNode8> nodes8;
  std::vector<DLeaf> leaves;
  std::vector<DTri> leaf_tris;
  std::vector<int32_t> leaf_prims;
};

// Appends `t` (node indices and leaf offsets rebased) and its wide collapse; returns the roots.
// For a mesh tree pass the mesh (vertices/indices) and its geometry index so that triangles are
// gathered in leaf order; for the top-level tree pass mesh = nullptr.
DTree append_tree(const Tree2& t, LayoutOut& out, const pb2_mesh* mesh = nullptr, int geom = -1);

}  // namespace pb2
