// bvh_build.cpp -- see bvh_build.h. Restatement of BVHImpl::build_sah_binned
// (reference src/pine/impl/accel/bvh.cpp:43-147), same arithmetic, same tie-breaks.
#include "bvh_build.h"

#include <utility>

namespace pb2 {
namespace {

constexpr int kBuckets = 16;

struct Builder {
  Tree2& tree;

  int make_leaf(const BuildPrim* first, const BuildPrim* last) {
    Node2 node;
    node.leaf_offset = int(tree.leaf_prims.size());
    node.leaf_count = int(last - first);
    for (const BuildPrim* p = first; p != last; ++p) tree.leaf_prims.push_back(p->index);
    for (const BuildPrim* p = first; p != last; ++p) node.box[0].extend(p->box);
    node.box[1] = node.box[0];
    tree.nodes.push_back(node);
    return int(tree.nodes.size()) - 1;
  }

  static int bucket_of(const Box3& cbounds, const BuildPrim& prim, int axis) {
    return int(kBuckets * cbounds.relative_position(prim.box.centroid(axis), axis));
  }

  int split(BuildPrim* first, BuildPrim* last, const Box3& bounds) {
    const int count = int(last - first);
    if (count == 1) return make_leaf(first, last);

    Box3 cbounds;
    for (int i = 0; i < count; i++) cbounds.extend(first[i].box.centroid());
    const float parent_area = bounds.surface_area();

    float best_cost = kFloatMax;
    int best_axis = -1, best_bucket = -1;
    for (int axis = 0; axis < 3; axis++) {
      if (cbounds.degenerated(axis)) continue;
      int n_in[kBuckets] = {};
      Box3 box_in[kBuckets];
      for (int i = 0; i < count; i++) {
        int b = pmin(bucket_of(cbounds, first[i], axis), kBuckets - 1);
        n_in[b]++;
        box_in[b].extend(first[i].box);
      }
      float cost[kBuckets - 1] = {};
      Box3 sweep;
      int n = 0;
      for (int i = 0; i < kBuckets - 1; i++) {  // left side of a split after bucket i
        sweep.extend(box_in[i]);
        n += n_in[i];
        cost[i] += n * sweep.surface_area();
      }
      sweep = Box3();
      n = 0;
      for (int i = kBuckets - 1; i >= 1; i--) {  // right side
        sweep.extend(box_in[i]);
        n += n_in[i];
        cost[i - 1] += n * sweep.surface_area();
      }
      for (int i = 0; i < kBuckets - 1; i++) cost[i] = 1.0f + cost[i] / parent_area;

      float axis_cost = kFloatMax;
      int axis_bucket = -1;
      for (int i = 0; i < kBuckets - 1; i++)
        if (cost[i] < axis_cost) {
          axis_cost = cost[i];
          axis_bucket = i;
        }
      if (axis_cost < best_cost) {
        best_cost = axis_cost;
        best_axis = axis;
        best_bucket = axis_bucket;
      }
    }

    if (best_cost > float(count)) return make_leaf(first, last);

    // psl::partition (src/psl/algorithm.h:394-402): forward scan, swap matches to the front
    BuildPrim* mid = first;
    for (BuildPrim* it = first; it != last; ++it) {
      int b = bucket_of(cbounds, *it, best_axis);
      if (b == kBuckets) b = kBuckets - 1;
      if (b <= best_bucket) {
        std::swap(*mid, *it);
        ++mid;
      }
    }

    Node2 node;
    for (BuildPrim* p = first; p != mid; ++p) node.box[0].extend(p->box);
    for (BuildPrim* p = mid; p != last; ++p) node.box[1].extend(p->box);
    node.child[0] = split(first, mid, node.box[0]);
    node.child[1] = split(mid, last, node.box[1]);
    tree.nodes.push_back(node);
    return int(tree.nodes.size()) - 1;
  }
};

}  // namespace

void build_sah_binary(std::vector<BuildPrim> prims, Tree2& out) {
  out = Tree2();
  if (prims.empty()) return;
  Box3 bounds;
  for (auto& p : prims) bounds.extend(p.box);
  out.nodes.reserve(prims.size());
  out.leaf_prims.reserve(prims.size());
  Builder b{out};
  b.split(prims.data(), prims.data() + prims.size(), bounds);
  out.root = int(out.nodes.size()) - 1;
}

}  // namespace pb2
