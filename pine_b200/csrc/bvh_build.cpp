// bvh_build.cpp -- see bvh_build.h. Restatement of BVHImpl::build_sah_binned
// (reference src/pine/impl/accel/bvh.cpp:43-147), same arithmetic, same tie-breaks.
//
// Parallelism: the recursion is a pure function of its primitive range, so big ranges hand their
// left half to another thread. Each half is built into its own fragment (nodes numbered from 0 in
// the reference's post-order) and the fragments are concatenated left, right, parent -- which is
// exactly the numbering the sequential recursion produces, so the tree is identical node for node
// whatever the thread count.
#include "bvh_build.h"

#include <atomic>
#include <future>
#include <thread>
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

  // Decides the split of [first,last) (bvh.cpp:50-131). Returns nullptr for "make a leaf",
  // else the partition point, with the two child boxes filled in.
  BuildPrim* choose_split(BuildPrim* first, BuildPrim* last, const Box3& bounds, Box3 (&child_box)[2]) {
    const int count = int(last - first);
    if (count == 1) return nullptr;

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

    if (best_cost > float(count)) return nullptr;

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

    for (BuildPrim* p = first; p != mid; ++p) child_box[0].extend(p->box);
    for (BuildPrim* p = mid; p != last; ++p) child_box[1].extend(p->box);
    return mid;
  }

  int split(BuildPrim* first, BuildPrim* last, const Box3& bounds) {
    Node2 node;
    BuildPrim* mid = choose_split(first, last, bounds, node.box);
    if (!mid) return make_leaf(first, last);
    node.child[0] = split(first, mid, node.box[0]);
    node.child[1] = split(mid, last, node.box[1]);
    tree.nodes.push_back(node);
    return int(tree.nodes.size()) - 1;
  }
};

constexpr int kParallelMin = 1 << 15;  // ranges below this are built by the calling thread
std::atomic<int> g_spare_threads{0};   // threads the builder may still start

void append_fragment(Tree2& dst, Tree2&& src) {
  const int node_base = int(dst.nodes.size()), prim_base = int(dst.leaf_prims.size());
  dst.leaf_prims.insert(dst.leaf_prims.end(), src.leaf_prims.begin(), src.leaf_prims.end());
  dst.nodes.reserve(dst.nodes.size() + src.nodes.size() + 1);
  for (Node2 nd : src.nodes) {
    if (nd.leaf_count) nd.leaf_offset += prim_base;
    else { nd.child[0] += node_base; nd.child[1] += node_base; }
    dst.nodes.push_back(nd);
  }
}

// Builds [first,last) into `out` (which must be empty); the subtree root is out.nodes.back().
void split_parallel(BuildPrim* first, BuildPrim* last, const Box3& bounds, Tree2& out) {
  Builder b{out};
  if (last - first < kParallelMin) { b.split(first, last, bounds); return; }
  Node2 node;
  BuildPrim* mid = b.choose_split(first, last, bounds, node.box);
  if (!mid) { b.make_leaf(first, last); return; }
  Tree2 right;
  if (g_spare_threads.fetch_sub(1) > 0) {
    auto fut = std::async(std::launch::async, [&] { split_parallel(first, mid, node.box[0], out); });
    split_parallel(mid, last, node.box[1], right);
    fut.get();
    g_spare_threads.fetch_add(1);
  } else {
    g_spare_threads.fetch_add(1);
    split_parallel(first, mid, node.box[0], out);
    split_parallel(mid, last, node.box[1], right);
  }
  node.child[0] = int(out.nodes.size()) - 1;
  append_fragment(out, std::move(right));
  node.child[1] = int(out.nodes.size()) - 1;
  out.nodes.push_back(node);
}

}  // namespace

void set_build_threads(int n) { g_spare_threads.store(n > 1 ? n - 1 : 0); }

void build_sah_binary(std::vector<BuildPrim> prims, Tree2& out) {
  out = Tree2();
  if (prims.empty()) return;
  Box3 bounds;
  for (auto& p : prims) bounds.extend(p.box);
  if (prims.size() < size_t(kParallelMin)) {
    out.nodes.reserve(prims.size());
    out.leaf_prims.reserve(prims.size());
  }
  split_parallel(prims.data(), prims.data() + prims.size(), bounds, out);
  out.root = int(out.nodes.size()) - 1;
}

}  // namespace pb2
