// bvh_layout.cpp -- device layouts of the host trees (see bvh_layout.h / device_scene.h).
//
// Wide tree. The reference's two-level structure (top tree over {mesh, analytic shape}, one tree
// per mesh: bvh.cpp:453-495) is viewed as ONE tree and collapsed top-down into 8-wide nodes, always
// opening the open child with the largest surface area. Every wide child box therefore CONTAINS
// the reference boxes of everything below it, and the kernel's (conservative) slab test passes
// whenever the reference's test of any of those boxes would: the wide walk can only test MORE
// primitives than the reference does, never fewer. What the reference additionally requires -- a
// primitive is only tested when its own leaf's exact box passes (bvh.cpp:409-418) -- is kept as a
// per-primitive `gate` (DLeaf) that the kernel checks for the winner only (traverse.cuh).
// Leaf children hold at most 3 primitives (unary count in `meta`); bigger reference leaves are
// spread over several slots with the same gate.
#include "bvh_layout.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace pb2 {

namespace {
// Runs fn(0..n-1) on up to `threads` host threads (dynamic distribution).
template <class F>
void parallel_for_index(size_t n, int threads, F&& fn) {
  if (n == 0) return;
  std::atomic<size_t> next{0};
  auto work = [&] { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads && size_t(t) < n; t++) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
}
int count_internal(const Tree2& t) {
  int n = 0;
  for (const Node2& nd : t.nodes) n += nd.leaf_count == 0;
  return n;
}
// Writes tree `t` into nodes2[node_base ...] / leaf_prims[leaf_base ...] (ranges sized by the caller).
// Internal nodes only get a DNode2; children that are leaves are folded into the parent's link/count.
// Node indices are remapped densely over internal nodes.
DTree fill_binary(const Tree2& t, int node_base, int leaf_base, DNode2* nodes2, int32_t* leaf_prims) {
  DTree d{-1, 0, 0, 0};
  if (t.root < 0) return d;
  std::memcpy(leaf_prims + leaf_base, t.leaf_prims.data(), t.leaf_prims.size() * sizeof(int32_t));
  std::vector<int> remap(t.nodes.size(), -1);
  int n_internal = 0;
  for (size_t i = 0; i < t.nodes.size(); i++)
    if (!t.nodes[i].leaf_count) remap[i] = node_base + n_internal++;
  for (size_t i = 0; i < t.nodes.size(); i++) {
    const Node2& nd = t.nodes[i];
    if (nd.leaf_count) continue;
    DNode2& o = nodes2[remap[i]];
    for (int c = 0; c < 2; c++) {
      for (int k = 0; k < 3; k++) { o.box[c][k] = nd.box[c].lower[k]; o.box[c][3 + k] = nd.box[c].upper[k]; }
      const Node2& ch = t.nodes[nd.child[c]];
      if (ch.leaf_count) { o.link[c] = leaf_base + ch.leaf_offset; o.count[c] = ch.leaf_count; }
      else { o.link[c] = remap[nd.child[c]]; o.count[c] = 0; }
    }
  }
  const Node2& root = t.nodes[t.root];
  if (root.leaf_count) {
    d.root = -1;
    d.root_count = root.leaf_count;
    d.root_link = leaf_base + root.leaf_offset;
  } else {
    d.root = remap[t.root];
  }
  return d;
}
}  // namespace

DTree append_binary(const Tree2& t, LayoutOut& out) {
  if (t.root < 0) return DTree{-1, 0, 0, 0};
  const int node_base = int(out.nodes2.size()), leaf_base = int(out.leaf_prims.size());
  out.nodes2.resize(size_t(node_base) + count_internal(t));
  out.leaf_prims.resize(size_t(leaf_base) + t.leaf_prims.size());
  return fill_binary(t, node_base, leaf_base, out.nodes2.data(), out.leaf_prims.data());
}

std::vector<DTree> append_binary_many(const std::vector<const Tree2*>& trees, LayoutOut& out, int threads) {
  const size_t n = trees.size();
  std::vector<int> n_internal(n, 0);
  parallel_for_index(n, threads, [&](size_t i) { if (trees[i]->root >= 0) n_internal[i] = count_internal(*trees[i]); });
  std::vector<int> node_base(n), leaf_base(n);
  size_t nodes = out.nodes2.size(), leaves = out.leaf_prims.size();
  for (size_t i = 0; i < n; i++) {
    node_base[i] = int(nodes);
    leaf_base[i] = int(leaves);
    if (trees[i]->root < 0) continue;
    nodes += size_t(n_internal[i]);
    leaves += trees[i]->leaf_prims.size();
  }
  out.nodes2.resize(nodes);
  out.leaf_prims.resize(leaves);
  std::vector<DTree> roots(n, DTree{-1, 0, 0, 0});
  parallel_for_index(n, threads, [&](size_t i) {
    roots[i] = fill_binary(*trees[i], node_base[i], leaf_base[i], out.nodes2.data(), out.leaf_prims.data());
  });
  return roots;
}

namespace {

enum HandleKind : int { H_TOP_NODE, H_TOP_LEAF, H_MESH_NODE, H_MESH_LEAF, H_SHAPE };
struct Handle {
  int kind = H_TOP_NODE;
  int tree = -1;   // mesh tree index (H_MESH_*)
  int node = -1;   // node index in its tree; H_SHAPE: top primitive id
  int lo = 0, hi = 0;  // H_TOP_LEAF / H_MESH_LEAF: sub-range of the leaf's primitives
  int gate = -1;   // DLeaf index of the reference leaf that gates the terminals below (-1: none)
};
struct Slot {
  Handle h;
  Box3 box;
};

struct Frag {
  raw_vector<DNode8> nodes;
  raw_vector<DPrim> prims;
};

struct Wide {
  const WideInput& in;
  std::vector<int> top_leaf_gate;                // top node -> DLeaf (leaves only)
  std::vector<std::vector<int>> mesh_leaf_gate;  // tree -> node -> DLeaf (leaves only)
  std::vector<int> tree_mesh, tree_geom;         // mesh tree -> mesh id / geometry index
  std::vector<char> top_unbounded;               // top node -> a Cone somewhere below it

  // Shapes that can be hit outside their own box (traverse.cuh (iv)).
  // shapes that can be hit outside their own BVH box: a Cone below its base, a Plane beyond +-100, a Cylinder
  // beyond the radius-r box around p0 the reference gives it (geometry.h:141,147)
  static bool unbounded_kind(int kind) {
    return kind == PB2_SHAPE_CONE || kind == PB2_SHAPE_PLANE || kind == PB2_SHAPE_CYLINDER;
  }
  void mark_unbounded() {
    const Tree2& t = top();
    top_unbounded.assign(t.nodes.size(), 0);
    // children precede their parent or not -- do not rely on the numbering: iterate to a fixpoint
    for (size_t i = 0; i < t.nodes.size(); i++) {
      const Node2& nd = t.nodes[i];
      for (int j = 0; j < nd.leaf_count; j++)
        if (unbounded_kind((*in.top_prims)[t.leaf_prims[nd.leaf_offset + j]].kind)) top_unbounded[i] = 1;
    }
    for (bool changed = true; changed;) {
      changed = false;
      for (size_t i = 0; i < t.nodes.size(); i++) {
        const Node2& nd = t.nodes[i];
        if (nd.leaf_count || top_unbounded[i]) continue;
        if (top_unbounded[nd.child[0]] || top_unbounded[nd.child[1]]) { top_unbounded[i] = 1; changed = true; }
      }
    }
  }
  bool has_unbounded(const Handle& h) const {
    switch (h.kind) {
      case H_TOP_NODE: return top_unbounded[h.node];
      case H_TOP_LEAF: {
        const Node2& nd = top().nodes[h.node];
        for (int j = h.lo; j < h.hi; j++)
          if (unbounded_kind((*in.top_prims)[top().leaf_prims[nd.leaf_offset + j]].kind)) return true;
        return false;
      }
      case H_SHAPE: return unbounded_kind((*in.top_prims)[h.node].kind);
      default: return false;
    }
  }

  const Tree2& top() const { return *in.top; }
  const Tree2& mtree(int k) const { return (*in.mesh_trees)[k]; }

  static bool terminal(const Handle& h) {
    return h.kind == H_SHAPE || (h.kind == H_MESH_LEAF && h.hi - h.lo <= 3);
  }
  static int fanout(const Handle& h) {
    switch (h.kind) {
      case H_TOP_NODE: case H_MESH_NODE: return 2;
      case H_TOP_LEAF: return std::min(8, h.hi - h.lo);
      case H_MESH_LEAF: return std::min(8, (h.hi - h.lo + 2) / 3);
      default: return 0;
    }
  }
  Box3 tri_box(int k, int tri) const {
    const pb2_mesh& m = in.meshes[tree_mesh[k]];
    const uint32_t* f = m.indices + 3 * size_t(tri);
    return aabb_of_tri(ld3(m.vertices + 3 * size_t(f[0])), ld3(m.vertices + 3 * size_t(f[1])),
                       ld3(m.vertices + 3 * size_t(f[2])));
  }
  // One top-level primitive as a slot: a mesh becomes the root of its tree, a shape a terminal.
  Slot prim_slot(int top_prim, const Box3& box, int gate) const {
    const DTopPrim& tp = (*in.top_prims)[top_prim];
    Slot s;
    s.box = box;
    if (tp.kind == PB2_SHAPE_MESH) {
      const int k = (*in.mesh_tree_of)[tp.index];
      const Tree2& t = mtree(k);
      const Node2& root = t.nodes[t.root];
      s.h.tree = k;
      s.h.node = t.root;
      if (root.leaf_count) {  // single-leaf mesh: no box test of its own (bvh.cpp:398-400)
        s.h.kind = H_MESH_LEAF; s.h.lo = 0; s.h.hi = root.leaf_count; s.h.gate = gate;
      } else {
        s.h.kind = H_MESH_NODE;
      }
    } else {
      s.h.kind = H_SHAPE; s.h.node = top_prim; s.h.gate = gate;
    }
    return s;
  }
  Slot top_leaf_slot(int leaf, int lo, int hi, const Box3& box) const {
    const Node2& nd = top().nodes[leaf];
    if (hi - lo == 1) return prim_slot(top().leaf_prims[nd.leaf_offset + lo], box, top_leaf_gate[leaf]);
    Slot s;
    s.h.kind = H_TOP_LEAF; s.h.node = leaf; s.h.lo = lo; s.h.hi = hi;
    s.box = box;
    return s;
  }
  void expand(const Handle& h, std::vector<Slot>& out) const {
    switch (h.kind) {
      case H_TOP_NODE: {
        const Node2& nd = top().nodes[h.node];
        for (int c = 0; c < 2; c++) {
          const Node2& ch = top().nodes[nd.child[c]];
          if (ch.leaf_count) out.push_back(top_leaf_slot(nd.child[c], 0, ch.leaf_count, nd.box[c]));
          else { Slot s; s.h.kind = H_TOP_NODE; s.h.node = nd.child[c]; s.box = nd.box[c]; out.push_back(s); }
        }
      } break;
      case H_TOP_LEAF: {
        const Node2& nd = top().nodes[h.node];
        const int n = h.hi - h.lo, parts = std::min(8, n);
        for (int i = 0; i < parts; i++) {
          const int a = h.lo + int(int64_t(n) * i / parts), b = h.lo + int(int64_t(n) * (i + 1) / parts);
          // A mesh stays inside its own box, but an analytic shape need not (a Cone is unbounded
          // below its base, geometry.cpp:415-446): the reference tests it whenever the LEAF's box is
          // hit, so that is the box it gets here.
          Box3 box;
          for (int j = a; j < b; j++) {
            const int prim = top().leaf_prims[nd.leaf_offset + j];
            box.extend((*in.top_prims)[prim].kind == PB2_SHAPE_MESH ? (*in.top_boxes)[prim] : nd.box[0]);
          }
          out.push_back(top_leaf_slot(h.node, a, b, box));
        }
      } break;
      case H_MESH_NODE: {
        const Tree2& t = mtree(h.tree);
        const Node2& nd = t.nodes[h.node];
        for (int c = 0; c < 2; c++) {
          const Node2& ch = t.nodes[nd.child[c]];
          Slot s;
          s.h.tree = h.tree; s.h.node = nd.child[c]; s.box = nd.box[c];
          if (ch.leaf_count) {
            s.h.kind = H_MESH_LEAF; s.h.lo = 0; s.h.hi = ch.leaf_count;
            s.h.gate = mesh_leaf_gate[h.tree][nd.child[c]];
          } else {
            s.h.kind = H_MESH_NODE;
          }
          out.push_back(s);
        }
      } break;
      case H_MESH_LEAF: {  // more than 3 triangles: chunks of <= 3 (or 8 sub-ranges of whole chunks)
        const Tree2& t = mtree(h.tree);
        const Node2& nd = t.nodes[h.node];
        const int chunks = (h.hi - h.lo + 2) / 3, parts = std::min(8, chunks);
        for (int i = 0; i < parts; i++) {
          const int a = h.lo + 3 * int(int64_t(chunks) * i / parts);
          const int b = std::min(h.hi, h.lo + 3 * int(int64_t(chunks) * (i + 1) / parts));
          Slot s;
          s.h = h; s.h.lo = a; s.h.hi = b;
          for (int j = a; j < b; j++) s.box.extend(tri_box(h.tree, t.leaf_prims[nd.leaf_offset + j]));
          out.push_back(s);
        }
      } break;
      default: break;
    }
  }

  void emit_prims(const Handle& h, raw_vector<DPrim>& prims) const {
    if (h.kind == H_SHAPE) {
      const DTopPrim& tp = (*in.top_prims)[h.node];
      DPrim d;
      std::memset(&d, 0, sizeof(d));
      std::memcpy(&d.v0[0], &tp.index, 4);
      d.prim = -1 - tp.kind; d.geom = tp.geom; d.leaf = h.gate;
      prims.push_back(d);
      return;
    }
    const Tree2& t = mtree(h.tree);
    const Node2& nd = t.nodes[h.node];
    const pb2_mesh& m = in.meshes[tree_mesh[h.tree]];
    for (int j = h.lo; j < h.hi; j++) {
      const int tri = t.leaf_prims[nd.leaf_offset + j];
      const uint32_t* f = m.indices + 3 * size_t(tri);
      DPrim d;
      std::memcpy(d.v0, m.vertices + 3 * size_t(f[0]), 12);
      std::memcpy(d.v1, m.vertices + 3 * size_t(f[1]), 12);
      std::memcpy(d.v2, m.vertices + 3 * size_t(f[2]), 12);
      d.prim = tri; d.geom = tree_geom[h.tree]; d.leaf = h.gate;
      prims.push_back(d);
    }
  }

  // Quantisation frame of one axis: scale = 2^e and an origin that is a multiple of the scale,
  // such that origin + q * scale is exactly representable for q in [0,255] and covers [lo,hi].
  static void frame(float lo, float hi, float& origin, uint8_t& exp_out) {
    int e;
    const double extent = double(hi) - double(lo);
    if (extent > 0) { std::frexp(extent / 255.0, &e); } else { e = -126; }
    // exact representability: |origin / scale| + 255 < 2^24
    const double amax = std::max(std::fabs(double(lo)), std::fabs(double(hi)));
    if (amax > 0) {
      int ea;
      std::frexp(amax, &ea);  // amax < 2^ea
      e = std::max(e, ea - 23);
    }
    e = std::min(std::max(e, -126), 127);
    for (;; e++) {
      const double scale = std::ldexp(1.0, e);
      const double o = std::floor(double(lo) / scale) * scale;
      if (o + 255.0 * scale >= double(hi) && std::fabs(o) / scale + 256.0 < 16777216.0) {
        origin = float(o);
        exp_out = uint8_t(e + 127);
        return;
      }
    }
  }

  // Builds the wide node for `slots` (collapsed already) into `node`; appends terminals to `prims`
  // and the handles of its internal children, in child order, to `kids`.
  void make_node(std::vector<Slot>& slots, DNode8& node, raw_vector<DPrim>& prims,
                 std::vector<Handle>& kids, uint32_t child_base) const {
    const int n = int(slots.size());
    Box3 bounds;
    for (auto& s : slots) bounds.extend(s.box);
    // slot assignment: child c prefers the slot whose corner direction its centroid points to
    const V3 center = bounds.centroid();
    float cost[8][8];
    for (int c = 0; c < n; c++) {
      const V3 v = slots[c].box.centroid() - center;
      for (int s = 0; s < 8; s++)
        cost[c][s] = (s & 1 ? v.x : -v.x) + (s & 2 ? v.y : -v.y) + (s & 4 ? v.z : -v.z);
    }
    int slot_of[8], child_in[8];
    for (int i = 0; i < 8; i++) { slot_of[i] = -1; child_in[i] = -1; }
    for (int round = 0; round < n; round++) {
      int bc = -1, bs = -1;
      float best = -kFloatMax;
      for (int c = 0; c < n; c++) {
        if (slot_of[c] >= 0) continue;
        for (int s = 0; s < 8; s++)
          if (child_in[s] < 0 && (bc < 0 || cost[c][s] > best)) { best = cost[c][s]; bc = c; bs = s; }
      }
      slot_of[bc] = bs;
      child_in[bs] = bc;
    }
    std::memset(&node, 0, sizeof(node));
    double scale[3];
    for (int k = 0; k < 3; k++) {
      frame(bounds.lower[k], bounds.upper[k], node.origin[k], node.exp[k]);
      scale[k] = std::ldexp(1.0, int(node.exp[k]) - 127);
    }
    node.child_base = child_base;
    node.prim_base = uint32_t(prims.size());
    for (auto& sl : slots)
      if (has_unbounded(sl.h)) node.prim_base |= 0x80000000u;  // kNoCullBit (traverse.cuh)
    int prim_offset = 0;
    for (int s = 0; s < 8; s++) {
      const int c = child_in[s];
      if (c < 0) continue;
      const Slot& sl = slots[c];
      for (int k = 0; k < 3; k++) {
        double ql = std::floor((double(sl.box.lower[k]) - double(node.origin[k])) / scale[k]);
        double qh = std::ceil((double(sl.box.upper[k]) - double(node.origin[k])) / scale[k]);
        ql = std::min(std::max(ql, 0.0), 255.0);
        qh = std::min(std::max(qh, 0.0), 255.0);
        // the planes are exact floats by construction; make sure they enclose the box
        auto plane = [&](double q) { return float(double(node.origin[k]) + q * scale[k]); };
        while (ql > 0 && plane(ql) > sl.box.lower[k]) ql -= 1;
        while (qh < 255 && plane(qh) < sl.box.upper[k]) qh += 1;
        node.qlo[k][s] = uint8_t(ql);
        node.qhi[k][s] = uint8_t(qh);
      }
      if (terminal(sl.h)) {
        const int before = int(prims.size());
        emit_prims(sl.h, prims);
        const int cnt = int(prims.size()) - before;  // 1..3
        node.meta[s] = uint8_t((((1 << cnt) - 1) << 5) | prim_offset);
        prim_offset += cnt;
      } else {
        node.imask |= uint8_t(1 << s);
        node.meta[s] = uint8_t((1 << 5) | (24 + s));
        kids.push_back(sl.h);
      }
    }
  }

  // Slots of the wide node that represents `h`.
  void collapse(const Handle& h, std::vector<Slot>& slots) const {
    slots.clear();
    if (terminal(h)) { Slot s; s.h = h; slots.push_back(s); return; }  // only ever the root
    expand(h, slots);
    bool closed[8] = {};
    while (slots.size() < 8) {
      int best = -1;
      float best_area = -1.0f;
      for (size_t i = 0; i < slots.size(); i++) {
        if (closed[i] || terminal(slots[i].h)) continue;
        if (int(slots.size()) - 1 + fanout(slots[i].h) > 8) { closed[i] = true; continue; }
        const float a = slots[i].box.surface_area();
        if (a > best_area) { best_area = a; best = int(i); }
      }
      if (best < 0) break;
      const Handle open = slots[best].h;
      slots[best] = slots.back();
      closed[best] = closed[slots.size() - 1];
      closed[slots.size() - 1] = false;
      slots.pop_back();
      expand(open, slots);
    }
  }

  // Breadth-first build of everything below `root` into `f` (local indices, root = node 0).
  // Stops after `max_nodes` nodes have been PROCESSED (0 = no limit); the unprocessed tail of the
  // queue is returned in `pending` as (handle, local node index).
  void build_fragment(const Handle& root, Frag& f, size_t max_nodes,
                      std::vector<std::pair<Handle, uint32_t>>* pending) const {
    std::vector<Handle> queue{root};
    f.nodes.emplace_back();
    std::vector<Slot> slots;
    std::vector<Handle> kids;
    size_t head = 0;
    for (; head < queue.size(); head++) {
      if (max_nodes && head >= max_nodes) break;
      collapse(queue[head], slots);
      if (head == 0 && slots.size() == 1 && terminal(slots[0].h)) {
        // degenerate scene: a single terminal; give the root the terminal's own box
        slots[0].box = Box3();
        if (root.kind == H_SHAPE) slots[0].box = (*in.top_boxes)[root.node];
        else for (int j = root.lo; j < root.hi; j++)
          slots[0].box.extend(tri_box(root.tree, mtree(root.tree).leaf_prims[mtree(root.tree).nodes[root.node].leaf_offset + j]));
      }
      kids.clear();
      DNode8 nd;
      make_node(slots, nd, f.prims, kids, uint32_t(f.nodes.size()));
      f.nodes[head] = nd;
      for (auto& k : kids) { queue.push_back(k); f.nodes.emplace_back(); }
    }
    if (pending)
      for (; head < queue.size(); head++) pending->push_back({queue[head], uint32_t(head)});
  }
};

}  // namespace

void build_wide(const WideInput& in, LayoutOut& out, int threads) {
  const bool verbose = std::getenv("PB2_VERBOSE") != nullptr;
  auto t0 = std::chrono::steady_clock::now();
  auto stage = [&](const char* name) {
    if (verbose)
      fprintf(stderr, "[build_wide] %-24s +%.1f ms\n", name,
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  };
  out.nodes8.clear();
  out.leaves.clear();
  out.wide_prims.clear();
  const Tree2& top = *in.top;
  if (top.root < 0) return;
  threads = std::max(1, threads);
  Wide w{in};
  w.mark_unbounded();
  auto push_gate = [&](const Box3& b) {
    DLeaf L;
    for (int k = 0; k < 3; k++) { L.box[k] = b.lower[k]; L.box[3 + k] = b.upper[k]; }
    L.pad[0] = L.pad[1] = 0;
    out.leaves.push_back(L);
    return int(out.leaves.size()) - 1;
  };
  // gates: a leaf that is the ROOT of its tree is tested without a box test (bvh.cpp:398-400)
  w.top_leaf_gate.assign(top.nodes.size(), -1);
  for (size_t i = 0; i < top.nodes.size(); i++)
    if (top.nodes[i].leaf_count && int(i) != top.root) w.top_leaf_gate[i] = push_gate(top.nodes[i].box[0]);
  out.box_shapes.clear();
  for (size_t i = 0; i < top.nodes.size(); i++) {
    const Node2& nd = top.nodes[i];
    for (int j = 0; j < nd.leaf_count; j++) {
      const DTopPrim& tp = (*in.top_prims)[top.leaf_prims[nd.leaf_offset + j]];
      if (tp.kind == PB2_SHAPE_AABB || tp.kind == PB2_SHAPE_OBB)
        out.box_shapes.push_back(DBoxShape{tp.kind, tp.index, w.top_leaf_gate[i], tp.geom});
    }
  }
  const size_t n_trees = in.mesh_trees->size();
  w.tree_mesh.assign(n_trees, -1);
  w.tree_geom.assign(n_trees, -1);
  for (const DTopPrim& tp : *in.top_prims)
    if (tp.kind == PB2_SHAPE_MESH) {
      const int k = (*in.mesh_tree_of)[tp.index];
      w.tree_mesh[k] = tp.index;
      w.tree_geom[k] = tp.geom;
    }
  // the mesh trees' gates: sizes by prefix sum, then every tree fills its own range of `leaves` in parallel
  w.mesh_leaf_gate.resize(n_trees);
  {
    std::vector<size_t> first(n_trees + 1, out.leaves.size());
    parallel_for_index(n_trees, threads, [&](size_t k) {
      const Tree2& t = (*in.mesh_trees)[k];
      size_t n = 0;
      for (size_t i = 0; i < t.nodes.size(); i++) n += t.nodes[i].leaf_count && int(i) != t.root;
      first[k + 1] = n;
    });
    for (size_t k = 0; k < n_trees; k++) first[k + 1] += first[k];
    out.leaves.resize(first[n_trees]);
    parallel_for_index(n_trees, threads, [&](size_t k) {
      const Tree2& t = (*in.mesh_trees)[k];
      w.mesh_leaf_gate[k].assign(t.nodes.size(), -1);
      size_t at = first[k];
      for (size_t i = 0; i < t.nodes.size(); i++)
        if (t.nodes[i].leaf_count && int(i) != t.root) {
          const Box3& b = t.nodes[i].box[0];
          DLeaf& L = out.leaves[at];
          for (int c = 0; c < 3; c++) { L.box[c] = b.lower[c]; L.box[3 + c] = b.upper[c]; }
          L.pad[0] = L.pad[1] = 0;
          w.mesh_leaf_gate[k][i] = int(at++);
        }
    });
  }

  stage("gates");
  Handle root;
  const Node2& rn = top.nodes[top.root];
  if (rn.leaf_count) {
    Slot s = w.top_leaf_slot(top.root, 0, rn.leaf_count, rn.box[0]);
    root = s.h;
  } else {
    root.kind = H_TOP_NODE;
    root.node = top.root;
  }

  // top part sequentially (breadth first), then one fragment per pending node in parallel
  Frag head;
  std::vector<std::pair<Handle, uint32_t>> pending;
  w.build_fragment(root, head, threads > 1 ? size_t(threads) * 8 : 0, &pending);
  stage("head fragment");
  std::vector<Frag> frags(pending.size());
  if (!pending.empty()) {
    std::atomic<size_t> next{0};
    auto work = [&] {
      for (size_t i; (i = next.fetch_add(1)) < pending.size();) w.build_fragment(pending[i].first, frags[i], 0, nullptr);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
  }
  stage("fragments");
  // merge: fragment i's nodes 1.. land at node_base[i] .. (node 0 fills the slot reserved for it in the head
  // part), its primitives at prim_base[i]; offsets by prefix sum, copies in parallel
  std::vector<size_t> node_base(frags.size() + 1), prim_base(frags.size() + 1);
  node_base[0] = head.nodes.size();
  prim_base[0] = head.prims.size();
  for (size_t i = 0; i < frags.size(); i++) {
    node_base[i + 1] = node_base[i] + frags[i].nodes.size() - 1;
    prim_base[i + 1] = prim_base[i] + frags[i].prims.size();
  }
  out.nodes8 = std::move(head.nodes);
  out.wide_prims = std::move(head.prims);
  out.nodes8.resize(node_base[frags.size()]);
  out.wide_prims.resize(prim_base[frags.size()]);
  parallel_for_index(frags.size(), threads, [&](size_t i) {
    Frag& f = frags[i];
    for (size_t j = 0; j < f.nodes.size(); j++) {
      DNode8 nd = f.nodes[j];
      nd.child_base += uint32_t(node_base[i]) - 1;
      nd.prim_base += uint32_t(prim_base[i]);  // the kNoCullBit in the top bit is untouched: bases stay below 2^31
      if (j == 0) out.nodes8[pending[i].second] = nd;
      else out.nodes8[node_base[i] + j - 1] = nd;
    }
    if (!f.prims.empty()) std::memcpy(&out.wide_prims[prim_base[i]], f.prims.data(), f.prims.size() * sizeof(DPrim));
    Frag().nodes.swap(f.nodes);
    Frag().prims.swap(f.prims);
  });
  stage("merge");
}

}  // namespace pb2
