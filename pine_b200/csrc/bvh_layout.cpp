// bvh_layout.cpp -- device layouts of a host Tree2 (see bvh_layout.h / device_scene.h).
//
// Wide tree: the binary tree is collapsed top-down, always opening the internal child with the
// largest surface area until eight slots are used. Reference leaves are never merged or split --
// each becomes one DLeaf carrying its exact fp32 box, because in the reference a primitive is only
// tested when its own leaf's box passes (bvh.cpp:409-418); inner boxes are only ever looser.
// Child boxes are stored as 8-bit offsets from the node's lower corner in units of a per-axis
// power of two, rounded outward and verified with the same float expression the kernel evaluates.
#include "bvh_layout.h"

#include <cmath>
#include <cstring>

namespace pb2 {
namespace {

struct Slot {
  Box3 box;
  int node;       // binary node index (internal) or -1
  int leaf_node;  // binary leaf node index or -1
};

inline float plane(float origin, uint8_t q, uint8_t e) {
  return origin + float(q) * u2f(uint32_t(e) << 23);
}

struct WideBuilder {
  const Tree2& t;
  LayoutOut& out;
  const pb2_mesh* mesh;
  int geom;
  int leaf_base;  // rebased offset of this tree's leaf_prims in out.leaf_prims
  std::vector<int> leaf_of_node;  // binary leaf node -> DLeaf index

  int make_leaf(int bnode, const Box3* gate) {
    if (leaf_of_node[bnode] >= 0) return leaf_of_node[bnode];
    const Node2& nd = t.nodes[bnode];
    DLeaf L;
    Box3 b = gate ? *gate : nd.box[0];
    for (int k = 0; k < 3; k++) { L.box[k] = b.lower[k]; L.box[3 + k] = b.upper[k]; }
    if (mesh) {
      L.link = int32_t(out.leaf_tris.size());
      L.count = int32_t(uint32_t(nd.leaf_count) | 0x80000000u);
      for (int i = 0; i < nd.leaf_count; i++) {
        int tri = t.leaf_prims[nd.leaf_offset + i];
        const uint32_t* f = mesh->indices + 3 * size_t(tri);
        DTri d;
        std::memcpy(d.v0, mesh->vertices + 3 * size_t(f[0]), 12);
        std::memcpy(d.v1, mesh->vertices + 3 * size_t(f[1]), 12);
        std::memcpy(d.v2, mesh->vertices + 3 * size_t(f[2]), 12);
        d.prim = tri; d.geom = geom; d.pad = 0;
        out.leaf_tris.push_back(d);
      }
    } else {
      L.link = leaf_base + nd.leaf_offset;
      L.count = nd.leaf_count;
    }
    out.leaves.push_back(L);
    return leaf_of_node[bnode] = int(out.leaves.size()) - 1;
  }

  // Builds the wide node for binary internal node `bnode`; returns its DNode8 index.
  int build(int bnode) {
    Slot slots[8];
    int n = 0;
    auto push_children = [&](int b) {
      const Node2& nd = t.nodes[b];
      for (int c = 0; c < 2; c++) {
        const Node2& ch = t.nodes[nd.child[c]];
        Slot s;
        s.box = nd.box[c];
        s.node = ch.leaf_count ? -1 : nd.child[c];
        s.leaf_node = ch.leaf_count ? nd.child[c] : -1;
        slots[n++] = s;
      }
    };
    push_children(bnode);
    while (n < 8) {
      int best = -1;
      float best_area = -1.0f;
      for (int i = 0; i < n; i++)
        if (slots[i].node >= 0) {
          float a = slots[i].box.surface_area();
          if (a > best_area) { best_area = a; best = i; }
        }
      if (best < 0) break;
      int b = slots[best].node;
      slots[best] = slots[n - 1];
      n--;
      push_children(b);
    }
    int idx = int(out.nodes8.size());
    out.nodes8.emplace_back();
    DNode8 nd;
    std::memset(&nd, 0, sizeof(nd));
    Box3 bounds;
    for (int i = 0; i < n; i++) bounds.extend(slots[i].box);
    for (int k = 0; k < 3; k++) {
      nd.origin[k] = bounds.lower[k];
      float extent = bounds.upper[k] - bounds.lower[k];
      int e = 0;
      if (extent > 0) {
        std::frexp(extent / 255.0f, &e);  // extent/255 = m * 2^e, m in [0.5,1) => 2^e >= extent/255
        // guard against rounding in origin + 255*scale falling short of upper
        while (plane(nd.origin[k], 255, uint8_t(pmin(pmax(e + 127, 1), 254))) < bounds.upper[k]) e++;
      } else {
        e = -126;
      }
      nd.exp[k] = uint8_t(pmin(pmax(e + 127, 1), 254));
    }
    nd.n_children = uint8_t(n);
    for (int i = 0; i < n; i++) {
      for (int k = 0; k < 3; k++) {
        float scale = u2f(uint32_t(nd.exp[k]) << 23);
        int ql = int(std::floor((slots[i].box.lower[k] - nd.origin[k]) / scale));
        int qh = int(std::ceil((slots[i].box.upper[k] - nd.origin[k]) / scale));
        ql = pmin(pmax(ql, 0), 255);
        qh = pmin(pmax(qh, 0), 255);
        while (ql > 0 && plane(nd.origin[k], uint8_t(ql), nd.exp[k]) > slots[i].box.lower[k]) ql--;
        while (qh < 255 && plane(nd.origin[k], uint8_t(qh), nd.exp[k]) < slots[i].box.upper[k]) qh++;
        nd.qlo[k][i] = uint8_t(ql);
        nd.qhi[k][i] = uint8_t(qh);
      }
    }
    for (int i = n; i < 8; i++)
      for (int k = 0; k < 3; k++) { nd.qlo[k][i] = 255; nd.qhi[k][i] = 0; }  // inverted = never hit
    for (int i = 0; i < n; i++)
      nd.child[i] = slots[i].node >= 0 ? build(slots[i].node) : ~make_leaf(slots[i].leaf_node, nullptr);
    out.nodes8[idx] = nd;
    return idx;
  }
};

}  // namespace

DTree append_tree(const Tree2& t, LayoutOut& out, const pb2_mesh* mesh, int geom) {
  DTree d{-1, 0, 0, 0x7fffffff};
  if (t.root < 0) return d;
  const int node_base = int(out.nodes2.size());
  const int leaf_base = int(out.leaf_prims.size());
  out.leaf_prims.insert(out.leaf_prims.end(), t.leaf_prims.begin(), t.leaf_prims.end());

  // binary layout: internal nodes only get a DNode2; children that are leaves are folded into the
  // parent's link/count. Node indices are remapped densely over internal nodes.
  std::vector<int> remap(t.nodes.size(), -1);
  int n_internal = 0;
  for (size_t i = 0; i < t.nodes.size(); i++)
    if (!t.nodes[i].leaf_count) remap[i] = node_base + n_internal++;
  out.nodes2.resize(size_t(node_base) + n_internal);
  for (size_t i = 0; i < t.nodes.size(); i++) {
    const Node2& nd = t.nodes[i];
    if (nd.leaf_count) continue;
    DNode2& o = out.nodes2[remap[i]];
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
  if (!mesh) return d;  // the top-level tree is only ever walked in reference order (traverse.cuh)
  WideBuilder wb{t, out, mesh, geom, leaf_base, std::vector<int>(t.nodes.size(), -1)};
  if (root.leaf_count) {
    Box3 everything;  // a root leaf is tested without any box test (bvh.cpp:398-400)
    everything.lower = V3(-kFloatMax);
    everything.upper = V3(kFloatMax);
    d.wide_root = ~wb.make_leaf(t.root, &everything);
  } else {
    d.wide_root = wb.build(t.root);
  }
  return d;
}

}  // namespace pb2
