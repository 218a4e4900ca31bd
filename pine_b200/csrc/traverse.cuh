// traverse.cuh -- BVH traversal on the device.
//
// (1) `*_ref_order`: walks the reference-identical binary tree in the reference's order
//     (BVHImpl::Intersect / BVHImpl::hit, src/pine/impl/accel/bvh.cpp:321-451, two-level wrapper
//     BVH::intersect / BVH::hit :497-548), reproducing its quirks:
//       * the slab test is AABB::hit(RayOctant) (bbox.h:59-72): p*inv - o*inv, compare-select
//         max/min, `tmin <= tmax`;
//       * t0/t1 are read from ray.tmax once per node BEFORE either child is processed, and what
//         the ordering `t0 > t1` compares afterwards is each child's slab EXIT distance;
//       * a leaf child is intersected immediately, inside the node visit, before the sibling's box
//         test; internal children go near-first with the other pushed on a stack.
//     This is the verification path and the fallback of (2); results are bit-identical.
// (2) `*_wide`: 8-wide quantised tree (device_scene.h DNode8) for production.
#pragma once
#include "shapes.cuh"

namespace pb2 {

struct Octant {
  int ox, oy, oz;  // 0 or 3: RayOctant::octantx3 (bbox.h:19-21), psl::signbit(x) = x < 0
  V3 inv, ood;     // safe_rcp(d), o * inv
};
__device__ __forceinline__ Octant make_octant(const Ray& r) {
  Octant q;
  q.ox = r.d.x < 0 ? 3 : 0;
  q.oy = r.d.y < 0 ? 3 : 0;
  q.oz = r.d.z < 0 ? 3 : 0;
  q.inv = V3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
  q.ood = r.o * q.inv;
  return q;
}
// AABB::hit(const RayOctant&, float tmin, float& tmax): on return tmax = slab exit distance.
__device__ __forceinline__ bool slab(const float* p, const Octant& q, float tmin, float& tmax) {
  float tmin0 = p[0 + q.ox] * q.inv.x - q.ood.x;
  float tmin1 = p[1 + q.oy] * q.inv.y - q.ood.y;
  float tmin2 = p[2 + q.oz] * q.inv.z - q.ood.z;
  float tmax0 = p[3 - q.ox] * q.inv.x - q.ood.x;
  float tmax1 = p[4 - q.oy] * q.inv.y - q.ood.y;
  float tmax2 = p[5 - q.oz] * q.inv.z - q.ood.z;
  tmin = pmax(tmin0, pmax(tmin1, pmax(tmin2, tmin)));
  tmax = pmin(tmax0, pmin(tmax1, pmin(tmax2, tmax)));
  return tmin <= tmax;
}

__device__ __forceinline__ void load_node(const DNode2* nodes, int idx, float (&b)[2][6],
                                          int (&link)[2], int (&count)[2]) {
  const float4* p = reinterpret_cast<const float4*>(nodes + idx);
  float4 a0 = __ldg(p), a1 = __ldg(p + 1), a2 = __ldg(p + 2), a3 = __ldg(p + 3);
  b[0][0] = a0.x; b[0][1] = a0.y; b[0][2] = a0.z; b[0][3] = a0.w; b[0][4] = a1.x; b[0][5] = a1.y;
  b[1][0] = a1.z; b[1][1] = a1.w; b[1][2] = a2.x; b[1][3] = a2.y; b[1][4] = a2.z; b[1][5] = a2.w;
  link[0] = __float_as_int(a3.x); link[1] = __float_as_int(a3.y);
  count[0] = __float_as_int(a3.z); count[1] = __float_as_int(a3.w);
}

constexpr int kStack = 64;  // the reference uses 32 and has undefined behaviour beyond it

// Generic reference-order walk. `leaf(prim_id)` returns true on an accepted hit; for closest-hit it
// also shrinks ray.tmax; when ANY is set the walk returns at the first true.
template <bool ANY, class Leaf>
__device__ __forceinline__ bool walk_ref_order(const DScene& sc, const DTree& tree, const Ray& ray,
                                               const Octant& q, Leaf&& leaf) {
  if (tree.root < 0 && tree.root_count == 0) return false;
  bool hit = false;
  if (tree.root_count > 0) {
    for (int i = 0; i < tree.root_count; i++)
      if (leaf(__ldg(sc.leaf_prims + tree.root_link + i))) {
        if (ANY) return true;
        hit = true;
      }
    return hit;
  }
  int stack[kStack];
  int ptr = 0;
  int next = tree.root;
  while (true) {
    float b[2][6];
    int link[2], count[2];
    load_node(sc.nodes2, next, b, link, count);
    int child[2] = {-1, -1};
    float t[2] = {ray.tmax, ray.tmax};
#pragma unroll
    for (int c = 0; c < 2; c++) {
      if (slab(b[c], q, ray.tmin, t[c])) {
        if (count[c] == 0) {
          child[c] = link[c];
        } else {
          for (int i = 0; i < count[c]; i++)
            if (leaf(__ldg(sc.leaf_prims + link[c] + i))) {
              if (ANY) return true;
              hit = true;
            }
        }
      }
    }
    if (child[0] != -1) {
      if (child[1] != -1) {
        if (t[0] > t[1]) {
          if (ptr < kStack) stack[ptr++] = child[0];
          next = child[1];
        } else {
          if (ptr < kStack) stack[ptr++] = child[1];
          next = child[0];
        }
      } else {
        next = child[0];
      }
    } else if (child[1] != -1) {
      next = child[1];
    } else {
      if (ptr == 0) break;
      next = stack[--ptr];
    }
  }
  return hit;
}

// BVH::intersect (bvh.cpp:513-548). ray.tmax is shrunk to t; returns geom/prim of the winner.
__device__ __forceinline__ bool closest_ref_order(const DScene& sc, Ray& ray, Hit& h) {
  Octant q = make_octant(ray);
  h.geom = -1;
  h.prim = -1;
  return walk_ref_order<false>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if (tp.kind == PB2_SHAPE_MESH) {
      const DMesh& m = sc.meshes[tp.index];
      int prim = -1;
      bool hit = walk_ref_order<false>(sc, m.tree, ray, q, [&](int tri) {
        V3 v0, v1, v2;
        mesh_tri(sc, m, tri, v0, v1, v2);
        bool ok = tri_intersect(ray, v0, v1, v2);
        if (ok) prim = tri;
        return ok;
      });
      if (hit) { h.geom = tp.geom; h.prim = prim; }
      return hit;
    }
    bool hit = shape_intersect(sc, tp.kind, tp.index, ray);
    if (hit) { h.geom = tp.geom; h.prim = 0; }
    return hit;
  });
}

// BVH::hit (bvh.cpp:497-511).
__device__ __forceinline__ bool any_ref_order(const DScene& sc, const Ray& ray) {
  Octant q = make_octant(ray);
  return walk_ref_order<true>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if (tp.kind == PB2_SHAPE_MESH) {
      const DMesh& m = sc.meshes[tp.index];
      return walk_ref_order<true>(sc, m.tree, ray, q, [&](int tri) {
        V3 v0, v1, v2;
        mesh_tri(sc, m, tri, v0, v1, v2);
        return tri_hit(ray, v0, v1, v2);
      });
    }
    return shape_hit(sc, tp.kind, tp.index, ray);
  });
}


// ====================================================================================================
// (2) wide traversal
// ====================================================================================================
// Used for the triangles of one mesh; results must equal the reference-order walk of that mesh bit
// for bit. A wide walk visits triangles in a different order, which only matters when two candidates
// are (nearly) tied. Let (p1, t1) be the strict minimum over all triangles the reference could
// accept. The reference returns (p1, t1) if
//   (i)  no other triangle reports exactly t1, and
//   (ii) p1's own leaf box passes the reference slab test with tmax = t1
// (every tmax the reference ever holds is >= t1, slab tests are monotone in tmax and in the box, so
// p1's leaf and all its ancestors pass whenever they are reached; p1 is then accepted -- the
// triangle test is strict, geometry.cpp:557 -- and nothing later can replace it). The wide walk
// culls with a slightly enlarged tmax (`slack`) so that a triangle whose t ties or beats the current
// best is never hidden behind a box whose entry distance differs from t by rounding, detects exact
// ties and checks (ii) for its winner. Ambiguous rays (a handful per million) re-trace the mesh in
// reference order, so the result is the reference's in all cases.
constexpr int kWideStack = 64;
constexpr int kNoTree = 0x7fffffff;

__device__ __forceinline__ float slack_of(float t) { return t + (t * 1.220703125e-4f + 1e-6f); }
__device__ __forceinline__ float next_up(float t) {
  return (t >= kFloatMax || t < 0) ? t : __uint_as_float(__float_as_uint(t) + 1u);
}

struct Node8Regs {
  float4 a;        // origin.xyz, {exp x,y,z, n_children}
  uint4 q0, q1, q2;  // qlo[x][8] qlo[y][8] | qlo[z][8] qhi[x][8] | qhi[y][8] qhi[z][8]
  int4 c0, c1;
};
__device__ __forceinline__ Node8Regs load_node8(const DNode8* nodes, int idx) {
  const uint4* p = reinterpret_cast<const uint4*>(nodes + idx);
  Node8Regs r;
  uint4 a = __ldg(p);
  r.a = make_float4(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(a.w));
  r.q0 = __ldg(p + 1); r.q1 = __ldg(p + 2); r.q2 = __ldg(p + 3);
  uint4 c0 = __ldg(p + 4), c1 = __ldg(p + 5);
  r.c0 = make_int4(c0.x, c0.y, c0.z, c0.w);
  r.c1 = make_int4(c1.x, c1.y, c1.z, c1.w);
  return r;
}
__device__ __forceinline__ uint32_t byte_of(uint32_t lo, uint32_t hi, int i) {
  return ((i < 4 ? lo : hi) >> ((i & 3) * 8)) & 0xffu;
}

// Slab test of child i of a wide node with the reference's expression on the dequantised planes.
__device__ __forceinline__ bool child_slab(const Node8Regs& nd, int i, const Octant& q, float tmin,
                                           float tmax, float& tnear) {
  uint32_t ew = __float_as_uint(nd.a.w);
  float sx = __uint_as_float((ew & 0xffu) << 23), sy = __uint_as_float(((ew >> 8) & 0xffu) << 23),
        sz = __uint_as_float(((ew >> 16) & 0xffu) << 23);
  float p[6];
  p[0] = nd.a.x + float(byte_of(nd.q0.x, nd.q0.y, i)) * sx;
  p[1] = nd.a.y + float(byte_of(nd.q0.z, nd.q0.w, i)) * sy;
  p[2] = nd.a.z + float(byte_of(nd.q1.x, nd.q1.y, i)) * sz;
  p[3] = nd.a.x + float(byte_of(nd.q1.z, nd.q1.w, i)) * sx;
  p[4] = nd.a.y + float(byte_of(nd.q2.x, nd.q2.y, i)) * sy;
  p[5] = nd.a.z + float(byte_of(nd.q2.z, nd.q2.w, i)) * sz;
  float tmin0 = p[0 + q.ox] * q.inv.x - q.ood.x;
  float tmin1 = p[1 + q.oy] * q.inv.y - q.ood.y;
  float tmin2 = p[2 + q.oz] * q.inv.z - q.ood.z;
  float tmax0 = p[3 - q.ox] * q.inv.x - q.ood.x;
  float tmax1 = p[4 - q.oy] * q.inv.y - q.ood.y;
  float tmax2 = p[5 - q.oz] * q.inv.z - q.ood.z;
  tnear = pmax(tmin0, pmax(tmin1, pmax(tmin2, tmin)));
  float tfar = pmin(tmax0, pmin(tmax1, pmin(tmax2, tmax)));
  return tnear <= tfar;
}
__device__ __forceinline__ bool leaf_gate(const DLeaf& L, const Octant& q, float tmin, float tmax) {
  float t = tmax;
  return slab(L.box, q, tmin, t);
}
__device__ __forceinline__ DLeaf load_leaf(const DLeaf* leaves, int idx) {
  const float4* p = reinterpret_cast<const float4*>(leaves + idx);
  float4 a = __ldg(p), b = __ldg(p + 1);
  DLeaf L;
  L.box[0] = a.x; L.box[1] = a.y; L.box[2] = a.z; L.box[3] = a.w; L.box[4] = b.x; L.box[5] = b.y;
  L.link = __float_as_int(b.z); L.count = __float_as_int(b.w);
  return L;
}
__device__ __forceinline__ void load_tri(const DTri* tris, int idx, V3& v0, V3& v1, V3& v2, int& prim,
                                         int& geom) {
  const float4* p = reinterpret_cast<const float4*>(tris + idx);
  float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  v0 = V3(a.x, a.y, a.z); v1 = V3(a.w, b.x, b.y); v2 = V3(b.z, b.w, c.x);
  prim = __float_as_int(c.y); geom = __float_as_int(c.z);
}

// Closest hit inside one mesh (the reference's lbvh[k].Intersect, bvh.cpp:523-527) over the wide
// tree. `ray.tmax` on entry is whatever the top-level walk holds; a triangle reporting exactly that
// value is rejected by the reference's strict `t >= tmax` test (geometry.cpp:557), so it is not a tie.
__device__ inline bool mesh_closest_wide(const DScene& sc, const DMesh& mesh, Ray& ray, const Octant& q,
                                         int& prim_out) {
  const int root = mesh.tree.wide_root;
  if (root == kNoTree) return false;
  const float tmax_in = ray.tmax;
  int refs[kWideStack];
  float dists[kWideStack];
  int ptr = 0;
  refs[ptr] = root; dists[ptr] = ray.tmin; ptr++;
  bool have = false, ambiguous = false;
  float tie_t = -1.0f;
  int win_leaf = -1;
  while (ptr > 0) {
    --ptr;
    int ref = refs[ptr];
    const float lim = slack_of(ray.tmax);
    if (dists[ptr] > lim) continue;
    if (ref >= 0) {
      Node8Regs nd = load_node8(sc.nodes8, ref);
      int n = int((__float_as_uint(nd.a.w) >> 24) & 0xffu);
      int cref[8];
      float ct[8];
      int m = 0;
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (i < n) {
          float tn;
          if (child_slab(nd, i, q, ray.tmin, lim, tn)) {
            int c = i < 4 ? (&nd.c0.x)[i] : (&nd.c1.x)[i - 4];
            int j = m++;
            while (j > 0 && ct[j - 1] < tn) { ct[j] = ct[j - 1]; cref[j] = cref[j - 1]; j--; }
            ct[j] = tn; cref[j] = c;  // descending: farthest first, nearest pushed last
          }
        }
      }
      if (ptr + m > kWideStack) { ambiguous = true; break; }
      for (int j = 0; j < m; j++) { refs[ptr] = cref[j]; dists[ptr] = ct[j]; ptr++; }
    } else {
      int leaf_id = ~ref;
      DLeaf L = load_leaf(sc.leaves, leaf_id);
      if (!leaf_gate(L, q, ray.tmin, lim)) continue;
      int count = L.count & 0x7fffffff;
      for (int i = 0; i < count; i++) {
        V3 v0, v1, v2;
        int prim, geom;
        load_tri(sc.leaf_tris, L.link + i, v0, v1, v2, prim, geom);
        Ray r2 = ray;
        r2.tmax = next_up(ray.tmax);  // lets t == current best through so that ties are seen
        if (tri_intersect(r2, v0, v1, v2)) {
          if (r2.tmax < ray.tmax) {
            ray.tmax = r2.tmax; prim_out = prim; win_leaf = leaf_id; have = true;
          } else if (have) {
            tie_t = r2.tmax;
          }
        }
      }
    }
  }
  if (have && !ambiguous) {
    if (tie_t == ray.tmax) ambiguous = true;
    else if (!leaf_gate(load_leaf(sc.leaves, win_leaf), q, ray.tmin, ray.tmax)) ambiguous = true;
  }
  if (ambiguous) {  // re-trace this mesh in reference order
    ray.tmax = tmax_in;
    return walk_ref_order<false>(sc, mesh.tree, ray, q, [&](int tri) {
      V3 v0, v1, v2;
      mesh_tri(sc, mesh, tri, v0, v1, v2);
      bool ok = tri_intersect(ray, v0, v1, v2);
      if (ok) prim_out = tri;
      return ok;
    });
  }
  return have;
}

// Any-hit inside one mesh (lbvh[k].hit, bvh.cpp:503-505): an OR over triangles whose leaf box passes
// with the ray's fixed tmax, so visiting order is irrelevant and no fallback is needed.
__device__ inline bool mesh_any_wide(const DScene& sc, const DMesh& mesh, const Ray& ray, const Octant& q) {
  const int root = mesh.tree.wide_root;
  if (root == kNoTree) return false;
  int refs[kWideStack];
  int ptr = 0;
  refs[ptr++] = root;
  while (ptr > 0) {
    int ref = refs[--ptr];
    if (ref >= 0) {
      Node8Regs nd = load_node8(sc.nodes8, ref);
      int n = int((__float_as_uint(nd.a.w) >> 24) & 0xffu);
      if (ptr + n > kWideStack) {  // cannot happen for sane trees; stay exact if it does
        return walk_ref_order<true>(sc, mesh.tree, ray, q, [&](int tri) {
          V3 v0, v1, v2;
          mesh_tri(sc, mesh, tri, v0, v1, v2);
          return tri_hit(ray, v0, v1, v2);
        });
      }
#pragma unroll
      for (int i = 0; i < 8; i++) {
        float tn;
        if (i < n && child_slab(nd, i, q, ray.tmin, ray.tmax, tn))
          refs[ptr++] = i < 4 ? (&nd.c0.x)[i] : (&nd.c1.x)[i - 4];
      }
    } else {
      DLeaf L = load_leaf(sc.leaves, ~ref);
      if (!leaf_gate(L, q, ray.tmin, ray.tmax)) continue;
      int count = L.count & 0x7fffffff;
      for (int i = 0; i < count; i++) {
        V3 v0, v1, v2;
        int prim, geom;
        load_tri(sc.leaf_tris, L.link + i, v0, v1, v2, prim, geom);
        if (tri_hit(ray, v0, v1, v2)) return true;
      }
    }
  }
  return false;
}

// Production traversal: the top-level tree (one primitive per geometry) is walked in the reference's
// order -- analytic shapes carry order-dependent quirks (a scaled Box compares a local-space entry
// distance with the world-space tmax, bbox.cpp:149-161; an origin inside a Box reports
// min(exit, tmax)), so their visiting order must be the reference's -- and every mesh below it uses
// the wide tree.
__device__ __forceinline__ bool closest_wide(const DScene& sc, Ray& ray, Hit& h) {
  Octant q = make_octant(ray);
  h.geom = -1;
  h.prim = -1;
  return walk_ref_order<false>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if (tp.kind == PB2_SHAPE_MESH) {
      int prim = -1;
      bool hit = mesh_closest_wide(sc, sc.meshes[tp.index], ray, q, prim);
      if (hit) { h.geom = tp.geom; h.prim = prim; }
      return hit;
    }
    bool hit = shape_intersect(sc, tp.kind, tp.index, ray);
    if (hit) { h.geom = tp.geom; h.prim = 0; }
    return hit;
  });
}
__device__ __forceinline__ bool any_wide(const DScene& sc, const Ray& ray) {
  Octant q = make_octant(ray);
  return walk_ref_order<true>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if (tp.kind == PB2_SHAPE_MESH) return mesh_any_wide(sc, sc.meshes[tp.index], ray, q);
    return shape_hit(sc, tp.kind, tp.index, ray);
  });
}

}  // namespace pb2
