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
// (2) `*_wide`: ONE 8-wide compressed tree over the whole scene (device_scene.h DNode8) for
//     production, made exact by detection + fallback to (1).
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

constexpr int kStack = 64;  // the reference uses 32 and has undefined behaviour beyond it; pb2_upload_scene refuses
                            // scenes with a deeper tree (scene_upload.cu kRefOrderStack), so no push is ever dropped

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
// SHAPES: a kShapes* mask (shapes.cuh) that contains the scene's geometry kinds
template <int SHAPES = kShapesAll>
__device__ __forceinline__ bool closest_ref_order(const DScene& sc, Ray& ray, Hit& h) {
  Octant q = make_octant(ray);
  h.geom = -1;
  h.prim = -1;
  return walk_ref_order<false>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if ((SHAPES & kShapesMesh) && tp.kind == PB2_SHAPE_MESH) {
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
    bool hit = shape_intersect<SHAPES>(sc, tp.kind, tp.index, ray);
    if (hit) { h.geom = tp.geom; h.prim = 0; }
    return hit;
  });
}

// BVH::hit (bvh.cpp:497-511).
template <int SHAPES = kShapesAll>
__device__ __forceinline__ bool any_ref_order(const DScene& sc, const Ray& ray) {
  Octant q = make_octant(ray);
  return walk_ref_order<true>(sc, sc.top, ray, q, [&](int top_id) {
    DTopPrim tp = sc.top_prims[top_id];
    if ((SHAPES & kShapesMesh) && tp.kind == PB2_SHAPE_MESH) {
      const DMesh& m = sc.meshes[tp.index];
      return walk_ref_order<true>(sc, m.tree, ray, q, [&](int tri) {
        V3 v0, v1, v2;
        mesh_tri(sc, m, tri, v0, v1, v2);
        return tri_hit(ray, v0, v1, v2);
      });
    }
    return shape_hit<SHAPES>(sc, tp.kind, tp.index, ray);
  });
}


// ====================================================================================================
// (2) wide traversal
// ====================================================================================================
// The wide walk visits primitives in a different order than the reference and tests a superset of
// the primitives the reference tests (bvh_layout.cpp). Its result equals the reference's whenever
//   (i)   no two accepted candidates report exactly the same t as the winner (ties are broken by
//         visiting order in the reference: strict `t < tmax` for Rect/Triangle, geometry.cpp:296,557),
//   (ii)  the winner's own reference leaf box ("gate") passes the reference slab test with
//         tmax = t_winner -- then the reference, whose tmax is always >= t_winner and whose
//         ancestor boxes contain the leaf box, reaches and accepts the winner, and nothing it sees
//         later can replace it (slab tests are monotone in tmax and in the box),
//   (ii') or, when that gate fails: it passes with tmax = U = min(incoming tmax, t2, half_slack_of(t))
//         where t2 is the second-smallest t among the candidates seen. The primitive test accepts
//         every candidate with t < slack_of(best), so every primitive whose t lies within half
//         the slack of the winner has been seen (boxes are culled against slack_of(best) and a box's
//         computed entry never exceeds the t of a primitive inside it by more than the other
//         half); the reference's tmax when it reaches the winner's leaf is the incoming tmax or
//         the t of a primitive it accepted earlier, i.e. >= U, and its slab test is monotone in
//         tmax. This certifies the common case of FLAT leaf boxes (a Rect, an axis-aligned
//         triangle), whose slab entry and the shape's own t are two float expressions of the
//         same distance and differ by a few ulps: without it 2.4% (C5), 9% (C3) of the closest-hit
//         rays took the fallback,
//   (iii) no primitive's reported t depends on the incoming tmax. True for everything but AABB/OBB
//         shapes (bbox.cpp:149-161: a hit from inside reports the exit clipped by tmax, and a
//         scaled OBB can even RAISE tmax). Their acceptance is monotone in tmax, so a ray for which
//         every box shape is rejected at the ray's incoming tmax (shape test AND leaf gate) never
//         has one accepted by the reference either: such rays take the wide walk with the box
//         shapes left out, all others go to the reference-order walk (wide_persistent.cuh),
//   (iv)  every primitive the reference accepts is also tested here. For primitives that lie
//         inside their leaf box this follows from box containment; a Cone does not (it is
//         unbounded below its base, geometry.cpp:415-446; likewise a Plane beyond the +-100 box and a
//         Cylinder beyond the radius-r box around p0 the reference bounds them by) and can be hit BEFORE the ray enters its
//         leaf box, which the reference accepts if it reaches the leaf while tmax is still large.
//         Nodes with a Cone anywhere below them (kNoCullBit) are therefore tested against the
//         ray's incoming tmax, never the current best; a Cone hit outside its box then fails (ii).
// Boxes are culled against a slightly enlarged tmax (`slack_of`), so a primitive whose t ties or
// beats the current best is never hidden behind a box whose computed entry distance exceeds t by
// rounding. Rays for which (i) or (ii) fails, or whose stack overflows (a handful per million),
// are re-traced in reference order: the result is the reference's in all cases.
// Any-hit (shadow rays) is an OR over primitives whose gate passes with the ray's fixed tmax, so the
// visiting order is irrelevant and no fallback is needed; it is valid for AABB/OBB shapes too.
constexpr int kWideStack = 28;

__device__ __forceinline__ float slack_of(float t) { return t + (t * 1.220703125e-4f + 1e-6f); }
__device__ __forceinline__ float half_slack_of(float t) { return t + (t * 6.103515625e-5f + 5e-7f); }
__device__ __forceinline__ float next_up(float t) {
  return (t >= kFloatMax || t < 0) ? t : __uint_as_float(__float_as_uint(t) + 1u);
}
__device__ __forceinline__ bool leaf_gate(const DLeaf* leaves, int leaf, const Octant& q, float tmin, float tmax) {
  if (leaf < 0) return true;
  const float4* p = reinterpret_cast<const float4*>(leaves + leaf);
  float4 a = __ldg(p), b = __ldg(p + 1);
  float box[6] = {a.x, a.y, a.z, a.w, b.x, b.y};
  float t = tmax;
  return slab(box, q, tmin, t);
}

// Per-ray constants of the conservative child-box test. For axis k the exact plane q of a node is
// P_q = origin + q * 2^e; for a box plane P inside the quantised box the reference evaluates
// R = fl(fl(P * inv) - ood) (bbox.h:59-72), which is monotone in P. We evaluate
//   t = fma(2^15 + q, a, base -/+ m),  a = 2^e * inv (exact),  base = fma(-2^15, a, fma(origin, inv, -ood))
// (the bias 2^15 + q is what byte_biased builds with one PRMT). With u = 2^-24 and X = q a + origin inv - ood:
//   t  = X -/+ m + e1,  |e1| <= u (|b| + |base| + |base -/+ m| + |t|)             (four roundings of ours)
//   R_q = X + e2,       |e2| <= u (|P_q inv| + |R_q|)                               (two of the reference's)
// and |b| <= |origin inv| + |ood|, |base| <= |b| + 2^15 |a|, |P_q inv| <= |origin inv| + 255 |a|, so
//   |e1| + |e2| <= u (6 |origin inv| + 5 |ood| + 66301 |a| + 2 m).
// m = 2^-21 (8447 |a|) + err with err >= 2^-21 (|origin inv| + |ood|), i.e. m >= u (8 |origin inv| + 8 |ood| +
// 67576 |a|), exceeds that bound, so near planes come out no larger and far planes no smaller than the
// reference's value for any box inside the quantised one. err depends on the node only through
// |origin inv| <= absmax|origin| * |inv| and is therefore computed ONCE per ray (WideRay::err); the margin is
// ~1e-6 of the ray's coordinate range, two orders below one quantisation step of a leaf-level node.
struct WideRay {
  V3 inv, ood;
  V3 err;            // 2^-21 (absmax |node origin| * |inv| + |ood|) per axis: the part of delta that does not
                     // depend on the node (|origin * inv| <= absmax * |inv|), hoisted out of the node test
  uint32_t octinv4;  // (7 ^ octant) replicated into 4 bytes
  uint32_t neg;      // bit k: d[k] < 0
};
// Byte k of w as the float 2^15 + byte (bits 0x4700bb00), built by ONE PRMT. The bias is folded into
// the per-node constants (b - 2^15 a), so a plane distance costs PRMT + FMA instead of
// PRMT + FADD + FMA: 48 instructions fewer per wide-node test.
__device__ __forceinline__ float byte_biased(uint32_t w, int k) {
  // constant first: SASS PRMT takes ONE immediate, and it must be the selector (4 different ones per
  // word) rather than the constant, or every PRMT is preceded by a MOV of its selector
  return __uint_as_float(__byte_perm(0x47000000u, w, 0x3140u + (uint32_t(k) << 4)));
}
constexpr float kByteBias = 32768.0f;

struct PrimRec {
  V3 v0, v1, v2;
  int prim, geom, leaf;
};
__device__ __forceinline__ PrimRec load_prim(const DPrim* prims, uint32_t idx) {
  const float4* p = reinterpret_cast<const float4*>(prims + idx);
  float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  PrimRec r;
  r.v0 = V3(a.x, a.y, a.z); r.v1 = V3(a.w, b.x, b.y); r.v2 = V3(b.z, b.w, c.x);
  r.prim = __float_as_int(c.y); r.geom = __float_as_int(c.z); r.leaf = __float_as_int(c.w);
  return r;
}

// Tests the 8 children of node `idx`; returns the hit mask (bits 24..31: internal children in
// traversal order, bits 0..23: primitives) and the node's child_base / prim_base / imask.
// `lim` culls by the current best t; nodes flagged kNoCullBit (an unbounded shape somewhere below,
// see the header comment) are tested against `lim_nocull` = the ray's incoming tmax instead.
constexpr uint32_t kNoCullBit = 0x80000000u;  // in DNode8::prim_base
// PB2_SMEM_TOP = N (north_star: "node data staged through shared memory"): the first N nodes of the breadth-first
// array -- the top levels -- are copied into shared memory by every block of the persistent kernels and read
// from there. Measured on B200 (profiles/r2/s7_smem_variants.txt, two runs each, C5 / C4 Msamples/s):
//   N = 0: 531 / 274    N = 9 (root + its children): 544 / 277    N = 73: 540 / 274    N = 201: 540 / 273
//   N = 585 (four levels, 47 KB per block): 359 / 189 -- the copies (one per block, 8 per SM) cost more L1
//   capacity and occupancy than the few lines the top levels take when the SM's blocks share them through L1.
// Kept: N = 9 -- every ray starts at the root and nearly every ray visits several of its children, so these
// are the node fetches with the most reuse and their latency is on every ray's critical path.
#ifndef PB2_SMEM_TOP
#define PB2_SMEM_TOP 9
#endif
// The test on a node already in registers (n0..n4 = its five 128-bit words).
// NOCULL = false: the tree holds no unbounded shape (no node carries kNoCullBit), the switch of limits is compiled out.
template <bool NOCULL = true>
__device__ __forceinline__ uint32_t test_node8_regs(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3,
                                                    const uint4 n4, const WideRay& w, float tmin, float lim,
                                                    float lim_nocull, uint32_t& child_base, uint32_t& prim_base,
                                                    uint32_t& imask) {
  const float ox = __uint_as_float(n0.x), oy = __uint_as_float(n0.y), oz = __uint_as_float(n0.z);
  const float ax = __uint_as_float((n0.w & 0xffu) << 23) * w.inv.x;
  const float ay = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * w.inv.y;
  const float az = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * w.inv.z;
  imask = n0.w >> 24;
  child_base = n1.x;
  prim_base = NOCULL ? (n1.y & ~kNoCullBit) : n1.y;
  if (NOCULL && (n1.y & kNoCullBit)) lim = lim_nocull;
  // b = origin * inv - ood with ONE rounding (the reference's fl(fl(P inv) - ood) has two, and delta below
  // bounds the distance of both expressions from the exact value), then the byte bias: base = b - 2^15 a.
  // delta = w.err + 2^-21 * 8447 |a| (see the derivation above; w.err >= 2^-21 (|origin inv| + |ood|)), and
  // the whole margin is ONE fma per axis because its node-independent part is per ray.
  const float mx = __fmaf_rn(fabsf(ax), 8447.0f * 4.76837158203125e-7f, w.err.x);
  const float my = __fmaf_rn(fabsf(ay), 8447.0f * 4.76837158203125e-7f, w.err.y);
  const float mz = __fmaf_rn(fabsf(az), 8447.0f * 4.76837158203125e-7f, w.err.z);
  const float basex = __fmaf_rn(-kByteBias, ax, __fmaf_rn(ox, w.inv.x, -w.ood.x));
  const float basey = __fmaf_rn(-kByteBias, ay, __fmaf_rn(oy, w.inv.y, -w.ood.y));
  const float basez = __fmaf_rn(-kByteBias, az, __fmaf_rn(oz, w.inv.z, -w.ood.z));
  const float bnx = basex - mx, bfx = basex + mx, bny = basey - my, bfy = basey + my, bnz = basez - mz, bfz = basez + mz;
  // qlo: x = n2.xy, y = n2.zw, z = n3.xy; qhi: x = n3.zw, y = n4.xy, z = n4.zw. near = lo unless d < 0
  const bool nx = w.neg & 1, ny = w.neg & 2, nz = w.neg & 4;
  uint32_t hitmask = 0;
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const uint32_t meta4 = j ? n1.w : n1.z;
    const uint32_t lox = j ? n2.y : n2.x, loy = j ? n2.w : n2.z, loz = j ? n3.y : n3.x;
    const uint32_t hix = j ? n3.w : n3.z, hiy = j ? n4.y : n4.x, hiz = j ? n4.w : n4.z;
    const uint32_t qnx = nx ? hix : lox, qfx = nx ? lox : hix;
    const uint32_t qny = ny ? hiy : loy, qfy = ny ? loy : hiy;
    const uint32_t qnz = nz ? hiz : loz, qfz = nz ? loz : hiz;
    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
    // byte k: count bits (unary, 3) << 5 | bit index (5): internal children get slot ^ octant in the low 3 bits
    const uint32_t m4 = meta4 ^ (w.octinv4 & inner_mask4);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float tnx = __fmaf_rn(byte_biased(qnx, k), ax, bnx), tfx = __fmaf_rn(byte_biased(qfx, k), ax, bfx);
      const float tny = __fmaf_rn(byte_biased(qny, k), ay, bny), tfy = __fmaf_rn(byte_biased(qfy, k), ay, bfy);
      const float tnz = __fmaf_rn(byte_biased(qnz, k), az, bnz), tfz = __fmaf_rn(byte_biased(qfz, k), az, bfz);
      const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
      const float tf = fminf(fminf(tfx, tfy), fminf(tfz, lim));
      const uint32_t b = (m4 >> (8 * k)) & 0xffu;
      if (tn <= tf) hitmask |= (b >> 5) << (b & 31u);
    }
  }
  return hitmask;
}
template <bool NOCULL = true>
__device__ __forceinline__ uint32_t test_node8(const DNode8* nodes, uint32_t idx, const WideRay& w, float tmin,
                                               float lim, float lim_nocull, uint32_t& child_base,
                                               uint32_t& prim_base, uint32_t& imask, const uint4* s_top = nullptr) {
  const uint4* p = reinterpret_cast<const uint4*>(nodes + idx);
#if PB2_SMEM_TOP > 0
  uint4 n0, n1, n2, n3, n4;
  if (idx < uint32_t(PB2_SMEM_TOP)) {
    const uint4* q = s_top + idx * 5u;
    n0 = q[0]; n1 = q[1]; n2 = q[2]; n3 = q[3]; n4 = q[4];
  } else {
    n0 = __ldg(p); n1 = __ldg(p + 1); n2 = __ldg(p + 2); n3 = __ldg(p + 3); n4 = __ldg(p + 4);
  }
#else
  const uint4 n0 = __ldg(p), n1 = __ldg(p + 1), n2 = __ldg(p + 2), n3 = __ldg(p + 3), n4 = __ldg(p + 4);
#endif
  return test_node8_regs<NOCULL>(n0, n1, n2, n3, n4, w, tmin, lim, lim_nocull, child_base, prim_base, imask);
}

__device__ __forceinline__ WideRay make_wide_ray(const DScene& sc, const Ray& ray, const Octant& q) {
  WideRay w;
  w.inv = q.inv;
  w.ood = q.ood;
  // rounded UP a little (1 + 2^-10): the three roundings of this expression must not shrink the bound
  w.err = V3((sc.node_origin_absmax[0] * fabsf(q.inv.x) + fabsf(q.ood.x)) * 4.7730281949043274e-7f,
             (sc.node_origin_absmax[1] * fabsf(q.inv.y) + fabsf(q.ood.y)) * 4.7730281949043274e-7f,
             (sc.node_origin_absmax[2] * fabsf(q.inv.z) + fabsf(q.ood.z)) * 4.7730281949043274e-7f);
  w.neg = (ray.d.x < 0 ? 1u : 0u) | (ray.d.y < 0 ? 2u : 0u) | (ray.d.z < 0 ? 4u : 0u);
  w.octinv4 = (7u ^ w.neg) * 0x01010101u;
  return w;
}

}  // namespace pb2
