// shapes.cuh -- per-shape closest-hit (`intersect`), any-hit (`hit`) and surface reconstruction,
// device side. Each function performs the reference's float operations in the reference's order
// (compile with -fmad=false), so `t` is bit-identical to the reference's.
//
// Reference: Rect geometry.cpp:280-312; AABB bbox.cpp:68-124; OBB bbox.cpp:143-175;
// Sphere geometry.cpp:73-101; Disk :129-158; Triangle :532-565; Cone (+quadratic) :20-29,415-460;
// Mesh :616-646.
#pragma once
#include "device_scene.h"

namespace pb2 {

__device__ __forceinline__ V3 ldv(const float* p) { return V3(p[0], p[1], p[2]); }
__device__ __forceinline__ V3 ray_at(const Ray& r, float t) { return r.o + t * r.d; }

// ---- Rect -----------------------------------------------------------------------------------------
__device__ __forceinline__ bool rect_test(const pb2_rect& s, const Ray& ray, float& t_out) {
  V3 n = ldv(s.n), pos = ldv(s.position);
  float denom = dot(ray.d, n);
  if (denom == 0.0f) return false;
  float t = dot(pos - ray.o, n) / denom;
  if (t <= ray.tmin || t >= ray.tmax) return false;
  V3 p = ray_at(ray, t) - pos;
  float u = dot(p, ldv(s.rx));
  if (u < -0.5f || u > 0.5f) return false;
  float v = dot(p, ldv(s.ry));
  if (v < -0.5f || v > 0.5f) return false;
  t_out = t;
  return true;
}

// ---- AABB as a shape ------------------------------------------------------------------------------
__device__ __forceinline__ bool aabb_slabs(V3 lo, V3 hi, V3 o, V3 d, float& tmin, float& tmax) {
  for (int i = 0; i < 3; i++) {
    if (pabs(d[i]) < 1e-6f) {
      if (o[i] < lo[i] || o[i] > hi[i]) return false;
      continue;
    }
    float inv_d = 1.0f / d[i];
    float t_near = (lo[i] - o[i]) * inv_d;
    float t_far = (hi[i] - o[i]) * inv_d;
    if (inv_d < 0.0f) { float s = t_far; t_far = t_near; t_near = s; }
    tmin = pmax(t_near, tmin);
    tmax = pmin(t_far, tmax);
    if (tmin > tmax) return false;
  }
  return true;
}
// `inside` (optional) reports the origin-inside-the-box case, where the reported t is
// min(exit, ray.tmax) and therefore depends on what was hit before (order dependent).
__device__ __forceinline__ bool aabb_intersect(const pb2_aabb& s, Ray& ray, bool* inside = nullptr) {
  float tmin = ray.tmin, tmax = ray.tmax;
  if (!aabb_slabs(ldv(s.lower), ldv(s.upper), ray.o, ray.d, tmin, tmax)) return false;
  if (inside) *inside = !(tmin > ray.tmin);
  ray.tmax = tmin > ray.tmin ? tmin : tmax;
  return true;
}
__device__ __forceinline__ bool aabb_hit(const pb2_aabb& s, V3 o, V3 d, float tmin, float tmax) {
  if (tmin > tmax) return false;
  return aabb_slabs(ldv(s.lower), ldv(s.upper), o, d, tmin, tmax);
}

// ---- OBB ------------------------------------------------------------------------------------------
struct ObbXform {
  M4 m, m_inv;
};
__device__ __forceinline__ M4 ld_m4_dev(const float* p) {
  return M4(V4(p[0], p[1], p[2], p[3]), V4(p[4], p[5], p[6], p[7]), V4(p[8], p[9], p[10], p[11]),
            V4(p[12], p[13], p[14], p[15]));
}
__device__ __forceinline__ bool obb_intersect(const pb2_obb& s, Ray& ray, bool* inside = nullptr) {
  M4 m_inv = ld_m4_dev(s.m_inv);
  V3 org = ray.o;
  V3 o = mul_point(m_inv, ray.o);
  V3 d = normalize(upper3(m_inv) * ray.d);
  float tmin = ray.tmin, tmax = ray.tmax;
  if (!aabb_slabs(ldv(s.base.lower), ldv(s.base.upper), o, d, tmin, tmax)) return false;
  M4 m = ld_m4_dev(s.m);
  V3 ps = o + tmin * d;
  V3 pe = o + tmax * d;
  tmin = distance(mul_point(m, ps), org);
  tmax = distance(mul_point(m, pe), org);
  if (inside) *inside = !(tmin > ray.tmin);
  ray.tmax = tmin > ray.tmin ? tmin : tmax;
  return true;
}
__device__ __forceinline__ bool obb_hit(const pb2_obb& s, const Ray& ray) {
  M4 m_inv = ld_m4_dev(s.m_inv);
  V3 o = mul_point(m_inv, ray.o);
  V3 d = normalize(upper3(m_inv) * ray.d);
  return aabb_hit(s.base, o, d, ray.tmin, ray.tmax);
}

// ---- Sphere ---------------------------------------------------------------------------------------
__device__ __forceinline__ float sphere_t(V3 ro, V3 rd, float tmin, V3 c, float r) {
  V3 ro_p = ro - c;
  float b = dot(ro_p, rd);
  float cc = dot(ro_p, ro_p) - r * r;
  float d = b * b - cc;
  if (d <= 0.0f) return -1.0f;
  d = sqrtf(d);
  float t = -b - d;
  if (t < tmin) t = -b + d;
  return t;
}
__device__ __forceinline__ bool sphere_intersect(const pb2_sphere& s, Ray& ray) {
  float t = sphere_t(ray.o, ray.d, ray.tmin, ldv(s.c), s.r);
  if (t < ray.tmin || t > ray.tmax) return false;
  ray.tmax = t;
  return true;
}
__device__ __forceinline__ bool sphere_hit(const pb2_sphere& s, const Ray& ray) {
  float t = sphere_t(ray.o, ray.d, ray.tmin, ldv(s.c), s.r);
  return t > ray.tmin && t < ray.tmax;
}

// ---- Disk -----------------------------------------------------------------------------------------
__device__ __forceinline__ bool disk_intersect(const pb2_disk& s, Ray& ray) {
  V3 n = ldv(s.n), pos = ldv(s.position);
  float denom = dot(ray.d, n);
  if (denom == 0.0f) return false;
  float t = (dot(pos, n) - dot(ray.o, n)) / denom;
  if (t < ray.tmin || t > ray.tmax) return false;
  V3 p = ray_at(ray, t) - pos;
  if (length_squared(p) > sqr(s.r)) return false;
  ray.tmax = t;
  return true;
}
__device__ __forceinline__ bool disk_hit(const pb2_disk& s, const Ray& ray) {
  V3 n = ldv(s.n), pos = ldv(s.position);
  float denom = dot(ray.d, n);
  if (denom == 0.0f) return false;
  float t = (dot(pos, n) - dot(ray.o, n)) / denom;
  if (t < ray.tmin) return false;
  if (t >= ray.tmax) return false;
  V3 p = ray_at(ray, t) - pos;
  if (length_squared(p) > sqr(s.r)) return false;
  return true;
}

// ---- Triangle (Moller-Trumbore with three divisions) -----------------------------------------------
// The reference divides three times (t, u, v = ... / D) and rejects on t, u, v, u + v in that order; the outcome
// is a pure AND of the four conditions, so they may be evaluated in any order. Most candidates fail the u / v
// range, which can be decided WITHOUT dividing whenever the numerator is clearly outside [0, D]:
//   * a / D rounds to a negative float iff a and D differ in sign and the quotient does not underflow to -0
//     (|a| > |D| * 1e-30 guarantees that);
//   * a / D rounds to more than 1 if |a| > |D| (1 + 2^-22) (the quotient then exceeds 1 + ulp);
//   * for u, v in [0, 1], fl(u + v) differs from (a + b) / D by less than 2e-7, so (a + b) > |D| (1 + 2e-6)
//     means u + v > 1.
// These are certain rejections; everything else takes the reference's own divisions below, so accepted hits
// and their t are the reference's bit for bit (the parity tests compare every ray of whole renders).
__device__ __forceinline__ bool tri_clearly_missed(float D, float a, float b) {
  const float s = D < 0.0f ? -1.0f : 1.0f;
  const float Dp = fabsf(D), ap = a * s, bp = b * s;  // exact sign flips
  const float lo = -(Dp * 1e-30f), hi = Dp * 1.000001f;
  return ap < lo || bp < lo || ap > hi || bp > hi || ap + bp > Dp * 1.000002f;
}
__device__ __forceinline__ bool tri_intersect(Ray& ray, V3 v0, V3 v1, V3 v2) {
  V3 E1 = v1 - v0, E2 = v2 - v0;
  V3 T = ray.o - v0;
  V3 P = cross(ray.d, E2);
  V3 Q = cross(T, E1);
  float D = dot(P, E1);
  if (D == 0.0f) return false;
  const float a = dot(P, T), b = dot(Q, ray.d);
  if (tri_clearly_missed(D, a, b)) return false;
  float t = dot(Q, E2) / D;
  if (t <= ray.tmin || t >= ray.tmax) return false;
  float u = a / D;
  if (u < 0.0f || u > 1.0f) return false;
  float v = b / D;
  if (v < 0.0f || v > 1.0f) return false;
  if (u + v > 1.0f) return false;
  ray.tmax = t;
  return true;
}
__device__ __forceinline__ bool tri_hit(const Ray& ray, V3 v0, V3 v1, V3 v2) {
  V3 E1 = v1 - v0, E2 = v2 - v0;
  V3 T = ray.o - v0;
  V3 P = cross(ray.d, E2);
  V3 Q = cross(T, E1);
  float D = dot(P, E1);
  if (D == 0.0f) return false;
  const float a = dot(P, T), b = dot(Q, ray.d);
  if (tri_clearly_missed(D, a, b)) return false;
  float t = dot(Q, E2) / D;
  if (t < ray.tmin || t > ray.tmax) return false;
  float u = a / D;
  if (u < 0.0f || u > 1.0f) return false;
  float v = b / D;
  if (v < 0.0f || v > 1.0f) return false;
  return u + v < 1.0f;
}

// ---- Cone (no base cap) ----------------------------------------------------------------------------
__device__ __forceinline__ bool quadratic_root(float a, float b, float c, float tmin, float& tmax) {
  float d = b * b - 4 * a * c;
  if (d <= 0.0f) return false;
  d = sqrtf(d);
  float t = (-b - d) / (2 * a);
  if (t < tmin) t += d / a;
  if (t < tmin || t > tmax) return false;
  tmax = t;
  return true;
}
__device__ __forceinline__ bool cone_root(const pb2_cone& s, const Ray& ray, float& tmax, V3& o) {
  V3 n = ldv(s.n);
  o = ray.o - ldv(s.p);
  V3 d = ray.d;
  float A2 = s.A2;
  float a = -A2 * sqr(dot(d, n)) + dot(d, d);
  float b = 2 * (-A2 * dot(o, n) * dot(d, n) + dot(o, d));
  float c = -A2 * sqr(dot(o, n)) + dot(o, o);
  tmax = ray.tmax;
  return quadratic_root(a, b, c, ray.tmin, tmax);
}
__device__ __forceinline__ bool cone_intersect(const pb2_cone& s, Ray& ray) {
  float tmax;
  V3 o;
  if (cone_root(s, ray, tmax, o) && dot(o + tmax * ray.d, ldv(s.n)) < 0) {
    ray.tmax = tmax;
    return true;
  }
  return false;
}
__device__ __forceinline__ bool cone_hit(const pb2_cone& s, const Ray& ray) {
  float tmax;
  V3 o;
  return cone_root(s, ray, tmax, o) && dot(o + tmax * ray.d, ldv(s.n)) <= 0;
}

// ---- Plane (geometry.cpp:35-45) ---------------------------------------------------------------------
// intersect accepts t == tmin and t == tmax (it rejects `t < tmin || t > tmax` only) and -- both comparisons
// being false -- a NaN t (ray in the plane); hit is strict on tmin. Restated as written.
__device__ __forceinline__ float plane_t(const pb2_plane& s, const Ray& ray) {
  const V3 n = ldv(s.n);
  return (dot(ldv(s.position), n) - dot(ray.o, n)) / dot(ray.d, n);
}
__device__ __forceinline__ bool plane_intersect(const pb2_plane& s, Ray& ray) {
  const float t = plane_t(s, ray);
  if (t < ray.tmin || t > ray.tmax) return false;
  ray.tmax = t;
  return true;
}
__device__ __forceinline__ bool plane_hit(const pb2_plane& s, const Ray& ray) {
  const float t = plane_t(s, ray);
  if (t <= ray.tmin) return false;
  return t < ray.tmax;
}
// ---- Cylinder (geometry.cpp:466-517): the open tube; the far root is NOT re-checked against tmin ------
__device__ __forceinline__ bool cylinder_root(const pb2_cylinder& s, const Ray& ray, float& t) {
  const V3 p0 = ldv(s.p0), n = ldv(s.n);
  const V3 m = ray.o - p0;
  const V3 v = ray.d - dot(ray.d, n) * n;
  const V3 w = m - dot(m, n) * n;
  const float a = dot(v, v);
  const float b = 2 * dot(v, w);
  const float c = dot(w, w) - s.r * s.r;
  const float discriminant = b * b - 4 * a * c;
  if (discriminant < 0) return false;
  const float sqrt_disc = sqrtf(discriminant);
  t = (-b - sqrt_disc) / (2 * a);
  if (t < ray.tmin) t = (-b + sqrt_disc) / (2 * a);
  if (t > ray.tmax) return false;
  const V3 hit_point = ray.o + t * ray.d;
  const V3 projection = p0 + dot(hit_point - p0, n) * n;
  if (dot(projection - p0, n) < 0 || dot(projection - ldv(s.p1), n) > 0) return false;
  return true;
}
__device__ __forceinline__ bool cylinder_intersect(const pb2_cylinder& s, Ray& ray) {
  float t;
  if (!cylinder_root(s, ray, t)) return false;
  ray.tmax = t;
  return true;
}
__device__ __forceinline__ bool cylinder_hit(const pb2_cylinder& s, const Ray& ray) {
  float t;
  return cylinder_root(s, ray, t);
}

// ---- dispatch over the analytic shapes (Shape::intersect / Shape::hit, geometry.h:304-309) ---------
// Which analytic shapes a traversal kernel has compiled IN (SHAPES): kShapesCommon = everything but Plane and
// Cylinder, kShapesAll = everything, kShapesRect = Rect only. The production kernels are instantiated per level and
// a scene runs the tightest one that covers it, because the cases inlined here decide the register allocation of the
// whole persistent loop: with Plane and Cylinder inlined C5 (triangles + Rects) lost 2.4%; with the AABB / OBB code
// (incl. the per-ray box pre-test of wide_persistent.cuh) compiled out of both kernels and only the Rect case left
// in the shadow kernel C5 gains 3.2% (557 -> 575), C4 2.9%, C3 4.2% (profiles/r2/s26_s27_*, s28_s31_*). The
// reference-order kernels know two masks only (RARE: kShapesAll, else kShapesCommon).
// SHAPES is a bit mask over pb2_shape_kind. Instantiated: the masks below; a scene runs the first one (render.cu
// shapes_level) that contains the kinds of its geometries.
constexpr int kShapeBit(int kind) { return 1 << kind; }
constexpr int kShapesMesh = kShapeBit(PB2_SHAPE_MESH);  // triangles of Mesh geometries (tested inline, not through the switch)
constexpr int kShapesRect = kShapeBit(PB2_SHAPE_RECT);                               // cbox-indirect (C3)
constexpr int kShapesRectMesh = kShapesRect | kShapesMesh;                           // meshes + Rect emitters / walls (C4, C5)
constexpr int kShapesBoxes = kShapeBit(PB2_SHAPE_AABB) | kShapeBit(PB2_SHAPE_OBB);
constexpr int kShapesCbox = kShapesRect | kShapesBoxes | kShapeBit(PB2_SHAPE_CONE);  // the README Cornell box (C1)
constexpr int kShapesCommon = kShapesCbox | kShapesMesh | kShapeBit(PB2_SHAPE_SPHERE) | kShapeBit(PB2_SHAPE_TRIANGLE) |
                              kShapeBit(PB2_SHAPE_DISK);
constexpr int kShapesNoBox = kShapesCommon & ~kShapesBoxes;
constexpr int kShapesUnbounded = kShapeBit(PB2_SHAPE_CONE) | kShapeBit(PB2_SHAPE_PLANE) | kShapeBit(PB2_SHAPE_CYLINDER);
constexpr int kShapesFlat = kShapesNoBox & ~kShapesUnbounded;  // no boxes, nothing that leaves its BVH box (kNoCullBit)
constexpr int kShapesAll = kShapesCommon | kShapeBit(PB2_SHAPE_PLANE) | kShapeBit(PB2_SHAPE_CYLINDER);
constexpr int kShapesRare = kShapesAll & ~kShapesCommon;
template <int SHAPES = kShapesAll>
__device__ __forceinline__ bool shape_intersect(const DScene& sc, int kind, int index, Ray& ray,
                                                bool* inside = nullptr) {
  static_assert(SHAPES > 1, "SHAPES is a kShapes* mask, not a bool");
  if ((SHAPES & ~kShapesMesh) == kShapesRect) {
    float t;
    if (kind != PB2_SHAPE_RECT || !rect_test(sc.rects[index], ray, t)) return false;
    ray.tmax = t;
    return true;
  }
  switch (kind) {
    case PB2_SHAPE_RECT:
      if (SHAPES & kShapesRect) {
        float t;
        if (!rect_test(sc.rects[index], ray, t)) return false;
        ray.tmax = t;
        return true;
      }
      break;
    case PB2_SHAPE_CONE: if (SHAPES & kShapeBit(PB2_SHAPE_CONE)) return cone_intersect(sc.cones[index], ray); break;
    case PB2_SHAPE_OBB: if (SHAPES & kShapeBit(PB2_SHAPE_OBB)) return obb_intersect(sc.obbs[index], ray, inside); break;
    case PB2_SHAPE_AABB: if (SHAPES & kShapeBit(PB2_SHAPE_AABB)) return aabb_intersect(sc.aabbs[index], ray, inside); break;
    case PB2_SHAPE_SPHERE: if (SHAPES & kShapeBit(PB2_SHAPE_SPHERE)) return sphere_intersect(sc.spheres[index], ray); break;
    case PB2_SHAPE_DISK: if (SHAPES & kShapeBit(PB2_SHAPE_DISK)) return disk_intersect(sc.disks[index], ray); break;
    case PB2_SHAPE_TRIANGLE:
      if (SHAPES & kShapeBit(PB2_SHAPE_TRIANGLE)) {
        const pb2_triangle& t = sc.triangles[index];
        return tri_intersect(ray, ldv(t.v0), ldv(t.v1), ldv(t.v2));
      }
      break;
    case PB2_SHAPE_PLANE: if (SHAPES & kShapeBit(PB2_SHAPE_PLANE)) return plane_intersect(sc.planes[index], ray); break;
    case PB2_SHAPE_CYLINDER: if (SHAPES & kShapeBit(PB2_SHAPE_CYLINDER)) return cylinder_intersect(sc.cylinders[index], ray); break;
  }
  return false;
}
template <int SHAPES = kShapesAll>
__device__ __forceinline__ bool shape_hit(const DScene& sc, int kind, int index, const Ray& ray) {
  static_assert(SHAPES > 1, "SHAPES is a kShapes* mask, not a bool");
  if ((SHAPES & ~kShapesMesh) == kShapesRect) {
    float t;
    return kind == PB2_SHAPE_RECT && rect_test(sc.rects[index], ray, t);
  }
  switch (kind) {
    case PB2_SHAPE_RECT:
      if (SHAPES & kShapesRect) {
        float t;
        return rect_test(sc.rects[index], ray, t);
      }
      break;
    case PB2_SHAPE_CONE: if (SHAPES & kShapeBit(PB2_SHAPE_CONE)) return cone_hit(sc.cones[index], ray); break;
    case PB2_SHAPE_OBB: if (SHAPES & kShapeBit(PB2_SHAPE_OBB)) return obb_hit(sc.obbs[index], ray); break;
    case PB2_SHAPE_AABB: if (SHAPES & kShapeBit(PB2_SHAPE_AABB)) return aabb_hit(sc.aabbs[index], ray.o, ray.d, ray.tmin, ray.tmax); break;
    case PB2_SHAPE_SPHERE: if (SHAPES & kShapeBit(PB2_SHAPE_SPHERE)) return sphere_hit(sc.spheres[index], ray); break;
    case PB2_SHAPE_DISK: if (SHAPES & kShapeBit(PB2_SHAPE_DISK)) return disk_hit(sc.disks[index], ray); break;
    case PB2_SHAPE_TRIANGLE:
      if (SHAPES & kShapeBit(PB2_SHAPE_TRIANGLE)) {
        const pb2_triangle& t = sc.triangles[index];
        return tri_hit(ray, ldv(t.v0), ldv(t.v1), ldv(t.v2));
      }
      break;
    case PB2_SHAPE_PLANE: if (SHAPES & kShapeBit(PB2_SHAPE_PLANE)) return plane_hit(sc.planes[index], ray); break;
    case PB2_SHAPE_CYLINDER: if (SHAPES & kShapeBit(PB2_SHAPE_CYLINDER)) return cylinder_hit(sc.cylinders[index], ray); break;
  }
  return false;
}
__device__ __forceinline__ void mesh_tri(const DScene& sc, const DMesh& m, int prim, V3& v0, V3& v1,
                                         V3& v2) {
  const uint32_t* f = sc.tris + 3 * (size_t(m.tri_base) + prim);
  const float* vb = sc.verts + 3 * size_t(m.vert_base);
  v0 = ldv(vb + 3 * size_t(f[0]));
  v1 = ldv(vb + 3 * size_t(f[1]));
  v2 = ldv(vb + 3 * size_t(f[2]));
}

}  // namespace pb2
