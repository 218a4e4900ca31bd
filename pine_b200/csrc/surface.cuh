// surface.cuh -- surface reconstruction at a hit (compute_surface_info), shape sampling for
// next-event estimation (Shape::sample) and the hit-light pdf (Shape::pdf), device side.
//
// Reference: compute_surface_info -- Rect geometry.cpp:305-312, AABB bbox.cpp:125-132,
// OBB bbox.cpp:176-180, Sphere geometry.cpp:102-106, Disk :151-157, Triangle :566-572,
// Cone :454-459, Mesh :631-646; inverse(mat3) vecmath.cpp:81-103.
// Shape::sample/pdf -- geometry.h:327-347 (rejects pdf<=0 / inf); Rect geometry.cpp:313-320,368-369
// (pdf() MULTIPLIES by the cosine where sample() divides -- reproduced);
// Disk :159-171; Sphere :107-121; Triangle :573-581; Mesh geometry.h:166-182; Cone :461-464.
#pragma once
#include "shapes.cuh"

namespace pb2 {

struct Surf {
  V3 p, n;
  V2 uv;
};

__device__ __forceinline__ M3 inverse3(const M3& m) {
  M3 r;
  float det = m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2]) +
              m[1][0] * (m[2][1] * m[0][2] - m[0][1] * m[2][2]) +
              m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]);
  if (det == 0) return r;
  r[0][0] = m[1][1] * m[2][2] - m[2][1] * m[1][2];
  r[0][1] = m[2][1] * m[0][2] - m[0][1] * m[2][2];
  r[0][2] = m[0][1] * m[1][2] - m[1][1] * m[0][2];
  r[1][0] = m[1][2] * m[2][0] - m[2][2] * m[1][0];
  r[1][1] = m[2][2] * m[0][0] - m[0][2] * m[2][0];
  r[1][2] = m[0][2] * m[1][0] - m[1][2] * m[0][0];
  r[2][0] = m[1][0] * m[2][1] - m[2][0] * m[1][1];
  r[2][1] = m[2][0] * m[0][1] - m[0][0] * m[2][1];
  r[2][2] = m[0][0] * m[1][1] - m[1][0] * m[0][1];
  for (int c = 0; c < 3; c++)
    for (int k = 0; k < 3; k++) r[c][k] /= det;
  return r;
}

__device__ __forceinline__ float phi2pi(float x, float y) {
  float phi = atan2f(y, x);
  return phi < 0.0f ? kPi * 2 + phi : phi;
}

__device__ __forceinline__ void aabb_surface(const pb2_aabb& s, V3 p, Surf& it) {
  V3 lo = ldv(s.lower), hi = ldv(s.upper);
  it.p = p;
  V3 pu = (p - (lo + hi) / 2.0f) / (hi - lo);
  int axis = max_axis(vabs(pu));
  it.n = V3(0.0f);
  it.n[axis] = pu[axis] > 0 ? 1.0f : -1.0f;
  it.p[axis] = pu[axis] > 0 ? hi[axis] : lo[axis];
}

__device__ inline void surface_info(const DScene& sc, int geom, int prim, V3 p, Surf& it) {
  DGeom g = sc.geoms[geom];
  it.p = V3();
  it.n = V3();
  it.uv = V2();
  switch (g.kind) {
    case PB2_SHAPE_RECT: {
      const pb2_rect& s = sc.rects[g.index];
      V3 rp = p - ldv(s.position);
      float u = dot(rp, ldv(s.rx));
      float v = dot(rp, ldv(s.ry));
      it.p = ldv(s.position) + s.lx * ldv(s.ex) * u + s.ly * ldv(s.ey) * v;
      it.n = ldv(s.n);
      it.uv = V2(u, v) + V2(0.5f, 0.5f);
    } break;
    case PB2_SHAPE_AABB: aabb_surface(sc.aabbs[g.index], p, it); break;
    case PB2_SHAPE_OBB: {
      const pb2_obb& s = sc.obbs[g.index];
      M4 m = ld_m4_dev(s.m), m_inv = ld_m4_dev(s.m_inv);
      aabb_surface(s.base, mul_point(m_inv, p), it);
      it.p = mul_point(m, it.p);
      it.n = normalize(transpose(upper3(m_inv)) * it.n);
    } break;
    case PB2_SHAPE_SPHERE: {
      const pb2_sphere& s = sc.spheres[g.index];
      it.n = normalize(p - ldv(s.c));
      it.p = ldv(s.c) + it.n * s.r;
      it.uv = V2(phi2pi(it.n.x, it.n.y), acosf(it.n.z));
    } break;
    case PB2_SHAPE_DISK: {
      const pb2_disk& s = sc.disks[g.index];
      it.n = ldv(s.n);
      float ex = dot(p - ldv(s.position), ldv(s.u));
      float ey = dot(p - ldv(s.position), ldv(s.v));
      it.uv = V2(ex, ey);
      it.p = ldv(s.position) + ex * ldv(s.u) + ey * ldv(s.v);
    } break;
    case PB2_SHAPE_TRIANGLE: {
      const pb2_triangle& s = sc.triangles[g.index];
      V3 v0 = ldv(s.v0), v1 = ldv(s.v1), v2 = ldv(s.v2);
      float u = dot(p - v0, v1 - v0);
      float v = dot(p - v0, v2 - v0);
      it.uv = V2(u, v);
      it.p = lerp3(u, v, v0, v1, v2);
      it.n = ldv(s.n);
    } break;
    case PB2_SHAPE_CONE: {
      const pb2_cone& s = sc.cones[g.index];
      V3 apex = ldv(s.p), n = ldv(s.n);
      float l = length(p - apex) * s.A;
      V3 x = apex - n * l;
      it.n = normalize(p - x);
      it.p = x + it.n * l * s.S;
    } break;
    case PB2_SHAPE_PLANE: {  // Plane::compute_surface_info (geometry.cpp:46-51)
      const pb2_plane& s = sc.planes[g.index];
      const V3 pos = ldv(s.position), u = ldv(s.u), v = ldv(s.v);
      it.n = ldv(s.n);
      const V3 dp = p - pos;
      it.uv = V2(dot(dp, u), dot(dp, v));
      it.p = pos + it.uv.x * u + it.uv.y * v;
    } break;
    case PB2_SHAPE_CYLINDER: {  // what Cylinder::intersect leaves in `it` (geometry.cpp:511-513); uv stays 0
      const pb2_cylinder& s = sc.cylinders[g.index];
      const V3 p0 = ldv(s.p0), n = ldv(s.n);
      const V3 projection = p0 + dot(p - p0, n) * n;
      it.n = normalize(p - projection);
      it.p = p;
    } break;
    case PB2_SHAPE_MESH: {
      const DMesh& m = sc.meshes[g.index];
      const uint32_t* f = sc.tris + 3 * (size_t(m.tri_base) + prim);
      const float* vb = sc.verts + 3 * size_t(m.vert_base);
      V3 v0 = ldv(vb + 3 * size_t(f[0])), v1 = ldv(vb + 3 * size_t(f[1])), v2 = ldv(vb + 3 * size_t(f[2]));
      V3 e1 = v1 - v0, e2 = v2 - v0;
      it.n = cross(e1, e2);
      M3 tbn = inverse3(M3(e1, e2, it.n));
      V3 q = tbn * (p - v0);
      it.uv = V2(q.x, q.y);
      it.p = lerp3(it.uv.x, it.uv.y, v0, v1, v2);
      if (m.has_normals) {
        const float* nb = sc.normals + 3 * size_t(m.vert_base);
        it.n = normalize(lerp3(it.uv.x, it.uv.y, ldv(nb + 3 * size_t(f[0])), ldv(nb + 3 * size_t(f[1])),
                               ldv(nb + 3 * size_t(f[2]))));
      } else {
        it.n = normalize(it.n);
      }
      if (m.has_uvs) {
        const float* tb = sc.uvs + 2 * size_t(m.vert_base);
        V2 t0(tb[2 * size_t(f[0])], tb[2 * size_t(f[0]) + 1]);
        V2 t1(tb[2 * size_t(f[1])], tb[2 * size_t(f[1]) + 1]);
        V2 t2(tb[2 * size_t(f[2])], tb[2 * size_t(f[2]) + 1]);
        it.uv = lerp3(it.uv.x, it.uv.y, t0, t1, t2);
      }
    } break;
  }
}

// ---- Shape::sample(p, u, u1) ------------------------------------------------------------------------
struct ShapeSample {
  V3 p, n, w;
  float distance, pdf;
};
__device__ __forceinline__ V2 sample_disk_concentric(V2 u) {  // sampling.h:20-35
  u = V2(u.x * 2 - 1.0f, u.y * 2 - 1.0f);
  float theta, r;
  if (pabs(u.x) > pabs(u.y)) {
    r = u.x;
    theta = kPi / 4.0f * u.y / u.x;
  } else {
    r = u.y;
    theta = kPi / 2.0f - kPi / 4.0f * (u.x / u.y);
  }
  return r * V2(cosf(theta), sinf(theta));
}
__device__ __forceinline__ float tri_area(V3 v0, V3 v1, V3 v2) { return length(cross(v1 - v0, v2 - v0)) / 2; }
__device__ __forceinline__ bool tri_sample(V3 v0, V3 v1, V3 v2, V3 p, V2 u, ShapeSample& ss) {
  if (u.x + u.y > 1.0f) u = V2(1.0f - u.x, 1.0f - u.y);
  ss.p = lerp3(u.x, u.y, v0, v1, v2);
  V3 n = normalize(cross(v0 - v1, v0 - v2));  // Triangle(v0,v1,v2) ctor, geometry.cpp:536-539
  if (n == V3(0.0f)) n = V3(0, 0, 1);
  ss.n = n;
  ss.w = normalize(ss.p - p, ss.distance);
  ss.pdf = sqr(ss.distance) / pmax(absdot(ss.w, ss.n) * tri_area(v0, v1, v2), kEpsilon);
  return true;
}
__device__ inline bool shape_sample(const DScene& sc, int geom, V3 p, V2 u, float u1, ShapeSample& ss) {
  DGeom g = sc.geoms[geom];
  bool ok = false;
  ss.pdf = 0;
  switch (g.kind) {
    case PB2_SHAPE_RECT: {
      const pb2_rect& s = sc.rects[g.index];
      ss.p = ldv(s.position) + (u.x - 0.5f) * ldv(s.ex) * s.lx + (u.y - 0.5f) * ldv(s.ey) * s.ly;
      ss.n = ldv(s.n);
      ss.w = normalize(ss.p - p, ss.distance);
      ss.pdf = sqr(ss.distance) / (absdot(ss.w, ss.n) * (s.lx * s.ly));
      ok = true;
    } break;
    case PB2_SHAPE_DISK: {
      const pb2_disk& s = sc.disks[g.index];
      V2 uv = sample_disk_concentric(u);
      ss.p = ldv(s.position) + s.r * ldv(s.u) * uv.x + s.r * ldv(s.v) * uv.y;
      ss.n = ldv(s.n);
      ss.w = normalize(ss.p - p, ss.distance);
      ss.pdf = sqr(ss.distance) / pmax(absdot(ss.w, ss.n) * (kPi * s.r * s.r), kEpsilon);
      ok = true;
    } break;
    case PB2_SHAPE_SPHERE: {
      const pb2_sphere& s = sc.spheres[g.index];
      V3 c = ldv(s.c);
      float l = length(c - p);
      float cos_theta = sqrtf(1 - sqr(s.r / l));
      float S = 2 * kPi * (1 - cos_theta);
      float cos_wo = 1 - u.y * (1 - cos_theta);
      float sin_wo = sqrtf(1 - cos_wo * cos_wo);
      float phi = u.x * 2 * kPi;
      V3 w = V3(sin_wo * cosf(phi), sin_wo * sinf(phi), cos_wo);
      ss.w = coordinate_system((c - p) / l) * w;
      ss.distance = sphere_t(p, ss.w, 0.0f, c, s.r);
      ss.pdf = 1.0f / S;
      ss.p = p + ss.w * ss.distance;
      ss.n = (ss.p - c) / s.r;
      ok = true;
    } break;
    case PB2_SHAPE_TRIANGLE: {
      const pb2_triangle& s = sc.triangles[g.index];
      ok = tri_sample(ldv(s.v0), ldv(s.v1), ldv(s.v2), p, u, ss);
      ss.n = ldv(s.n);
    } break;
    case PB2_SHAPE_MESH: {
      const DMesh& m = sc.meshes[g.index];
      if (m.n_tris == 0) return false;
      V3 v0, v1, v2;
      mesh_tri(sc, m, int(m.n_tris * u1), v0, v1, v2);
      ok = tri_sample(v0, v1, v2, p, u, ss);
      ss.pdf /= m.n_tris;
    } break;
    default: return false;  // Cone::sample returns nullopt; Box samples have pdf 0 and are rejected
  }
  if (!ok || ss.pdf <= 0 || isinf(ss.pdf)) return false;
  return true;
}

// Shape::pdf(ray, ps, ns) with ray.tmax = hit distance (lightsampler.cpp:27-29).
__device__ inline float shape_pdf(const DScene& sc, int geom, V3 ray_o, V3 ray_d, float tmax, V3 ns) {
  DGeom g = sc.geoms[geom];
  switch (g.kind) {
    case PB2_SHAPE_RECT: {
      const pb2_rect& s = sc.rects[g.index];
      return sqr(tmax) / (s.lx * s.ly) * absdot(ns, ray_d);
    }
    case PB2_SHAPE_DISK: {
      const pb2_disk& s = sc.disks[g.index];
      return sqr(tmax) / ((kPi * s.r * s.r) * absdot(ns, ray_d));
    }
    case PB2_SHAPE_SPHERE: {
      const pb2_sphere& s = sc.spheres[g.index];
      float l = length(ldv(s.c) - ray_o);
      float cos_theta = sqrtf(1 - sqr(s.r / l));
      float S = 2 * kPi * (1 - cos_theta);
      return 1.0f / S;
    }
    case PB2_SHAPE_TRIANGLE: {
      const pb2_triangle& s = sc.triangles[g.index];
      return sqr(tmax) / (tri_area(ldv(s.v0), ldv(s.v1), ldv(s.v2)) * absdot(ns, ray_d));
    }
    case PB2_SHAPE_CONE: {
      const pb2_cone& s = sc.cones[g.index];
      float area = sqrtf(s.r * s.r + s.h * s.h) * kPi * s.r + kPi * s.bottom.r * s.bottom.r;
      return sqr(tmax) / area * absdot(ns, ray_d);
    }
    case PB2_SHAPE_MESH: {
      const DMesh& m = sc.meshes[g.index];
      V3 v0, v1, v2;
      mesh_tri(sc, m, 0, v0, v1, v2);
      float area = tri_area(v0, v1, v2) * m.n_tris;  // Mesh::area(): first triangle x count
      return sqr(tmax) / (area * absdot(ns, ray_d));
    }
  }
  return 0.0f;
}

}  // namespace pb2
