// host_shapes.h -- host-side mirror of the reference's shape / camera constructors and matrix
// helpers: turns constructor arguments (what a PRL script writes) into the member records of
// include/pine_b200.h with the same float operations the reference constructors perform, and gives
// the per-shape AABBs BVH::build feeds its builder. Host only; g++ -ffp-contract=off.
//
// Reference: Rect ctor geometry.cpp:255-265; Disk ctor :124-128; Cone ctor :409-414;
// OBB ctor bbox.cpp:143 + inverse(mat4) vecmath.cpp:104-131; ThinLenCamera ctor camera.cpp:7-15;
// translate/scale/rotate_* vecmath.h:1088-1151; look_at :1172-1180; get_aabb(): Rect
// geometry.cpp:400-407, Disk/Sphere :122,172, Cone geometry.h:126, AABB bbox.h:85, OBB bbox.cpp:8-17,
// Triangle geometry.cpp:582-588.
#pragma once
#include "../../include/pine_b200.h"
#include "pmath.h"

namespace pb2 {

inline V3 ld3(const float* p) { return V3(p[0], p[1], p[2]); }
inline void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
inline M4 ld_m4(const float* p) {  // column-major
  return M4(V4(p[0], p[1], p[2], p[3]), V4(p[4], p[5], p[6], p[7]), V4(p[8], p[9], p[10], p[11]),
            V4(p[12], p[13], p[14], p[15]));
}
inline void st_m4(float* p, const M4& m) {
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) p[c * 4 + r] = m[c][r];
}
// mat4(x0,y0,z0,w0, x1,...) takes ROW-major arguments (vecmath.h:559-562)
inline M4 m4_rows(float a0, float a1, float a2, float a3, float b0, float b1, float b2, float b3,
                  float c0, float c1, float c2, float c3, float d0, float d1, float d2, float d3) {
  return M4(V4(a0, b0, c0, d0), V4(a1, b1, c1, d1), V4(a2, b2, c2, d2), V4(a3, b3, c3, d3));
}
inline M4 m4_translate(V3 v) { return m4_rows(1, 0, 0, v.x, 0, 1, 0, v.y, 0, 0, 1, v.z, 0, 0, 0, 1); }
inline M4 m4_scale(V3 v) { return m4_rows(v.x, 0, 0, 0, 0, v.y, 0, 0, 0, 0, v.z, 0, 0, 0, 0, 1); }
inline M4 m4_rotate_x(float r) {
  return m4_rows(1, 0, 0, 0, 0, cosf(r), -sinf(r), 0, 0, sinf(r), cosf(r), 0, 0, 0, 0, 1);
}
inline M4 m4_rotate_y(float r) {
  return m4_rows(cosf(r), 0, sinf(r), 0, 0, 1, 0, 0, -sinf(r), 0, cosf(r), 0, 0, 0, 0, 1);
}
inline M4 m4_rotate_z(float r) {
  return m4_rows(cosf(r), -sinf(r), 0, 0, sinf(r), cosf(r), 0, 0, 0, 0, 1, 0, 0, 0, 0, 1);
}
// Matrix4 product: ret[c][r] += lhs[i][r] * rhs[c][i], i ascending from 0 (vecmath.h:592-599)
inline M4 m4_mul(const M4& a, const M4& b) {
  M4 ret;
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) {
      float s = 0.0f;
      for (int i = 0; i < 4; i++) s += a[i][r] * b[c][i];
      ret[c][r] = s;
    }
  return ret;
}
// q2m(a, b, c, d) of the glTF loader: unit quaternion (w = a, x = b, y = c, z = d) to a rotation matrix
// (fileio.cpp:127-144); every sum is evaluated left to right in binary32.
inline M4 m4_from_quat(float a, float b, float c, float d) {
  return m4_rows(a * a + b * b - c * c - d * d, 2 * b * c - 2 * a * d, 2 * b * d + 2 * a * c, 0,
                 2 * b * c + 2 * a * d, a * a - b * b + c * c - d * d, 2 * c * d - 2 * a * b, 0,
                 2 * b * d - 2 * a * c, 2 * c * d + 2 * a * b, a * a - b * b - c * c + d * d, 0,
                 0, 0, 0, 1);
}
inline M4 m4_transpose(const M4& m) {
  M4 r;
  for (int c = 0; c < 4; c++)
    for (int k = 0; k < 4; k++) r[c][k] = m[k][c];
  return r;
}
// inverse(mat3) (vecmath.cpp:81-103): adjugate entry by entry, then every entry divided by det
inline M3 m3_inverse(const M3& m) {
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
// Cofactor inverse with the reference's index pattern and summation order (vecmath.cpp:104-131).
inline M4 m4_inverse(const M4& m) {
  M4 r;
  auto minor3 = [&](int i, int v) {
    int i1 = (1 + i) % 4, i2 = (2 + i) % 4, i3 = (3 + i) % 4;
    int v1 = (1 + v) % 4, v2 = (2 + v) % 4, v3 = (3 + v) % 4;
    return m[i1][v1] * (m[i2][v2] * m[i3][v3] - m[i3][v2] * m[i2][v3]) +
           m[i2][v1] * (m[i3][v2] * m[i1][v3] - m[i1][v2] * m[i3][v3]) +
           m[i3][v1] * (m[i1][v2] * m[i2][v3] - m[i2][v2] * m[i1][v3]);
  };
  float det = 0;
  for (int i = 0; i < 4; i++) det += minor3(i, 3) * m[i % 4][3] * float(i % 2 ? -1 : 1);
  if (det == 0) return r;
  for (int v = 0; v < 4; v++)
    for (int i = 0; i < 4; i++) r[v][i] = minor3(i, v) * float((v + i) % 2 ? 1 : -1);
  for (int c = 0; c < 4; c++)
    for (int rr = 0; rr < 4; rr++) r[c][rr] /= det;
  return r;
}
inline M4 m4_look_at(V3 from, V3 at, V3 up = V3(0, 1, 0)) {
  V3 z = normalize(at - from);
  if (pabs(dot(z, up)) > 0.999f) z = normalize(z + V3(0.0f, 0.0f, 1e-5f));
  V3 x = normalize(cross(up, z));
  V3 y = cross(z, x);
  return M4(V4(x, 0), V4(y, 0), V4(z, 0), V4(from, 1.0f));
}

struct Box3 {  // pine::AABB value semantics (bbox.h:29-92)
  V3 lower = V3(kFloatMax), upper = V3(-kFloatMax);
  void extend(V3 p) { lower = vmin(lower, p); upper = vmax(upper, p); }
  void extend(const Box3& b) { lower = vmin(lower, b.lower); upper = vmax(upper, b.upper); }
  V3 centroid() const { return (lower + upper) / 2.0f; }
  float centroid(int d) const { return (lower[d] + upper[d]) / 2.0f; }
  bool degenerated(int d) const { return upper[d] <= lower[d]; }
  float relative_position(float p, int d) const {
    float o = p - lower[d];
    float dd = upper[d] - lower[d];
    return dd > 0.0f ? o / dd : o;
  }
  float surface_area() const {
    V3 d = upper - lower;
    return 2.0f * (d.x * d.y + d.x * d.z + d.y * d.z);
  }
};
inline Box3 box_union(const Box3& a, const Box3& b) {
  Box3 r;
  r.lower = vmin(a.lower, b.lower);
  r.upper = vmax(a.upper, b.upper);
  return r;
}

inline pb2_rect make_rect(V3 position, V3 ex_in, V3 ey_in, bool flip) {
  pb2_rect r;
  V3 ex = normalize(ex_in), ey = normalize(ey_in);
  V3 n = normalize(cross(ex, ey)) * float(flip ? -1 : 1);
  float lx = length(ex_in), ly = length(ey_in);
  st3(r.position, position); st3(r.ex, ex); st3(r.ey, ey); st3(r.n, n);
  r.lx = lx; r.ly = ly;
  st3(r.rx, ex / lx); st3(r.ry, ey / ly);
  return r;
}
inline pb2_disk make_disk(V3 position, V3 normal, float radius) {
  pb2_disk d;
  V3 n = normalize(normal);
  M3 tbn = coordinate_system(n);
  st3(d.position, position); st3(d.n, n); st3(d.u, tbn.x); st3(d.v, tbn.y);
  d.r = radius;
  return d;
}
inline pb2_cone make_cone(V3 p, V3 n, float r, float h) {
  pb2_cone c;
  c.bottom = make_disk(p, n, r);
  st3(c.p, p + n * h); st3(c.n, normalize(n));
  c.r = r; c.h = h;
  c.A2 = sqr(r / h) + 1;
  c.A = sqrtf(c.A2);
  c.S = r / sqrtf(r * r + h * h);
  return c;
}
inline pb2_plane make_plane(V3 position, V3 normal) {  // Plane::Plane (geometry.cpp:31-34)
  pb2_plane q;
  V3 n = normalize(normal);
  M3 tbn = coordinate_system(n);
  st3(q.position, position); st3(q.n, n); st3(q.u, tbn.x); st3(q.v, tbn.y);
  return q;
}
inline pb2_cylinder make_cylinder(V3 p0, V3 p1, float r) {  // Cylinder::Cylinder (geometry.h:140-141)
  pb2_cylinder c;
  V3 n = normalize(p1 - p0);
  st3(c.p0, p0); st3(c.p1, p1); st3(c.n, n);
  c.r = r;
  c.bottom = make_disk(p0, -n, r);
  c.top = make_disk(p0, n, r);  // the reference puts BOTH disks at p0
  return c;
}
inline pb2_obb make_obb(V3 lo, V3 hi, const M4& m) {
  pb2_obb o;
  st3(o.base.lower, lo); st3(o.base.upper, hi);
  st_m4(o.m, m);
  st_m4(o.m_inv, m4_inverse(m));
  return o;
}
inline pb2_triangle make_triangle(V3 v0, V3 v1, V3 v2) {
  pb2_triangle t;
  V3 n = normalize(cross(v0 - v1, v0 - v2));
  if (n == V3(0.0f)) n = V3(0, 0, 1);
  st3(t.v0, v0); st3(t.v1, v1); st3(t.v2, v2); st3(t.n, n);
  return t;
}
inline pb2_camera make_camera(int w, int h, int tonemapper, V3 from, V3 to, float fov,
                              float len_radius, float focus) {
  pb2_camera c;
  st3(c.position, from);
  M3 c2w = upper3(m4_look_at(from, to));
  for (int k = 0; k < 3; k++) st3(c.c2w + 3 * k, c2w[k]);
  float aspect = float(w) / h;
  c.fov2d[0] = fov * aspect; c.fov2d[1] = fov;
  c.len_radius = len_radius; c.focus_distance = focus;
  c.width = w; c.height = h; c.tonemapper = tonemapper;
  return c;
}

// ---- AABBs as BVH::build sees them ---------------------------------------------------------------
inline Box3 aabb_of(const pb2_rect& r) {
  V3 p = ld3(r.position), ex = ld3(r.ex), ey = ld3(r.ey);
  Box3 b;
  b.extend(p - ex * r.lx / 2.0f - ey * r.ly / 2.0f);
  b.extend(p - ex * r.lx / 2.0f + ey * r.ly / 2.0f);
  b.extend(p + ex * r.lx / 2.0f - ey * r.ly / 2.0f);
  b.extend(p + ex * r.lx / 2.0f + ey * r.ly / 2.0f);
  return b;
}
inline Box3 aabb_of_sphere(V3 c, float r) {
  Box3 b;
  b.lower = c - V3(r); b.upper = c + V3(r);
  return b;
}
inline Box3 aabb_of(const pb2_sphere& s) { return aabb_of_sphere(ld3(s.c), s.r); }
inline Box3 aabb_of(const pb2_disk& d) { return aabb_of_sphere(ld3(d.position), d.r); }
inline Box3 aabb_of(const pb2_plane& q) {  // Plane::get_aabb (geometry.cpp:52)
  Box3 b;
  b.lower = ld3(q.position) + V3(-100.0f); b.upper = ld3(q.position) + V3(100.0f);
  return b;
}
inline Box3 aabb_of(const pb2_cylinder& c) {  // union_(bottom.get_aabb(), top.get_aabb()) (geometry.h:147)
  Box3 b = aabb_of(c.bottom);
  b.extend(aabb_of(c.top));
  return b;
}
inline Box3 aabb_of(const pb2_cone& c) {
  Box3 b = aabb_of(c.bottom);
  b.extend(ld3(c.p));
  return b;
}
inline Box3 aabb_of(const pb2_aabb& a) {  // AABB(*this).extend_by(epsilon)
  Box3 b;
  b.lower = ld3(a.lower) - V3(kEpsilon);
  b.upper = ld3(a.upper) + V3(kEpsilon);
  return b;
}
inline Box3 aabb_of(const pb2_obb& o) {
  Box3 b;
  M4 m = ld_m4(o.m);
  for (int i = 0; i < 8; i++) {
    V3 p = ld3(o.base.lower);
    if (i % 2 >= 1) p[0] = o.base.upper[0];
    if (i % 4 >= 2) p[1] = o.base.upper[1];
    if (i % 8 >= 4) p[2] = o.base.upper[2];
    b.extend(mul_point(m, p));
  }
  return b;
}
inline Box3 aabb_of_tri(V3 v0, V3 v1, V3 v2) {
  Box3 b;
  b.extend(v0); b.extend(v1); b.extend(v2);
  return b;
}
inline Box3 aabb_of(const pb2_triangle& t) { return aabb_of_tri(ld3(t.v0), ld3(t.v1), ld3(t.v2)); }

}  // namespace pb2
