// host_api.cpp -- C exports of the host-side constructor mirror (host_shapes.h) for language
// bindings; declared in include/pine_b200_host.h. Pure host code, no CUDA.
#include "../../include/pine_b200_host.h"

#include <math.h>

#include "host_shapes.h"

using namespace pb2;

extern "C" {
void pb2h_make_rect(const float* pos, const float* ex, const float* ey, int flip, pb2_rect* out) {
  *out = make_rect(ld3(pos), ld3(ex), ld3(ey), flip != 0);
}
void pb2h_make_obb(const float* lo, const float* hi, const float* m_colmajor, pb2_obb* out) {
  *out = make_obb(ld3(lo), ld3(hi), ld_m4(m_colmajor));
}
void pb2h_make_plane(const float* position, const float* normal, pb2_plane* out) {
  *out = make_plane(ld3(position), ld3(normal));
}
void pb2h_make_cylinder(const float* p0, const float* p1, float r, pb2_cylinder* out) {
  *out = make_cylinder(ld3(p0), ld3(p1), r);
}
void pb2h_make_disk(const float* p, const float* n, float r, pb2_disk* out) {
  *out = make_disk(ld3(p), ld3(n), r);
}
void pb2h_make_cone(const float* p, const float* n, float r, float h, pb2_cone* out) {
  *out = make_cone(ld3(p), ld3(n), r, h);
}
void pb2h_make_triangle(const float* a, const float* b, const float* c, pb2_triangle* out) {
  *out = make_triangle(ld3(a), ld3(b), ld3(c));
}
void pb2h_make_camera(int w, int h, int tonemapper, const float* from, const float* to, float fov,
                      float len_radius, float focus, pb2_camera* out) {
  *out = make_camera(w, h, tonemapper, ld3(from), ld3(to), fov, len_radius, focus);
}
void pb2h_mat4(int op, const float* a, float* out_colmajor) {
  M4 m;
  switch (op) {
    case PB2H_TRANSLATE: m = m4_translate(ld3(a)); break;
    case PB2H_SCALE: m = m4_scale(ld3(a)); break;
    case PB2H_ROTATE_X: m = m4_rotate_x(a[0]); break;
    case PB2H_ROTATE_Y: m = m4_rotate_y(a[0]); break;
    case PB2H_ROTATE_Z: m = m4_rotate_z(a[0]); break;
    case PB2H_INVERSE: m = m4_inverse(ld_m4(a)); break;
    case PB2H_LOOK_AT: m = m4_look_at(ld3(a), ld3(a + 3)); break;
    case PB2H_QUAT: m = m4_from_quat(a[0], a[1], a[2], a[3]); break;
    case PB2H_TRANSPOSE: m = m4_transpose(ld_m4(a)); break;
  }
  st_m4(out_colmajor, m);
}
void pb2h_mat4_mul(const float* a, const float* b, float* out) { st_m4(out, m4_mul(ld_m4(a), ld_m4(b))); }
void pb2h_mesh_apply(const float* m_colmajor, float* vertices, size_t n_vertices, float* normals,
                     size_t n_normals) {  // Mesh::apply (geometry.cpp:652-658)
  const M4 m = ld_m4(m_colmajor);
  for (size_t i = 0; i < n_vertices; i++) st3(vertices + 3 * i, mul_point(m, ld3(vertices + 3 * i)));
  const M3 m3 = transpose(m3_inverse(upper3(m)));
  for (size_t i = 0; i < n_normals; i++) st3(normals + 3 * i, normalize(m3 * ld3(normals + 3 * i)));
}
float pb2h_cosf(float x) { return cosf(x); }  // psl::cos (SpotLight ctor, light.cpp:23-24)
int pb2h_round_spp(int spp) {  // BlueSobolSampler ctor (sampler.cpp:115-121): clamp 256, roundup2
  if (spp > 256) spp = 256;
  if (spp < 1) return 0;
  int r = 1;
  while (r < spp) r <<= 1;
  return r;
}
}
