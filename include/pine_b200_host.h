/* pine_b200_host.h -- host-side mirror of the reference's constructors, exported with C linkage.
 *
 * include/pine_b200.h takes the MEMBERS of Pine's shape structs. A caller that owns real
 * pine::Scene objects copies the members (integration/cuda_path_integrator.cpp). A caller that
 * only has constructor arguments (what a PRL script writes: Rect(pos, ex, ey), Box(aabb, m),
 * Cone(p, n, r, h), ThinLenCamera(film, from, to, fov)) uses these helpers, which perform the
 * reference constructors' float operations in the same order (citations in
 * pine_b200/csrc/host_shapes.h). Matrices are column-major, 16 floats.
 */
#ifndef PINE_B200_HOST_H
#define PINE_B200_HOST_H
#include "pine_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
enum pb2h_mat_op {
  PB2H_TRANSLATE = 0, /* a = v[3]            vecmath.h:1098 */
  PB2H_SCALE = 1,     /* a = v[3]            vecmath.h:1117 */
  PB2H_ROTATE_X = 2,  /* a = {rad}           vecmath.h:1135 */
  PB2H_ROTATE_Y = 3,  /* a = {rad}           vecmath.h:1144 */
  PB2H_ROTATE_Z = 4,  /* a = {rad}           vecmath.h:1126 */
  PB2H_INVERSE = 5,   /* a = m[16]           vecmath.cpp:104 */
  PB2H_LOOK_AT = 6,   /* a = from[3], to[3]  vecmath.h:1172 */
  PB2H_QUAT = 7,      /* a = {w, x, y, z}    q2m, fileio.cpp:127-144 */
  PB2H_TRANSPOSE = 8  /* a = m[16] */
};
void pb2h_make_rect(const float* pos, const float* ex, const float* ey, int flip, pb2_rect* out);
void pb2h_make_obb(const float* lo, const float* hi, const float* m_colmajor, pb2_obb* out);
void pb2h_make_disk(const float* p, const float* n, float r, pb2_disk* out);
void pb2h_make_cone(const float* p, const float* n, float r, float h, pb2_cone* out);
void pb2h_make_plane(const float* position, const float* normal, pb2_plane* out);   /* geometry.cpp:31-34 */
void pb2h_make_cylinder(const float* p0, const float* p1, float r, pb2_cylinder* out); /* geometry.h:140-141 */
void pb2h_make_triangle(const float* a, const float* b, const float* c, pb2_triangle* out);
void pb2h_make_camera(int w, int h, int tonemapper, const float* from, const float* to, float fov,
                      float len_radius, float focus, pb2_camera* out);
void pb2h_mat4(int op, const float* a, float* out_colmajor);
void pb2h_mat4_mul(const float* a, const float* b, float* out_colmajor);
/* Mesh::apply(m) (geometry.cpp:652-658) in place: points by m, normals by transpose(inverse(mat3(m))),
 * renormalised. Used by the glTF loader (scene_from_gltf, fileio.cpp:146-327). */
void pb2h_mesh_apply(const float* m_colmajor, float* vertices, size_t n_vertices, float* normals,
                     size_t n_normals);
float pb2h_cosf(float x); /* the host libm's cosf, as SpotLight's constructor uses (light.cpp:23-24) */
int pb2h_round_spp(int spp);
/* Host evaluation of the device draw functions of SobolSampler (sampler.h:83-164, sampler.cpp:82-113)
 * and UniformSampler (the same __host__ __device__ code the kernels inline; test hook, no GPU):
 * init({width, height}), start_pixel((px, py), sample_index), then n_samples x the pattern of
 * '1' (get1d) / '2' (get2d) draws. sampler = PB2_SAMPLER_SOBOL or PB2_SAMPLER_UNIFORM. */
int pb2h_sampler_draws(int sampler, int spp, int sampler_seed, int width, int height, int px, int py,
                       int sample_index, const char* pattern, int n_samples, float* out_draws);
#ifdef __cplusplus
}
#endif
#endif
