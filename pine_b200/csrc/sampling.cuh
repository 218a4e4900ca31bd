// sampling.cuh -- per-pixel RNG, the blue-noise Sobol sampler, the camera and the sampling warps.
//
// Reference: murmur_hash64A / hash / split_mix_64 / RNG rng.h:9-45,60-65,74-79,97-144;
// Sampler wrapper sampler.h:286-309; BlueSobolSampler sampler.h:166-201 + bluenoise_<N>spp()
// src/contrib/bluesobol/bluenoise_16spp.cpp:14-34; ThinLenCamera::gen_ray camera.cpp:17-33;
// warps sampling.h:13-74; spawn_ray / offset_ray_origin interaction.cpp:6-13, ray.h:25-37.
#pragma once
#include "device_scene.h"

namespace pb2 {

// ---- RNG (u64 integer arithmetic, exact) ------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t murmur_pixel(int px, int py, int sample_index) {
  // hash(vec2i p, int sample_index): 12 bytes, seed 0
  const uint64_t m = 0xc6a4a7935bd1e995ull;
  const int r = 47;
  uint64_t h = 0 ^ (12ull * m);
  uint64_t k = uint64_t(uint32_t(px)) | (uint64_t(uint32_t(py)) << 32);
  k *= m; k ^= k >> r; k *= m;
  h ^= k; h *= m;
  h ^= uint64_t(uint32_t(sample_index));  // tail bytes 3..0
  h *= m;
  h ^= h >> r; h *= m; h ^= h >> r;
  return h;
}
__host__ __device__ __forceinline__ uint64_t split_mix_64(uint64_t& s) {
  uint64_t r = s += 0x9E3779B97f4A7C15ull;
  r = (r ^ (r >> 30)) * 0xBF58476D1CE4E5B9ull;
  r = (r ^ (r >> 27)) * 0x94D049BB133111EBull;
  return r ^ (r >> 31);
}
struct Rng {
  uint64_t s0, s1;
};
__host__ __device__ __forceinline__ Rng rng_seed(uint64_t seed) {
  Rng g;
  g.s0 = split_mix_64(seed);
  g.s1 = split_mix_64(seed);
  return g;
}
__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
__host__ __device__ __forceinline__ uint64_t rng_next(Rng& g) {
  const uint64_t a = g.s0;
  uint64_t b = g.s1;
  const uint64_t result = a + b;
  b ^= a;
  g.s0 = rotl64(a, 24) ^ b ^ (b << 16);
  g.s1 = rotl64(b, 37);
  return result;
}
__host__ __device__ __forceinline__ float rng_nextf(Rng& g) {
  uint64_t u = rng_next(g);
  return pmin(float(uint32_t(u ^ (u >> 32))) * 0x1p-32f, kOneMinusEpsilon);
}

// ---- BlueSobolSampler ---------------------------------------------------------------------------------
struct BlueSampler {
  int px, py, index, dim;
};
__device__ __forceinline__ float blue_dim(const DScene& sc, const BlueSampler& s, int dim) {
  int pi = s.px & 127, pj = s.py & 127;
  int si = s.index & 255, sd = dim & 255;
  int tile = (pi + pj * 128) * 8;
  int ranked = si ^ int(__ldg(sc.rank + ((sd + tile) % (128 * 128 * 8))));
  int value = int(__ldg(sc.sobol + sd + ranked * 256));
  value ^= int(__ldg(sc.scramble + (sd % 8) + tile));
  return (0.5f + value) / 256.0f;
}
__device__ __forceinline__ float blue_get1d(const DScene& sc, BlueSampler& s) {
  if (s.dim >= 256) s.dim = 2;
  return blue_dim(sc, s, s.dim++);
}
__device__ __forceinline__ V2 blue_get2d(const DScene& sc, BlueSampler& s) {
  if (s.dim + 1 >= 256) s.dim = 2;
  int d = s.dim;
  s.dim += 2;
  return V2(blue_dim(sc, s, d), blue_dim(sc, s, d + 1));
}

// ---- camera ---------------------------------------------------------------------------------------------
__device__ __forceinline__ V2 sample_disk_polar(V2 u) {
  float r = sqrtf(u.x);
  float theta = 2 * kPi * u.y;
  return V2(r * cosf(theta), r * sinf(theta));
}
__device__ __forceinline__ Ray camera_ray(const pb2_camera& c, V2 p_film, V2 u2) {
  M3 c2w(V3(c.c2w[0], c.c2w[1], c.c2w[2]), V3(c.c2w[3], c.c2w[4], c.c2w[5]),
         V3(c.c2w[6], c.c2w[7], c.c2w[8]));
  V3 position(c.position[0], c.position[1], c.position[2]);
  V2 pc = (p_film - V2(0.5f, 0.5f)) * 2.0f;
  pc = pc * V2(c.fov2d[0], c.fov2d[1]);
  Ray r;
  r.tmin = 0.0f;
  r.tmax = kFloatMax;
  if (c.len_radius == 0.0f) {
    r.o = position;
    r.d = normalize(c2w * V3(pc.x, pc.y, 1.0f));
  } else {
    V3 dir = normalize(V3(pc.x, pc.y, 1.0f));
    V3 p_focus = c.focus_distance * dir / dir.z;
    V2 d2 = c.len_radius * sample_disk_polar(u2);
    V3 p_len(d2.x, d2.y, 0.0f);
    r.o = position + p_len;
    r.d = c2w * normalize(p_focus - p_len);
  }
  return r;
}

// ---- spawn_ray --------------------------------------------------------------------------------------------
__device__ __forceinline__ V3 offset_ray_origin(V3 p, V3 n) {
  const float origin = 1.0f / 32.0f;
  const float float_scale = 1.0f / 65536.0f;
  const float int_scale = 256.0f;
  int ox = int(int_scale * n.x), oy = int(int_scale * n.y), oz = int(int_scale * n.z);
  float px = __int_as_float(__float_as_int(p.x) + (p.x < 0 ? -ox : ox));
  float py = __int_as_float(__float_as_int(p.y) + (p.y < 0 ? -oy : oy));
  float pz = __int_as_float(__float_as_int(p.z) + (p.z < 0 ? -oz : oz));
  return V3(pabs(p.x) < origin ? p.x + n.x * float_scale : px,
            pabs(p.y) < origin ? p.y + n.y * float_scale : py,
            pabs(p.z) < origin ? p.z + n.z * float_scale : pz);
}
__device__ __forceinline__ Ray spawn_ray(V3 p, V3 n, V3 wo, float tmax = kFloatMax) {
  Ray r;
  r.d = wo;
  r.o = offset_ray_origin(p, face_same_hemisphere(n, r.d));
  r.tmin = 0.0f;
  r.tmax = tmax * (1.0f - 1e-3f);
  return r;
}

// ---- warps ------------------------------------------------------------------------------------------------
__device__ __forceinline__ V2 disk_concentric(V2 u) {
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
__device__ __forceinline__ V3 cosine_weighted_hemisphere(V2 u) {
  V2 d = disk_concentric(u);
  float z = sqrtf(pmax(1.0f - d.x * d.x - d.y * d.y, 0.0f));
  return V3(d.x, d.y, z);
}
__device__ __forceinline__ V3 uniform_sphere(V2 u) {
  float phi = u.x * kPi * 2;
  float cos_theta = 1 - 2 * u.y;
  float sin_theta = sqrtf(1.0f - sqr(cos_theta));
  return V3(sin_theta * cosf(phi), sin_theta * sinf(phi), cos_theta);
}
__device__ __forceinline__ float balance_heuristic(float pF, float pG) { return pF / (pF + pG); }

}  // namespace pb2
