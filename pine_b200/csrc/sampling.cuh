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

// ---- samplers: state of pine::Sampler's active alternative for one (pixel, sample) ------------------------
// GENERIC = false compiles the BlueSobolSampler draws only (the sampler of every named configuration: its
// shade kernels carry no dispatch); GENERIC = true dispatches on DScene::sampler_kind.
template <bool GENERIC>
struct PixelSamplerT {
  int px, py, index, dim;
};
using PixelSampler = PixelSamplerT<true>;
using BlueOnlySampler = PixelSamplerT<false>;

// BlueSobolSampler (sampler.h:166-201, bluenoise_<N>spp.cpp:14-34)
template <bool G>
__device__ __forceinline__ float blue_dim(const DScene& sc, const PixelSamplerT<G>& s, int dim) {
  int pi = s.px & 127, pj = s.py & 127;
  int si = s.index & 255, sd = dim & 255;
  int tile = (pi + pj * 128) * 8;
  int ranked = si ^ int(__ldg(sc.rank + ((sd + tile) % (128 * 128 * 8))));
  int value = int(__ldg(sc.sobol + sd + ranked * 256));
  value ^= int(__ldg(sc.scramble + (sd % 8) + tile));
  return (0.5f + value) / 256.0f;
}

// SobolSampler (sampler.h:83-164, sampler.cpp:82-113, lowdiscrepancy.h:73-80): all integer until the
// final scale, so the draws are the reference's bit for bit.
__host__ __device__ __forceinline__ uint64_t mix_bits(uint64_t v) {  // rng.h:81-88
  v ^= (v >> 31); v *= 0x7fb5d329728ea185ull;
  v ^= (v >> 27); v *= 0x81dadef4bc2dd44dull;
  v ^= (v >> 33);
  return v;
}
__host__ __device__ __forceinline__ uint64_t murmur_int(int v) {  // hash(int): 4 tail bytes, seed 0 (rng.h:9-65)
  const uint64_t m = 0xc6a4a7935bd1e995ull;
  const int r = 47;
  uint64_t h = 0 ^ (4ull * m);
  h ^= uint64_t(uint32_t(v));
  h *= m;
  h ^= h >> r; h *= m; h ^= h >> r;
  return h;
}
__host__ __device__ __forceinline__ uint64_t spread_bits_64x2(uint64_t x) {  // left_shift_64x2 (vecmath.h)
  x &= 0xffffffffull;
  x = (x ^ (x << 16)) & 0x0000ffff0000ffffull;
  x = (x ^ (x << 8)) & 0x00ff00ff00ff00ffull;
  x = (x ^ (x << 4)) & 0x0f0f0f0f0f0f0f0full;
  x = (x ^ (x << 2)) & 0x3333333333333333ull;
  x = (x ^ (x << 1)) & 0x5555555555555555ull;
  return x;
}
__host__ __device__ __forceinline__ uint32_t reverse_bits32(uint32_t x) {
#ifdef __CUDA_ARCH__
  return __brev(x);
#else
  x = (x & 0x55555555u) << 1 | (x & 0xaaaaaaaau) >> 1;
  x = (x & 0x33333333u) << 2 | (x & 0xccccccccu) >> 2;
  x = (x & 0x0f0f0f0fu) << 4 | (x & 0xf0f0f0f0u) >> 4;
  x = (x & 0x00ff00ffu) << 8 | (x & 0xff00ff00u) >> 8;
  return (x << 16) | (x >> 16);
#endif
}
__host__ __device__ __forceinline__ uint32_t fast_owen(uint32_t v, uint32_t seed) {  // sampler.h:97-110
  v = reverse_bits32(v);
  v ^= v * 0x3d20adeau;
  v += seed;
  v *= (seed >> 16) | 1u;
  v ^= v * 0x05526c56u;
  v ^= v * 0x53a22864u;
  return reverse_bits32(v);
}
// SobolMatrices32 restricted to the two dimensions the sampler uses (sobolmatrices.cpp:40-58): dimension
// 0 is the bit reversal of the low 32 bits (rows >= 32 are zero), dimension 1 the Pascal-triangle
// matrix v[i] = v[i-1] ^ (v[i-1] >> 1) from 0x80000000 (which repeats with period 32, as the table does).
__host__ __device__ __forceinline__ uint32_t sobol_dim0(uint64_t a) { return reverse_bits32(uint32_t(a)); }
__host__ __device__ __forceinline__ uint32_t sobol_dim1(uint64_t a) {
  uint32_t v = 0, col = 0x80000000u;
  for (; a != 0; a >>= 1, col ^= col >> 1)
    if (a & 1) v ^= col;
  return v;
}
__host__ __device__ __forceinline__ uint64_t sobol_sample_index(uint64_t sobol_index, int dimension, int log2_spp,
                                                              int nbase4_digits) {  // sampler.cpp:87-113
  // the 24 permutations of {0,1,2,3} in the reference's order, 2 bits per image, digit 0 lowest
  const uint8_t perms[24] = {0xE4, 0xB4, 0xD8, 0x78, 0x6C, 0x9C, 0xE1, 0xB1, 0xC9, 0x39, 0x2D, 0x8D,
                             0xC6, 0x36, 0xD2, 0x72, 0x4E, 0x1E, 0x27, 0x87, 0x1B, 0x4B, 0x63, 0x93};
  uint64_t si = 0;
  const bool only_power_of_2 = log2_spp & 1;
  const int last_digit = only_power_of_2 ? 1 : 0;
  for (int i = nbase4_digits - 1; i >= last_digit; --i) {
    const int digit_shift = 2 * i - (only_power_of_2 ? 1 : 0);
    int digit = int(sobol_index >> digit_shift) & 3;
    const uint64_t higher_digits = sobol_index >> (digit_shift + 2);
    const int p = int((mix_bits(higher_digits ^ uint64_t(0x55555555u * uint32_t(dimension))) >> 24) % 24);
    digit = (perms[p] >> (2 * digit)) & 3;
    si |= uint64_t(digit) << digit_shift;
  }
  if (only_power_of_2) {
    const int digit = int(sobol_index & 1);
    si |= uint64_t(digit ^ int(mix_bits((sobol_index >> 1) ^ uint64_t(0x55555555u * uint32_t(dimension))) & 1));
  }
  return si;
}
__host__ __device__ __forceinline__ float sobol_to_float(uint32_t v) { return pmin(float(v) * 0x1p-32f, kOneMinusEpsilon); }
__host__ __device__ __forceinline__ uint64_t sobol_index_of(const PixelSampler& s, int log2_spp) {  // sampler.h:135-142
  const uint64_t morton = (spread_bits_64x2(uint32_t(s.py)) << 1) | spread_bits_64x2(uint32_t(s.px));
  // start_pixel(p, 0) then `sobol_index++` per sample (path.cpp:32-34): a sum, which carries into the
  // Morton bits when spp is not a power of two, exactly as the reference's does
  return (morton << log2_spp) + uint64_t(int64_t(s.index));
}

// SobolSampler::get1d / get2d (sampler.h:143-160): the dimension advances before the Owen seed is hashed
__host__ __device__ __forceinline__ float sobol_get1d(PixelSampler& s, int log2_spp, int nbase4_digits) {
  const uint64_t si = sobol_sample_index(sobol_index_of(s, log2_spp), s.dim, log2_spp, nbase4_digits);
  s.dim += 1;
  const uint64_t u = murmur_int(s.dim);
  return sobol_to_float(fast_owen(sobol_dim0(si), uint32_t(u)));
}
__host__ __device__ __forceinline__ V2 sobol_get2d(PixelSampler& s, int log2_spp, int nbase4_digits) {
  const uint64_t si = sobol_sample_index(sobol_index_of(s, log2_spp), s.dim, log2_spp, nbase4_digits);
  s.dim += 2;
  const uint64_t u = murmur_int(s.dim);
  return V2(sobol_to_float(fast_owen(sobol_dim0(si), uint32_t(u))),
            sobol_to_float(fast_owen(sobol_dim1(si), uint32_t(u >> 32))));
}

// UniformSampler: a draw per (pixel, sample, dimension) from the reference's generator
__host__ __device__ __forceinline__ float uniform_dim(const PixelSampler& s, int seed, int dim) {
  uint64_t key = murmur_pixel(s.px, s.py, s.index) ^ mix_bits(uint64_t(uint32_t(dim)) * 0x9E3779B97f4A7C15ull + uint64_t(uint32_t(seed)) + 1ull);
  Rng g = rng_seed(key);
  return rng_nextf(g);
}

// HaltonSampler (sampler.h:40-81, sampler.cpp:20-81, lowdiscrepancy.h:26-51): the pixel's first index comes
// from the inverse radical inverses of (x mod 128, y mod 128) in bases 2 and 3 combined by the Chinese
// remainder theorem, samples are `stride` = 128 * 243 apart, dimension d >= 2 is the radical inverse in the
// d-th prime with a fixed random digit permutation (tables made on the host, halton_tables below).
struct HaltonTables {
  const int32_t* primes;   // Primes[1000] (primes.cpp)
  const int32_t* sums;     // PrimeSums[1000]: offset of each base's permutation
  const uint16_t* perms;   // compute_radical_inverse_permutations(RNG()) (lowdiscrepancy.cpp:5-17)
};
constexpr int kHaltonPrimes = 1000;           // PrimeTablesize (primes.h:7)
constexpr int kHaltonScale2 = 128, kHaltonScale3 = 243, kHaltonExp2 = 7, kHaltonExp3 = 5;
constexpr int kHaltonStride = kHaltonScale2 * kHaltonScale3;
__host__ __device__ __forceinline__ uint64_t inverse_radical_inverse(uint64_t inverse, int base, int n_digits) {
  uint64_t index = 0;
  for (int i = 0; i < n_digits; i++) {
    const uint64_t digit = inverse % uint64_t(base);
    inverse /= uint64_t(base);
    index = index * uint64_t(base) + digit;
  }
  return index;
}
__host__ __device__ __forceinline__ uint64_t halton_index_of(const PixelSampler& s, int inv0, int inv1) {  // sampler.cpp:67-81
  const int pmx = ((s.px % 128) + 128) % 128, pmy = ((s.py % 128) + 128) % 128;  // psl::mod
  int64_t h = 0;
  h += int64_t(inverse_radical_inverse(uint64_t(pmx), 2, kHaltonExp2) * uint64_t(kHaltonScale3) * uint64_t(inv1));
  h += int64_t(inverse_radical_inverse(uint64_t(pmy), 3, kHaltonExp3) * uint64_t(kHaltonScale2) * uint64_t(inv0));
  h %= kHaltonStride;
  h += int64_t(s.index) * int64_t(kHaltonStride);
  return uint64_t(h);
}
__host__ __device__ __forceinline__ float halton_dim(const HaltonTables& t, uint64_t a, int dim) {  // lowdiscrepancy.h:26-40
  const int base = t.primes[dim];
  const uint16_t* perm = t.perms + t.sums[dim];
  const float inv_base = 1.0f / float(base);
  float inv_base_n = 1.0f;
  uint64_t reversed = 0;
  while (a) {
    const uint64_t next = a / uint64_t(base);
    const uint64_t digit = a - next * uint64_t(base);
    reversed = reversed * uint64_t(base) + perm[digit];
    inv_base_n *= inv_base;
    a = next;
  }
  const float series = float(int(perm[0])) / (float(base) + 1.0f);
  return pmin((float(reversed) + series) * inv_base_n, kOneMinusEpsilon);
}
template <class S>
__host__ __device__ __forceinline__ float halton_get1d(const HaltonTables& t, S& s, int inv0, int inv1) {
  if (s.dim < 2 || s.dim >= kHaltonPrimes) s.dim = 2;  // start_pixel / start_next_sample set 2 (sampler.h:49-52)
  return halton_dim(t, halton_index_of(PixelSampler{s.px, s.py, s.index, 0}, inv0, inv1), s.dim++);
}
template <class S>
__host__ __device__ __forceinline__ V2 halton_get2d(const HaltonTables& t, S& s, int inv0, int inv1) {
  if (s.dim < 2 || s.dim + 1 >= kHaltonPrimes) s.dim = 2;
  const uint64_t a = halton_index_of(PixelSampler{s.px, s.py, s.index, 0}, inv0, inv1);
  const int d = s.dim;
  s.dim += 2;
  return V2(halton_dim(t, a, d), halton_dim(t, a, d + 1));
}

template <bool G>
__device__ __forceinline__ float smp_get1d(const DScene& sc, PixelSamplerT<G>& s) {
  if constexpr (G) {
    if (sc.sampler_kind == PB2_SAMPLER_SOBOL) return sobol_get1d(s, sc.log2_spp, sc.nbase4_digits);
    if (sc.sampler_kind == PB2_SAMPLER_UNIFORM) return uniform_dim(s, sc.sampler_seed, s.dim++);
    if (sc.sampler_kind == PB2_SAMPLER_HALTON)
      return halton_get1d(HaltonTables{sc.halton_primes, sc.halton_sums, sc.halton_perms}, s, sc.halton_inv0, sc.halton_inv1);
  }
  if (s.dim >= 256) s.dim = 2;
  return blue_dim(sc, s, s.dim++);
}
template <bool G>
__device__ __forceinline__ V2 smp_get2d(const DScene& sc, PixelSamplerT<G>& s) {
  if constexpr (G) {
    if (sc.sampler_kind == PB2_SAMPLER_SOBOL) return sobol_get2d(s, sc.log2_spp, sc.nbase4_digits);
    if (sc.sampler_kind == PB2_SAMPLER_UNIFORM) {
      const int d = s.dim;
      s.dim += 2;
      return V2(uniform_dim(s, sc.sampler_seed, d), uniform_dim(s, sc.sampler_seed, d + 1));
    }
    if (sc.sampler_kind == PB2_SAMPLER_HALTON)
      return halton_get2d(HaltonTables{sc.halton_primes, sc.halton_sums, sc.halton_perms}, s, sc.halton_inv0, sc.halton_inv1);
  }
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
