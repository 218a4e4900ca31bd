// oracle/oracle_port.cpp -- TEST INFRASTRUCTURE ONLY (never linked into, imported or called by the
// product; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it).
//
// Plain C++ restatement of the arithmetic on Pine's PathIntegrator traversal path, written from the
// reference sources cited per function (paths relative to /root/reference/src). It needs no
// reference tree to compile. Scope: per-pixel RNG + hash, BlueSobol sampler indexing and dimension
// bookkeeping, ThinLenCamera::gen_ray, spawn_ray, the shapes' intersect()/hit(), the binned-SAH
// builder and the two-level reference-order closest-hit / any-hit walks -- i.e. everything that
// must be BIT-exact. It also counts nodes visited and primitives tested per ray, which is what
// SURVEY.md 8(d)'s algorithmic bytes-per-ray figure is made of (tests/tools/b_ray.py).
// Shading (BSDFs, light sampling, the estimator) is not restated here: its oracle is the compiled
// reference itself (oracle/_ref), against which this port is pinned too (tests/test_oracle_port.py).
// PINNING: the reference ships no golden vectors for this path, so every function
// here is checked against oracle/_ref outputs; tests/golden/ holds vectors generated from it.
//
// Float semantics: compile with -O2 -ffp-contract=off, no -march, no -ffast-math.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

namespace {

struct V3 {
  float x = 0, y = 0, z = 0;
  V3() = default;
  V3(float a, float b, float c) : x(a), y(b), z(c) {}
  explicit V3(float a) : x(a), y(a), z(a) {}
  float& operator[](int i) { return (&x)[i]; }
  float operator[](int i) const { return (&x)[i]; }
};
// pine/core/vecmath.h: component-wise operators, dot = x*x' + y*y' + z*z' left to right
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length(V3 v) { return std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z); }
inline V3 normalize(V3 v) {  // vecmath.h: returns v unchanged when its length is 0
  float l = length(v);
  return l == 0 ? v : v / l;
}
inline float fmin_(float a, float b) { return a < b ? a : b; }  // psl::min (psl/math.h:19-26)
inline float fmax_(float a, float b) { return a > b ? a : b; }  // psl::max
inline float sqr(float x) { return x * x; }
inline float safe_rcp(float v) { return v == 0.0f ? 1e+20f : 1.0f / v; }  // vecmath.h:1062-1068
constexpr float kFloatMax = 3.402823466e+38f;
constexpr float kOneMinusEps = 0x1.fffffep-1f;

struct Ray {
  V3 o, d;
  float tmin = 0, tmax = kFloatMax;
  V3 at(float t) const { return o + t * d; }
};

// ---- pine/core/rng.h:9-45 murmur_hash64A over {p.x, p.y, sample_index} (12 bytes, seed 0) --------
uint64_t murmur64a(const unsigned char* key, size_t len, uint64_t seed) {
  const uint64_t m = 0xc6a4a7935bd1e995ull;
  const int r = 47;
  uint64_t h = seed ^ (len * m);
  const unsigned char* end = key + 8 * (len / 8);
  while (key != end) {
    uint64_t k;
    std::memcpy(&k, key, 8);
    key += 8;
    k *= m; k ^= k >> r; k *= m;
    h ^= k; h *= m;
  }
  switch (len & 7) {
    case 7: h ^= uint64_t(key[6]) << 48; [[fallthrough]];
    case 6: h ^= uint64_t(key[5]) << 40; [[fallthrough]];
    case 5: h ^= uint64_t(key[4]) << 32; [[fallthrough]];
    case 4: h ^= uint64_t(key[3]) << 24; [[fallthrough]];
    case 3: h ^= uint64_t(key[2]) << 16; [[fallthrough]];
    case 2: h ^= uint64_t(key[1]) << 8; [[fallthrough]];
    case 1: h ^= uint64_t(key[0]); h *= m;
  }
  h ^= h >> r; h *= m; h ^= h >> r;
  return h;
}
uint64_t hash_pixel(int px, int py, int sample_index) {  // rng.h:60-65 hash(vec2i, int)
  int32_t buf[4] = {px, py, sample_index, 0};
  return murmur64a(reinterpret_cast<const unsigned char*>(buf), 12, 0);
}
struct Rng {  // rng.h:97-144
  uint64_t s[2];
  explicit Rng(uint64_t seed) {
    auto mix = [&]() {  // split_mix_64, rng.h:74-79
      uint64_t r = seed += 0x9E3779B97f4A7C15ull;
      r = (r ^ (r >> 30)) * 0xBF58476D1CE4E5B9ull;
      r = (r ^ (r >> 27)) * 0x94D049BB133111EBull;
      return r ^ (r >> 31);
    };
    s[0] = mix();
    s[1] = mix();
  }
  uint64_t next64() {
    const uint64_t s0 = s[0];
    uint64_t s1 = s[1];
    const uint64_t result = s0 + s1;
    s1 ^= s0;
    s[0] = ((s0 << 24) | (s0 >> 40)) ^ s1 ^ (s1 << 16);
    s[1] = (s1 << 37) | (s1 >> 27);
    return result;
  }
  float nextf() {
    uint64_t u = next64();
    return fmin_(float(uint32_t(u ^ (u >> 32))) * 0x1p-32f, kOneMinusEps);
  }
};

// ---- contrib/bluesobol/bluenoise_16spp.cpp:14-34 on u8 tables --------------------------------------
float bluenoise(const uint8_t* sobol, const uint8_t* scramble, const uint8_t* rank, int px, int py,
                int index, int dim) {
  px &= 127; py &= 127; index &= 255; dim &= 255;
  int ranked = index ^ rank[(dim + (px + py * 128) * 8) % (128 * 128 * 8)];
  int value = sobol[dim + ranked * 256];
  value ^= scramble[(dim % 8) + (px + py * 128) * 8];
  return (0.5f + value) / 256.0f;
}

// ---- shapes (records = the members of the reference structs, same layout as include/pine_b200.h) ----
struct Rect { float position[3], ex[3], ey[3], n[3], lx, ly, rx[3], ry[3]; };
struct Aabb { float lower[3], upper[3]; };
struct Obb { Aabb base; float m[16], m_inv[16]; };
struct Disk { float position[3], n[3], u[3], v[3], r; };
struct Cone { Disk bottom; float p[3], n[3], r, h, A, A2, S; };
struct Sphere { float c[3], r; };
struct Plane { float position[3], n[3], u[3], v[3]; };
struct Cylinder { float p0[3], p1[3], n[3], r; Disk bottom, top; };
struct Tri { float v0[3], v1[3], v2[3], n[3]; };
inline V3 ld(const float* p) { return {p[0], p[1], p[2]}; }

// geometry.cpp:280-304
bool rect_test(const Rect& s, const Ray& ray, float& t_out) {
  V3 n = ld(s.n), pos = ld(s.position);
  float denom = dot(ray.d, n);
  if (denom == 0.0f) return false;
  float t = dot(pos - ray.o, n) / denom;
  if (t <= ray.tmin || t >= ray.tmax) return false;
  V3 p = ray.at(t) - pos;
  float u = dot(p, ld(s.rx));
  if (u < -0.5f || u > 0.5f) return false;
  float v = dot(p, ld(s.ry));
  if (v < -0.5f || v > 0.5f) return false;
  t_out = t;
  return true;
}
// geometry.cpp:532-565; strict = intersect(), non-strict = hit()
bool tri_test(const Ray& ray, V3 v0, V3 v1, V3 v2, bool any, float& t_out) {
  V3 E1 = v1 - v0, E2 = v2 - v0, T = ray.o - v0;
  V3 P = cross(ray.d, E2), Q = cross(T, E1);
  float D = dot(P, E1);
  if (D == 0.0f) return false;
  float t = dot(Q, E2) / D;
  if (any ? (t < ray.tmin || t > ray.tmax) : (t <= ray.tmin || t >= ray.tmax)) return false;
  float u = dot(P, T) / D;
  if (u < 0.0f || u > 1.0f) return false;
  float v = dot(Q, ray.d) / D;
  if (v < 0.0f || v > 1.0f) return false;
  if (any ? !(u + v < 1.0f) : (u + v > 1.0f)) return false;
  t_out = t;
  return true;
}
// geometry.cpp:20-29
bool quadratic(float a, float b, float c, float tmin, float& tmax) {
  float d = b * b - 4 * a * c;
  if (d <= 0.0f) return false;
  d = std::sqrt(d);
  float t = (-b - d) / (2 * a);
  if (t < tmin) t += d / a;
  if (t < tmin || t > tmax) return false;
  tmax = t;
  return true;
}
// geometry.cpp:415-446 (no base cap)
bool cone_test(const Cone& s, const Ray& ray, bool any, float& t_out) {
  V3 o = ray.o - ld(s.p), d = ray.d, n = ld(s.n);
  float a = -s.A2 * sqr(dot(d, n)) + dot(d, d);
  float b = 2 * (-s.A2 * dot(o, n) * dot(d, n) + dot(o, d));
  float c = -s.A2 * sqr(dot(o, n)) + dot(o, o);
  float tmax = ray.tmax;
  if (!quadratic(a, b, c, ray.tmin, tmax)) return false;
  float side = dot(o + tmax * d, n);
  if (any ? !(side <= 0) : !(side < 0)) return false;
  t_out = tmax;
  return true;
}
// geometry.cpp:35-45: hit() is strict at tmin, intersect() non-strict at both ends (and lets a NaN t through)
bool plane_test(const Plane& s, const Ray& ray, bool any, float& t_out) {
  V3 n = ld(s.n);
  float t = (dot(ld(s.position), n) - dot(ray.o, n)) / dot(ray.d, n);
  if (any) {
    if (t <= ray.tmin) return false;
    if (!(t < ray.tmax)) return false;
  } else {
    if (t < ray.tmin || t > ray.tmax) return false;
  }
  t_out = t;
  return true;
}
// geometry.cpp:466-517: the open tube; the far root is not re-checked against tmin
bool cylinder_test(const Cylinder& s, const Ray& ray, float& t_out) {
  V3 p0 = ld(s.p0), n = ld(s.n);
  V3 m = ray.o - p0;
  V3 v = ray.d - dot(ray.d, n) * n;
  V3 w = m - dot(m, n) * n;
  float a = dot(v, v);
  float b = 2 * dot(v, w);
  float c = dot(w, w) - s.r * s.r;
  float discriminant = b * b - 4 * a * c;
  if (discriminant < 0) return false;
  float sqrt_disc = std::sqrt(discriminant);
  float t = (-b - sqrt_disc) / (2 * a);
  if (t < ray.tmin) t = (-b + sqrt_disc) / (2 * a);
  if (t > ray.tmax) return false;
  V3 hit_point = ray.at(t);
  V3 projection = p0 + dot(hit_point - p0, n) * n;
  if (dot(projection - p0, n) < 0 || dot(projection - ld(s.p1), n) > 0) return false;
  t_out = t;
  return true;
}
// geometry.cpp:73-92
bool sphere_test(const Sphere& s, const Ray& ray, bool any, float& t_out) {
  V3 ro_p = ray.o - ld(s.c);
  float b = dot(ro_p, ray.d);
  float c = dot(ro_p, ro_p) - s.r * s.r;
  float d = b * b - c;
  float t;
  if (d <= 0.0f) {
    t = -1.0f;
  } else {
    d = std::sqrt(d);
    t = -b - d;
    if (t < ray.tmin) t = -b + d;
  }
  if (any) return t > ray.tmin && t < ray.tmax;
  if (t < ray.tmin || t > ray.tmax) return false;
  t_out = t;
  return true;
}
// geometry.cpp:129-150
bool disk_test(const Disk& s, const Ray& ray, bool any, float& t_out) {
  V3 n = ld(s.n), pos = ld(s.position);
  float denom = dot(ray.d, n);
  if (denom == 0.0f) return false;
  float t = (dot(pos, n) - dot(ray.o, n)) / denom;
  if (any) { if (t < ray.tmin) return false; if (t >= ray.tmax) return false; }
  else if (t < ray.tmin || t > ray.tmax) return false;
  V3 p = ray.at(t) - pos;
  if (dot(p, p) > sqr(s.r)) return false;
  t_out = t;
  return true;
}
// bbox.cpp:68-113
bool aabb_slabs(const Aabb& s, V3 o, V3 d, float& tmin, float& tmax) {
  for (int i = 0; i < 3; i++) {
    if (std::fabs(d[i]) < 1e-6f) {
      if (o[i] < s.lower[i] || o[i] > s.upper[i]) return false;
      continue;
    }
    float inv_d = 1.0f / d[i];
    float t_near = (s.lower[i] - o[i]) * inv_d;
    float t_far = (s.upper[i] - o[i]) * inv_d;
    if (inv_d < 0.0f) std::swap(t_far, t_near);
    tmin = fmax_(t_near, tmin);
    tmax = fmin_(t_far, tmax);
    if (tmin > tmax) return false;
  }
  return true;
}
bool aabb_test(const Aabb& s, const Ray& ray, bool any, float& t_out) {
  float tmin = ray.tmin, tmax = ray.tmax;
  if (any && tmin > tmax) return false;
  if (!aabb_slabs(s, ray.o, ray.d, tmin, tmax)) return false;
  t_out = tmin > ray.tmin ? tmin : tmax;
  return true;
}
// Matrix4 * vec3 (vecmath.h:704-707): m.x*v.x + m.y*v.y + m.z*v.z + m.w, column-major storage
V3 mul_point(const float* m, V3 v) {
  V3 r;
  for (int k = 0; k < 3; k++) r[k] = m[0 + k] * v.x + m[4 + k] * v.y + m[8 + k] * v.z + m[12 + k];
  return r;
}
V3 mul_dir(const float* m, V3 v) {
  V3 r;
  for (int k = 0; k < 3; k++) r[k] = m[0 + k] * v.x + m[4 + k] * v.y + m[8 + k] * v.z;
  return r;
}
// bbox.cpp:143-175: world-space tmin/tmax fed to the local-space slab test, then distances
bool obb_test(const Obb& s, const Ray& ray, bool any, float& t_out) {
  V3 o = mul_point(s.m_inv, ray.o);
  V3 d = normalize(mul_dir(s.m_inv, ray.d));
  float tmin = ray.tmin, tmax = ray.tmax;
  if (any) {
    if (tmin > tmax) return false;
    return aabb_slabs(s.base, o, d, tmin, tmax);
  }
  if (!aabb_slabs(s.base, o, d, tmin, tmax)) return false;
  V3 ps = o + tmin * d, pe = o + tmax * d;
  tmin = length(mul_point(s.m, ps) - ray.o);
  tmax = length(mul_point(s.m, pe) - ray.o);
  t_out = tmin > ray.tmin ? tmin : tmax;
  return true;
}

}  // namespace

// ===================================================================================================
// BVH: builder (impl/accel/bvh.cpp:30-147) and reference-order walks (bvh.cpp:321-451, 497-548)
// ===================================================================================================
namespace {

struct Box {
  V3 lower{kFloatMax, kFloatMax, kFloatMax}, upper{-kFloatMax, -kFloatMax, -kFloatMax};
  void extend(V3 p) {
    for (int k = 0; k < 3; k++) { lower[k] = fmin_(lower[k], p[k]); upper[k] = fmax_(upper[k], p[k]); }
  }
  void extend(const Box& b) {
    for (int k = 0; k < 3; k++) { lower[k] = fmin_(lower[k], b.lower[k]); upper[k] = fmax_(upper[k], b.upper[k]); }
  }
  float centroid(int d) const { return (lower[d] + upper[d]) / 2; }
  V3 centroid() const { return (lower + upper) / 2; }
  bool degenerated(int d) const { return upper[d] <= lower[d]; }
  float relative_position(float p, int d) const {  // bbox.cpp: divides only when the extent is > 0
    float o = p - lower[d];
    float e = upper[d] - lower[d];
    return e > 0 ? o / e : o;
  }
  float surface_area() const {
    V3 d = upper - lower;
    return 2 * (d.x * d.y + d.x * d.z + d.y * d.z);
  }
};
struct Prim { Box box; int index; };
struct Node {
  Box box[2];
  int child[2] = {-1, -1};
  std::vector<int> prims;  // non-empty <=> leaf
};
struct Tree {
  std::vector<Node> nodes;
  int root = -1;

  int build_range(Prim* begin, Prim* end, const Box& bounds) {
    Node node;
    const int n = int(end - begin);
    auto make_leaf = [&]() {
      for (int i = 0; i < n; i++) node.prims.push_back(begin[i].index);
      for (Prim* p = begin; p != end; p++) node.box[0].extend(p->box);
      node.box[1] = node.box[0];
      nodes.push_back(node);
      return int(nodes.size()) - 1;
    };
    if (n == 1) return make_leaf();
    Box cb;
    for (int i = 0; i < n; i++) cb.extend(begin[i].box.centroid());
    const float area = bounds.surface_area();
    const int NB = 16;
    float min_cost = kFloatMax;
    int best_axis = -1, split_bucket = -1;
    for (int axis = 0; axis < 3; axis++) {
      if (cb.degenerated(axis)) continue;
      int count[NB] = {};
      Box bbox[NB];
      for (int i = 0; i < n; i++) {
        int b = int(NB * cb.relative_position(begin[i].box.centroid(axis), axis));
        if (b > NB - 1) b = NB - 1;
        count[b]++;
        bbox[b].extend(begin[i].box);
      }
      float cost[NB - 1] = {};
      Box fwd;
      int cf = 0;
      for (int i = 0; i < NB - 1; i++) { fwd.extend(bbox[i]); cf += count[i]; cost[i] += cf * fwd.surface_area(); }
      Box bwd;
      int cbk = 0;
      for (int i = NB - 1; i >= 1; i--) { bwd.extend(bbox[i]); cbk += count[i]; cost[i - 1] += cbk * bwd.surface_area(); }
      for (int i = 0; i < NB - 1; i++) cost[i] = 1.0f + cost[i] / area;
      float axis_min = kFloatMax;
      int axis_bucket = -1;
      for (int i = 0; i < NB - 1; i++)
        if (cost[i] < axis_min) { axis_min = cost[i]; axis_bucket = i; }
      if (axis_min < min_cost) { min_cost = axis_min; best_axis = axis; split_bucket = axis_bucket; }
    }
    if (min_cost > float(n)) return make_leaf();
    // psl::partition (psl/algorithm.h): forward scan, matching elements swapped to the front
    Prim* mid = begin;
    for (Prim* it = begin; it != end; ++it) {
      int b = int(NB * cb.relative_position(it->box.centroid(best_axis), best_axis));
      if (b == NB) b = NB - 1;
      if (b <= split_bucket) { std::swap(*mid, *it); ++mid; }
    }
    for (Prim* p = begin; p != mid; p++) node.box[0].extend(p->box);
    for (Prim* p = mid; p != end; p++) node.box[1].extend(p->box);
    Box b0 = node.box[0], b1 = node.box[1];
    int c0 = build_range(begin, mid, b0);
    int c1 = build_range(mid, end, b1);
    node.child[0] = c0;
    node.child[1] = c1;
    nodes.push_back(node);
    return int(nodes.size()) - 1;
  }
  void build(std::vector<Prim> prims) {
    nodes.clear();
    root = -1;
    if (prims.empty()) return;
    Box all;
    for (auto& p : prims) all.extend(p.box);
    build_range(prims.data(), prims.data() + prims.size(), all);
    root = int(nodes.size()) - 1;
  }
  Box bounds() const {
    Box b = nodes[root].box[0];
    b.extend(nodes[root].box[1]);
    return b;
  }
};

struct Octant {  // bbox.h:18-27
  int o3[3];
  V3 inv, ood;
  explicit Octant(const Ray& r) {
    for (int k = 0; k < 3; k++) { o3[k] = (r.d[k] < 0 ? 1 : 0) * 3; inv[k] = safe_rcp(r.d[k]); }
    ood = r.o * inv;
  }
};
// bbox.h:59-72; psl::max(a,b,c,d) = max(a, max(b, max(c, d))) (psl/math.h variadic)
bool slab(const Box& b, const Octant& q, float tmin, float& tmax) {
  const float* p = &b.lower.x;
  float p6[6] = {b.lower.x, b.lower.y, b.lower.z, b.upper.x, b.upper.y, b.upper.z};
  (void)p;
  float tmin0 = p6[0 + q.o3[0]] * q.inv[0] - q.ood[0];
  float tmin1 = p6[1 + q.o3[1]] * q.inv[1] - q.ood[1];
  float tmin2 = p6[2 + q.o3[2]] * q.inv[2] - q.ood[2];
  float tmax0 = p6[3 - q.o3[0]] * q.inv[0] - q.ood[0];
  float tmax1 = p6[4 - q.o3[1]] * q.inv[1] - q.ood[1];
  float tmax2 = p6[5 - q.o3[2]] * q.inv[2] - q.ood[2];
  tmin = fmax_(tmin0, fmax_(tmin1, fmax_(tmin2, tmin)));
  tmax = fmin_(tmax0, fmin_(tmax1, fmin_(tmax2, tmax)));
  return tmin <= tmax;
}

struct Counters { int nodes = 0, prim_tests = 0; long prim_bytes = 0; };
// information content of one primitive test under the reference traversal -- 48 B per
// triangle (3 x vec3 + 12 B index), the shape struct's own members otherwise.
static inline int shape_bytes(int kind) {
  static const int b[8] = {24, 152, 16, 48, 64, 76, 52, 48};  // AABB OBB SPHERE TRIANGLE CONE RECT DISK (MESH)
  return b[kind & 7];
}
static long g_last_nodes = 0, g_last_prim_tests = 0, g_last_prim_bytes = 0;

// BVHImpl::Intersect (ANY=false) / BVHImpl::hit (ANY=true). `leaf(prim)` returns true on an accepted
// hit (closest: after shrinking ray.tmax).
template <bool ANY, class F>
bool walk(const Tree& t, const Ray& ray, const Octant& q, Counters& cnt, F&& leaf) {
  if (t.root < 0) return false;
  bool hit = false;
  const Node& rootn = t.nodes[t.root];
  if (!rootn.prims.empty()) {
    for (int idx : rootn.prims)
      if (leaf(idx)) { if (ANY) return true; hit = true; }
    return hit;
  }
  int stack[64], ptr = 0, next = t.root;
  while (true) {
    const Node& node = t.nodes[next];
    cnt.nodes++;
    int ci[2] = {-1, -1};
    float tt[2] = {ray.tmax, ray.tmax};
    for (int c = 0; c < 2; c++) {
      if (slab(node.box[c], q, ray.tmin, tt[c])) {
        const Node& ch = t.nodes[node.child[c]];
        if (ch.prims.empty()) ci[c] = node.child[c];
        else
          for (int idx : ch.prims)
            if (leaf(idx)) { if (ANY) return true; hit = true; }
      }
    }
    if (ci[0] != -1) {
      if (ci[1] != -1) {
        if (tt[0] > tt[1]) { stack[ptr++] = ci[0]; next = ci[1]; }
        else { stack[ptr++] = ci[1]; next = ci[0]; }
      } else next = ci[0];
    } else if (ci[1] != -1) next = ci[1];
    else { if (ptr == 0) break; next = stack[--ptr]; }
  }
  return hit;
}

enum Kind { K_AABB = 0, K_OBB, K_SPHERE, K_TRIANGLE, K_CONE, K_RECT, K_DISK, K_MESH, K_PLANE, K_CYLINDER };
struct Geom { int kind, index; };
struct MeshData { std::vector<V3> verts; std::vector<uint32_t> idx; };
struct Scene {
  std::vector<Geom> geoms;
  std::vector<Aabb> aabbs; std::vector<Obb> obbs; std::vector<Sphere> spheres; std::vector<Tri> tris;
  std::vector<Cone> cones; std::vector<Rect> rects; std::vector<Disk> disks; std::vector<MeshData> meshes;
  std::vector<Plane> planes; std::vector<Cylinder> cylinders;
  Tree top;
  std::vector<Tree> lbvh;
  std::vector<int> indices;  // BVH::indices: top primitive -> geometry
  bool built = false;

  Box shape_box(const Geom& g) const {
    Box b;
    switch (g.kind) {
      case K_AABB: { const Aabb& a = aabbs[g.index]; b.lower = ld(a.lower) - V3(1.1920928955078125e-07f); b.upper = ld(a.upper) + V3(1.1920928955078125e-07f); } break;
      case K_OBB: {
        const Obb& o = obbs[g.index];
        for (int i = 0; i < 8; i++) {
          V3 p = ld(o.base.lower);
          if (i % 2 >= 1) p[0] = o.base.upper[0];
          if (i % 4 >= 2) p[1] = o.base.upper[1];
          if (i % 8 >= 4) p[2] = o.base.upper[2];
          b.extend(mul_point(o.m, p));
        }
      } break;
      case K_SPHERE: b.lower = ld(spheres[g.index].c) - V3(spheres[g.index].r); b.upper = ld(spheres[g.index].c) + V3(spheres[g.index].r); break;
      case K_DISK: b.lower = ld(disks[g.index].position) - V3(disks[g.index].r); b.upper = ld(disks[g.index].position) + V3(disks[g.index].r); break;
      case K_CONE: {
        const Cone& c = cones[g.index];
        b.lower = ld(c.bottom.position) - V3(c.bottom.r); b.upper = ld(c.bottom.position) + V3(c.bottom.r);
        b.extend(ld(c.p));
      } break;
      case K_PLANE: b.lower = ld(planes[g.index].position) + V3(-100.0f); b.upper = ld(planes[g.index].position) + V3(100.0f); break;  // geometry.cpp:52
      case K_CYLINDER: {  // union_(bottom.get_aabb(), top.get_aabb()) (geometry.h:147): both disks sit at p0
        const Cylinder& c = cylinders[g.index];
        b.lower = ld(c.bottom.position) - V3(c.bottom.r); b.upper = ld(c.bottom.position) + V3(c.bottom.r);
        b.extend(ld(c.top.position) - V3(c.top.r)); b.extend(ld(c.top.position) + V3(c.top.r));
      } break;
      case K_TRIANGLE: b.extend(ld(tris[g.index].v0)); b.extend(ld(tris[g.index].v1)); b.extend(ld(tris[g.index].v2)); break;
      case K_RECT: {
        const Rect& r = rects[g.index];
        V3 p = ld(r.position), ex = ld(r.ex), ey = ld(r.ey);
        b.extend(p - ex * r.lx / 2 - ey * r.ly / 2);
        b.extend(p - ex * r.lx / 2 + ey * r.ly / 2);
        b.extend(p + ex * r.lx / 2 - ey * r.ly / 2);
        b.extend(p + ex * r.lx / 2 + ey * r.ly / 2);
      } break;
    }
    return b;
  }
  void build() {  // BVH::build, bvh.cpp:453-495
    lbvh.clear();
    indices.clear();
    for (size_t i = 0; i < geoms.size(); i++) {
      if (geoms[i].kind != K_MESH) continue;
      const MeshData& m = meshes[geoms[i].index];
      const size_t nt = m.idx.size() / 3;
      if (nt == 0) continue;
      std::vector<Prim> prims(nt);
      for (size_t t = 0; t < nt; t++) {
        prims[t].box.extend(m.verts[m.idx[3 * t]]);
        prims[t].box.extend(m.verts[m.idx[3 * t + 1]]);
        prims[t].box.extend(m.verts[m.idx[3 * t + 2]]);
        prims[t].index = int(t);
      }
      lbvh.emplace_back();
      lbvh.back().build(std::move(prims));
      indices.push_back(int(i));
    }
    std::vector<Prim> prims;
    for (size_t k = 0; k < lbvh.size(); k++) prims.push_back(Prim{lbvh[k].bounds(), int(prims.size())});
    for (size_t i = 0; i < geoms.size(); i++) {
      if (geoms[i].kind == K_MESH) continue;
      prims.push_back(Prim{shape_box(geoms[i]), int(prims.size())});
      indices.push_back(int(i));
    }
    top.build(std::move(prims));
    built = true;
  }
  bool shape_test(const Geom& g, const Ray& ray, bool any, float& t) const {
    switch (g.kind) {
      case K_AABB: return aabb_test(aabbs[g.index], ray, any, t);
      case K_OBB: return obb_test(obbs[g.index], ray, any, t);
      case K_SPHERE: return sphere_test(spheres[g.index], ray, any, t);
      case K_TRIANGLE: { const Tri& s = tris[g.index]; return tri_test(ray, ld(s.v0), ld(s.v1), ld(s.v2), any, t); }
      case K_CONE: return cone_test(cones[g.index], ray, any, t);
      case K_RECT: {
        if (!any) return rect_test(rects[g.index], ray, t);
        return rect_test(rects[g.index], ray, t);  // Rect::hit == Rect::intersect's test
      }
      case K_DISK: return disk_test(disks[g.index], ray, any, t);
      case K_PLANE: return plane_test(planes[g.index], ray, any, t);
      case K_CYLINDER: return cylinder_test(cylinders[g.index], ray, t);
    }
    return false;
  }
  // BVH::intersect (bvh.cpp:513-536)
  bool closest(Ray& ray, int& geom, int& prim, Counters& cnt) const {
    Octant q(ray);
    geom = prim = -1;
    return walk<false>(top, ray, q, cnt, [&](int top_id) {
      const Geom& g = geoms[indices[top_id]];
      if (top_id < int(lbvh.size())) {
        const MeshData& m = meshes[g.index];
        int p = -1;
        bool hit = walk<false>(lbvh[top_id], ray, q, cnt, [&](int tri) {
          cnt.prim_tests++; cnt.prim_bytes += 48;
          float t;
          if (!tri_test(ray, m.verts[m.idx[3 * tri]], m.verts[m.idx[3 * tri + 1]], m.verts[m.idx[3 * tri + 2]], false, t)) return false;
          ray.tmax = t;
          p = tri;
          return true;
        });
        if (hit) { geom = indices[top_id]; prim = p; }
        return hit;
      }
      cnt.prim_tests++; cnt.prim_bytes += shape_bytes(g.kind);
      float t;
      if (!shape_test(g, ray, false, t)) return false;
      ray.tmax = t;
      geom = indices[top_id];
      prim = 0;
      return true;
    });
  }
  // BVH::hit (bvh.cpp:497-511)
  bool any(const Ray& ray, Counters& cnt) const {
    Octant q(ray);
    return walk<true>(top, ray, q, cnt, [&](int top_id) {
      const Geom& g = geoms[indices[top_id]];
      if (top_id < int(lbvh.size())) {
        const MeshData& m = meshes[g.index];
        return walk<true>(lbvh[top_id], ray, q, cnt, [&](int tri) {
          cnt.prim_tests++; cnt.prim_bytes += 48;
          float t;
          return tri_test(ray, m.verts[m.idx[3 * tri]], m.verts[m.idx[3 * tri + 1]], m.verts[m.idx[3 * tri + 2]], true, t);
        });
      }
      cnt.prim_tests++; cnt.prim_bytes += shape_bytes(g.kind);
      float t;
      return shape_test(g, ray, true, t);
    });
  }
};

}  // namespace

// ===================================================================================================
// C ABI (ctypes: oracle/port.py)
// ===================================================================================================
extern "C" {

void* port_scene_new() { return new Scene(); }
void port_scene_free(void* h) { delete static_cast<Scene*>(h); }
// kind: pb2_shape_kind numbering (0 AABB .. 6 DISK, 8 PLANE, 9 CYLINDER); rec: the shape record (layout of include/pine_b200.h)
int port_add_shape(void* h, int kind, const void* rec) {
  Scene& S = *static_cast<Scene*>(h);
  int index = -1;
  switch (kind) {
    case K_AABB: S.aabbs.push_back(*static_cast<const Aabb*>(rec)); index = int(S.aabbs.size()) - 1; break;
    case K_OBB: S.obbs.push_back(*static_cast<const Obb*>(rec)); index = int(S.obbs.size()) - 1; break;
    case K_SPHERE: S.spheres.push_back(*static_cast<const Sphere*>(rec)); index = int(S.spheres.size()) - 1; break;
    case K_TRIANGLE: S.tris.push_back(*static_cast<const Tri*>(rec)); index = int(S.tris.size()) - 1; break;
    case K_CONE: S.cones.push_back(*static_cast<const Cone*>(rec)); index = int(S.cones.size()) - 1; break;
    case K_RECT: S.rects.push_back(*static_cast<const Rect*>(rec)); index = int(S.rects.size()) - 1; break;
    case K_DISK: S.disks.push_back(*static_cast<const Disk*>(rec)); index = int(S.disks.size()) - 1; break;
    case K_PLANE: S.planes.push_back(*static_cast<const Plane*>(rec)); index = int(S.planes.size()) - 1; break;
    case K_CYLINDER: S.cylinders.push_back(*static_cast<const Cylinder*>(rec)); index = int(S.cylinders.size()) - 1; break;
    default: return -1;
  }
  S.geoms.push_back(Geom{kind, index});
  S.built = false;
  return int(S.geoms.size()) - 1;
}
int port_add_mesh(void* h, const float* verts, int nv, const uint32_t* idx, int nt) {
  Scene& S = *static_cast<Scene*>(h);
  MeshData m;
  m.verts.resize(nv);
  for (int i = 0; i < nv; i++) m.verts[i] = V3(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
  m.idx.assign(idx, idx + size_t(nt) * 3);
  S.meshes.push_back(std::move(m));
  S.geoms.push_back(Geom{K_MESH, int(S.meshes.size()) - 1});
  S.built = false;
  return int(S.geoms.size()) - 1;
}
void port_build(void* h) { static_cast<Scene*>(h)->build(); }

// rays: n x {o, d, tmin, tmax}. Optional n_nodes / n_prim_tests: per-ray counts of binary nodes
// visited and primitive intersect() calls under the reference traversal.
void port_trace_closest(void* h, const float* rays, int64_t n, int32_t* geom, int32_t* prim, float* t,
                        int32_t* n_nodes, int32_t* n_prim_tests) {
  Scene& S = *static_cast<Scene*>(h);
  if (!S.built) S.build();
  g_last_nodes = g_last_prim_tests = g_last_prim_bytes = 0;
  for (int64_t i = 0; i < n; i++) {
    const float* r = rays + 8 * i;
    Ray ray;
    ray.o = V3(r[0], r[1], r[2]); ray.d = V3(r[3], r[4], r[5]); ray.tmin = r[6]; ray.tmax = r[7];
    Counters c;
    int g, p;
    bool hit = S.closest(ray, g, p, c);
    geom[i] = hit ? g : -1;
    prim[i] = hit ? p : -1;
    t[i] = ray.tmax;
    if (n_nodes) n_nodes[i] = c.nodes;
    if (n_prim_tests) n_prim_tests[i] = c.prim_tests;
    g_last_nodes += c.nodes; g_last_prim_tests += c.prim_tests; g_last_prim_bytes += c.prim_bytes;
  }
}
void port_trace_any(void* h, const float* rays, int64_t n, uint8_t* occluded, int32_t* n_nodes,
                    int32_t* n_prim_tests) {
  Scene& S = *static_cast<Scene*>(h);
  if (!S.built) S.build();
  g_last_nodes = g_last_prim_tests = g_last_prim_bytes = 0;
  for (int64_t i = 0; i < n; i++) {
    const float* r = rays + 8 * i;
    Ray ray;
    ray.o = V3(r[0], r[1], r[2]); ray.d = V3(r[3], r[4], r[5]); ray.tmin = r[6]; ray.tmax = r[7];
    Counters c;
    occluded[i] = S.any(ray, c) ? 1 : 0;
    if (n_nodes) n_nodes[i] = c.nodes;
    if (n_prim_tests) n_prim_tests[i] = c.prim_tests;
    g_last_nodes += c.nodes; g_last_prim_tests += c.prim_tests; g_last_prim_bytes += c.prim_bytes;
  }
}
// Totals of the last port_trace_* call: {binary nodes visited, primitive tests, primitive bytes}.
void port_last_totals(int64_t* out) { out[0] = g_last_nodes; out[1] = g_last_prim_tests; out[2] = g_last_prim_bytes; }
// Tree dump in the layout of oracle/ref_harness.cpp:ref_bvh_nodes (which = -1 top, k = mesh tree k).
static const Tree* pick(Scene& S, int which) {
  if (which < 0) return &S.top;
  return which < int(S.lbvh.size()) ? &S.lbvh[which] : nullptr;
}
int port_bvh_info(void* h, int which, int32_t* root, int32_t* n_leaf_prims) {
  Scene& S = *static_cast<Scene*>(h);
  if (!S.built) S.build();
  const Tree* t = pick(S, which);
  if (!t) return 0;
  *root = t->root;
  int np = 0;
  for (auto& nd : t->nodes) np += int(nd.prims.size());
  *n_leaf_prims = np;
  return int(t->nodes.size());
}
void port_bvh_nodes(void* h, int which, float* aabbs, int32_t* children, int32_t* leaf_range, int32_t* leaf_prims) {
  Scene& S = *static_cast<Scene*>(h);
  const Tree* t = pick(S, which);
  if (!t) return;
  int off = 0;
  for (size_t i = 0; i < t->nodes.size(); i++) {
    const Node& nd = t->nodes[i];
    for (int c = 0; c < 2; c++)
      for (int k = 0; k < 3; k++) {
        aabbs[i * 12 + c * 6 + k] = nd.box[c].lower[k];
        aabbs[i * 12 + c * 6 + 3 + k] = nd.box[c].upper[k];
      }
    children[i * 2] = nd.child[0];
    children[i * 2 + 1] = nd.child[1];
    leaf_range[i * 2] = off;
    leaf_range[i * 2 + 1] = int(nd.prims.size());
    for (int p : nd.prims) leaf_prims[off++] = p;
  }
}

// ---- sampler / RNG / camera known answers -----------------------------------------------------------
uint64_t port_hash_pixel(int px, int py, int sample_index) { return hash_pixel(px, py, sample_index); }
void port_rng_u64(uint64_t seed, int n, uint64_t* out) {
  Rng g(seed);
  for (int i = 0; i < n; i++) out[i] = g.next64();
}
float port_bluenoise(const uint8_t* sobol, const uint8_t* scramble, const uint8_t* rank, int px, int py,
                     int index, int dim) {
  return bluenoise(sobol, scramble, rank, px, py, index, dim);
}
// Sampler wrapper stream (sampler.h:286-309 + BlueSobolSampler :166-201): start_pixel(p, sample_index),
// n_rand randf() draws, then per sample a pattern of '1' (get1d) / '2' (get2d) draws with
// start_next_sample between samples. Same contract as ref_sampler_stream.
void port_sampler_stream(const uint8_t* sobol, const uint8_t* scramble, const uint8_t* rank, int px, int py,
                         int sample_index, int n_rand, float* out_rand, const char* pattern, int n_samples,
                         float* out_draws) {
  Rng g(hash_pixel(px, py, sample_index));
  for (int i = 0; i < n_rand; i++) out_rand[i] = g.nextf();
  int index = sample_index, dim = 0;
  auto sample = [&](int d) { return bluenoise(sobol, scramble, rank, px, py, index, d); };
  for (int s = 0; s < n_samples; s++) {
    for (const char* c = pattern; *c; c++) {
      if (*c == '1') {
        if (dim >= 256) dim = 2;
        *out_draws++ = sample(dim++);
      } else {
        if (dim + 1 >= 256) dim = 2;
        int d = dim;
        dim += 2;
        *out_draws++ = sample(d);
        *out_draws++ = sample(d + 1);
      }
    }
    dim = 0;
    index++;
  }
}
// ThinLenCamera::gen_ray (camera.cpp:17-33). cam = {position3, c2w 9 (column-major mat3), fov2d 2,
// len_radius, focus_distance}; in: n x {p_film.xy, u2.xy}; out rays n x 8.
void port_gen_rays(const float* cam, const float* pfilm_u2, int64_t n, float* rays) {
  V3 position(cam[0], cam[1], cam[2]);
  V3 cx(cam[3], cam[4], cam[5]), cy(cam[6], cam[7], cam[8]), cz(cam[9], cam[10], cam[11]);
  float fx = cam[12], fy = cam[13], len_radius = cam[14], focus = cam[15];
  for (int64_t i = 0; i < n; i++) {
    const float* a = pfilm_u2 + 4 * i;
    float pcx = (a[0] - 0.5f) * 2 * fx, pcy = (a[1] - 0.5f) * 2 * fy;
    V3 o, d;
    if (len_radius == 0.0f) {
      o = position;
      d = normalize(cx * pcx + cy * pcy + cz * 1.0f);
    } else {
      V3 dir = normalize(V3(pcx, pcy, 1.0f));
      V3 p_focus = focus * dir / dir.z;
      float r = std::sqrt(a[2]), theta = 2 * 3.14159265358979323846f * a[3];  // sample_disk_polar, sampling.h:13-18
      V3 p_len(len_radius * (r * std::cos(theta)), len_radius * (r * std::sin(theta)), 0.0f);
      o = position + p_len;
      V3 q = normalize(p_focus - p_len);
      d = cx * q.x + cy * q.y + cz * q.z;
    }
    float* out = rays + 8 * i;
    out[0] = o.x; out[1] = o.y; out[2] = o.z; out[3] = d.x; out[4] = d.y; out[5] = d.z;
    out[6] = 0.0f; out[7] = kFloatMax;
  }
}
// The camera rays of PathIntegrator::render (path.cpp:32-35) for sample `si` of every pixel:
// gen_ray((p + rand2f()) / size, rand2f()) with g++'s right-to-left argument evaluation, i.e. the
// lens pair is drawn first.
void port_primary_rays(const float* cam, int w, int h, int si, float* rays) {
  for (int py = 0; py < h; py++)
    for (int px = 0; px < w; px++) {
      Rng g(hash_pixel(px, py, 0));
      float in[4] = {0, 0, 0, 0};
      for (int s = 0; s <= si; s++) {
        float u0 = g.nextf(), u1 = g.nextf(), j0 = g.nextf(), j1 = g.nextf();
        in[0] = (float(px) + j0) / float(w); in[1] = (float(py) + j1) / float(h);
        in[2] = u0; in[3] = u1;
      }
      port_gen_rays(cam, in, 1, rays + 8 * (size_t(px) + size_t(py) * w));
    }
}
// SurfaceInteraction::spawn_ray (interaction.cpp:6-13, ray.h:25-37): in n x {p3, n3, wo3, distance}
void port_spawn_rays(const float* in, int64_t n, float* rays) {
  for (int64_t i = 0; i < n; i++) {
    const float* a = in + 10 * i;
    V3 p(a[0], a[1], a[2]), nn(a[3], a[4], a[5]), wo(a[6], a[7], a[8]);
    if (dot(nn, wo) < 0) nn = -nn;  // face_same_hemisphere(n, d)
    int32_t of[3] = {int32_t(256.0f * nn.x), int32_t(256.0f * nn.y), int32_t(256.0f * nn.z)};
    V3 o;
    for (int k = 0; k < 3; k++) {
      int32_t bits;
      float pk = p[k];
      std::memcpy(&bits, &pk, 4);
      bits += pk < 0 ? -of[k] : of[k];
      float pi;
      std::memcpy(&pi, &bits, 4);
      o[k] = std::fabs(pk) < (1.0f / 32.0f) ? pk + nn[k] * (1.0f / 65536.0f) : pi;
    }
    float* out = rays + 8 * i;
    out[0] = o.x; out[1] = o.y; out[2] = o.z; out[3] = wo.x; out[4] = wo.y; out[5] = wo.z;
    out[6] = 0.0f; out[7] = a[9] * (1.0f - 1e-3f);
  }
}

}  // extern "C"
