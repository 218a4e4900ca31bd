// pmath.h -- float vector/matrix arithmetic shared by the host builder and the sm_100a kernels.
//
// Every operator evaluates in the same order as the reference's templates
// (src/pine/core/vecmath.h:160-300 Vector3, :700-790 dot/cross/length/normalize, :1062-1068
// safe_rcp, :1172-1195 look_at/coordinate_system) so that, compiled with -fmad=false (device) and
// -ffp-contract=off (host), results are bit-identical to the reference's IEEE binary32 results.
// min/max are compare-selects like psl::min/max (src/psl/math.h:19-26), NOT fminf/fmaxf.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define PB2_HD __host__ __device__ __forceinline__
#else
#define PB2_HD inline
#endif

namespace pb2 {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kEpsilon = 1.1920928955078125e-07f;  // numeric_limits<float>::epsilon()
constexpr float kOneMinusEpsilon = 0x1.fffffep-1f;
constexpr float kFloatMax = 3.402823466e+38f;

PB2_HD float pmin(float a, float b) { return a < b ? a : b; }
PB2_HD float pmax(float a, float b) { return a > b ? a : b; }
PB2_HD int pmin(int a, int b) { return a < b ? a : b; }
PB2_HD int pmax(int a, int b) { return a > b ? a : b; }
PB2_HD float sqr(float x) { return x * x; }
PB2_HD float pabs(float x) { return fabsf(x); }
PB2_HD float pclamp(float v, float lo, float hi) { return pmin(pmax(v, lo), hi); }  // psl math.h:110
PB2_HD float plerp(float t, float a, float b) { return a * (1.0f - t) + b * t; }     // psl math.h:119

struct V2 {
  float x, y;
  PB2_HD V2() : x(0), y(0) {}
  PB2_HD V2(float x_, float y_) : x(x_), y(y_) {}
};
struct V3 {
  float x, y, z;
  PB2_HD V3() : x(0), y(0), z(0) {}
  PB2_HD explicit V3(float v) : x(v), y(v), z(v) {}
  PB2_HD V3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
  PB2_HD float& operator[](int i) { return (&x)[i]; }
  PB2_HD const float& operator[](int i) const { return (&x)[i]; }
  PB2_HD V3 operator-() const { return V3(-x, -y, -z); }
};
PB2_HD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
PB2_HD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
PB2_HD V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
PB2_HD V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
PB2_HD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
PB2_HD V3 operator*(float s, V3 a) { return V3(s * a.x, s * a.y, s * a.z); }
PB2_HD V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
PB2_HD V3& operator+=(V3& a, V3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
PB2_HD V3& operator*=(V3& a, V3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
PB2_HD bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
PB2_HD V2 operator+(V2 a, V2 b) { return V2(a.x + b.x, a.y + b.y); }
PB2_HD V2 operator-(V2 a, V2 b) { return V2(a.x - b.x, a.y - b.y); }
PB2_HD V2 operator*(V2 a, V2 b) { return V2(a.x * b.x, a.y * b.y); }
PB2_HD V2 operator*(V2 a, float s) { return V2(a.x * s, a.y * s); }
PB2_HD V2 operator*(float s, V2 a) { return V2(s * a.x, s * a.y); }

PB2_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PB2_HD float absdot(V3 a, V3 b) { return pabs(dot(a, b)); }
PB2_HD V3 cross(V3 a, V3 b) {
  return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
PB2_HD float length_squared(V3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
PB2_HD float length_squared(V2 v) { return v.x * v.x + v.y * v.y; }
PB2_HD float length(V3 v) { return sqrtf(length_squared(v)); }
PB2_HD float distance(V3 a, V3 b) { return length(a - b); }
PB2_HD V3 normalize(V3 v) {
  float len = length(v);
  if (len == 0) return v;
  return v / len;
}
PB2_HD V3 normalize(V3 v, float& len) {
  len = length(v);
  if (len == 0) return v;
  return v / len;
}
PB2_HD V3 vmin(V3 a, V3 b) { return V3(pmin(a.x, b.x), pmin(a.y, b.y), pmin(a.z, b.z)); }
PB2_HD V3 vmax(V3 a, V3 b) { return V3(pmax(a.x, b.x), pmax(a.y, b.y), pmax(a.z, b.z)); }
PB2_HD V3 vabs(V3 a) { return V3(pabs(a.x), pabs(a.y), pabs(a.z)); }
PB2_HD bool is_zero(V3 v) { return v.x == 0 && v.y == 0 && v.z == 0; }
PB2_HD int max_axis(V3 v) {  // vecmath.h:1263-1268
  if (v[0] > v[1]) return v[0] > v[2] ? 0 : 2;
  return v[1] > v[2] ? 1 : 2;
}
// lerp(u, v, a, b, c) = (1 - u - v) * a + u * b + v * c  (vecmath.h:876-879)
PB2_HD V3 lerp3(float u, float v, V3 a, V3 b, V3 c) { return (1.0f - u - v) * a + u * b + v * c; }
PB2_HD V2 lerp3(float u, float v, V2 a, V2 b, V2 c) { return (1.0f - u - v) * a + u * b + v * c; }
PB2_HD V3 face_same_hemisphere(V3 v, V3 ref) { return dot(v, ref) < 0 ? -v : v; }
PB2_HD float safe_rcp(float v) { return v == 0.0f ? 1e+20f : 1.0f / v; }

// Column-major 3x3 / 4x4, columns x,y,z(,w) as in Matrix3/Matrix4 (vecmath.h:494-616).
struct M3 {
  V3 x, y, z;
  PB2_HD M3() : x(1, 0, 0), y(0, 1, 0), z(0, 0, 1) {}
  PB2_HD M3(V3 x_, V3 y_, V3 z_) : x(x_), y(y_), z(z_) {}
  PB2_HD V3& operator[](int i) { return (&x)[i]; }
  PB2_HD const V3& operator[](int i) const { return (&x)[i]; }
  PB2_HD V3 row(int i) const { return V3(x[i], y[i], z[i]); }
};
PB2_HD V3 operator*(const M3& m, V3 v) { return m.x * v.x + m.y * v.y + m.z * v.z; }
PB2_HD M3 transpose(const M3& m) { return M3(m.row(0), m.row(1), m.row(2)); }

struct V4 {
  float x, y, z, w;
  PB2_HD V4() : x(0), y(0), z(0), w(0) {}
  PB2_HD V4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
  PB2_HD V4(V3 v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
  PB2_HD float& operator[](int i) { return (&x)[i]; }
  PB2_HD const float& operator[](int i) const { return (&x)[i]; }
};
PB2_HD V4 operator+(V4 a, V4 b) { return V4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
PB2_HD V4 operator*(V4 a, float s) { return V4(a.x * s, a.y * s, a.z * s, a.w * s); }
struct M4 {
  V4 x, y, z, w;
  PB2_HD M4() : x(1, 0, 0, 0), y(0, 1, 0, 0), z(0, 0, 1, 0), w(0, 0, 0, 1) {}
  PB2_HD M4(V4 x_, V4 y_, V4 z_, V4 w_) : x(x_), y(y_), z(z_), w(w_) {}
  PB2_HD V4& operator[](int i) { return (&x)[i]; }
  PB2_HD const V4& operator[](int i) const { return (&x)[i]; }
};
// Matrix4 * Vector3 = Vector3(m.x*v.x + m.y*v.y + m.z*v.z + m.w)  (vecmath.h:704-707)
PB2_HD V3 mul_point(const M4& m, V3 v) {
  V4 r = m.x * v.x + m.y * v.y + m.z * v.z + m.w;
  return V3(r.x, r.y, r.z);
}
PB2_HD M3 upper3(const M4& m) {  // Matrix4::operator Matrix3 (vecmath.h:567)
  return M3(V3(m.x.x, m.x.y, m.x.z), V3(m.y.x, m.y.y, m.y.z), V3(m.z.x, m.z.y, m.z.z));
}

// coordinate_system(n) (vecmath.h:1182-1195): columns t, b, n
PB2_HD M3 coordinate_system(V3 n) {
  V3 t;
  if (pabs(n.x) > pabs(n.y))
    t = normalize(cross(n, V3(0, 1, 0)));
  else
    t = normalize(cross(n, V3(1, 0, 0)));
  V3 b = cross(n, t);
  return M3(t, b, n);
}

PB2_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
PB2_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

}  // namespace pb2
