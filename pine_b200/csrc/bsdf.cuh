// bsdf.cuh -- local-frame BSDFs: sample / f / pdf / is_delta, and Material::sample_bxdf.
//
// Reference: scattering.h:8-152 (frame helpers, reflection / refraction, dielectric and Schlick Fresnel,
// the Trowbridge-Reitz distribution); bxdf.cpp:11-324 (Diffuse, Conductor, Refractive,
// RefractiveDielectric, DiffusiveDielectric), :329-372 (BSSRDF lobe); material.h:30-121 and
// material.cpp:9-28 (material -> BxDF choice). Same expressions in the same order; transcendental
// functions (powf, cosf, sinf) come from the CUDA math library, so shading agrees with the
// reference to libm rounding, not bit for bit.
#pragma once
#include "sampling.cuh"

namespace pb2 {

// Directions live in the shading frame: z is the normal, so cos(theta) of a direction is its z component.
__device__ __forceinline__ float cos_n(V3 w) { return w.z; }
__device__ __forceinline__ float abs_cos_n(V3 w) { return pabs(w.z); }
__device__ __forceinline__ bool same_side(V3 a, V3 b) { return a.z * b.z > 0.0f; }
__device__ __forceinline__ V3 up_facing(V3 v) { return v.z < 0.0f ? -v : v; }
__device__ __forceinline__ V3 mirror_z(V3 w) { return V3(-w.x, -w.y, w.z); }
__device__ __forceinline__ V3 mirror_about(V3 w, V3 axis) { return 2.0f * dot(w, axis) * axis - w; }

// Polar quantities of a direction, evaluated once per use site (scattering.h:8-44 spells them as separate
// helpers; the float expressions are the same: sin^2 = 1 - z^2, tan^2 = sin^2 / max(z^2, eps), cos/sin(phi) =
// clamp(x / sin, -1, 1) with the reference's value 1 for BOTH at the pole).
struct Polar {
  float z2, tan2, cphi, sphi;
};
__device__ __forceinline__ Polar polar_of(V3 w) {
  Polar q;
  q.z2 = sqr(w.z);
  const float s2 = 1.0f - q.z2;
  q.tan2 = s2 / pmax(q.z2, kEpsilon);
  const float s = sqrtf(s2);
  const bool pole = s == 0;
  q.cphi = pole ? 1 : pclamp(w.x / s, -1.0f, 1.0f);
  q.sphi = pole ? 1 : pclamp(w.y / s, -1.0f, 1.0f);
  return q;
}

// Snell refraction of `w` through a surface with normal `axis` (scattering.h:54-72). `eta` is flipped when `w`
// arrives from below; the value used goes to *eta_used. false = total internal reflection.
__device__ __forceinline__ bool refract_through(V3 w, V3 axis, float eta, V3& out, float* eta_used = nullptr) {
  float ci = dot(axis, w);
  if (ci < 0) {
    eta = 1.0f / eta;
    ci = -ci;
    axis = -axis;
  }
  const float s2t = pmax(0.0f, 1.0f - sqr(ci)) / sqr(eta);
  if (s2t >= 1) return false;
  out = -w / eta + (ci / eta - sqrtf(1.0f - s2t)) * axis;
  if (eta_used) *eta_used = eta;
  return true;
}
// Unpolarised Fresnel reflectance of a dielectric boundary (scattering.h:74-92).
__device__ __forceinline__ float fresnel_unpolarized(float ci, float eta) {
  if (ci < 0) {
    eta = 1 / eta;
    ci = -ci;
  }
  const float s2t = (1.0f - sqr(ci)) / sqr(eta);
  if (s2t >= 1.0f) return 1.0f;
  const float ct = sqrtf(1.0f - s2t);
  const float r_par = (eta * ci - ct) / (eta * ci + ct);
  const float r_perp = (ci - eta * ct) / (ci + eta * ct);
  return (sqr(r_par) + sqr(r_perp)) / 2.0f;
}
__device__ __forceinline__ V3 schlick(V3 f0, float c) { return f0 + (V3(1.0f) - f0) * powf(1.0f - c, 5.0f); }

// Anisotropic GGX microfacet lobe (scattering.h:98-152, TrowbridgeReitzDistribution; every caller passes
// ax == ay == roughness^2).
struct Ggx {
  float ax, ay;
  __device__ __forceinline__ float ndf(V3 m) const {
    const Polar q = polar_of(m);
    const float z4 = sqr(q.z2);
    if (z4 < 1e-6f) return 0.0f;
    const float e = q.tan2 * (sqr(q.cphi / ax) + sqr(q.sphi / ay));
    return 1.0f / (kPi * ax * ay * z4 * sqr(1 + e));
  }
  __device__ __forceinline__ float lambda(V3 w) const {
    const Polar q = polar_of(w);
    const float a2 = sqr(q.cphi * ax) + sqr(q.sphi * ay);
    return (sqrtf(1.0f + a2 * q.tan2) - 1.0f) / 2.0f;
  }
  __device__ __forceinline__ float shadow1(V3 w) const { return 1.0f / (1.0f + lambda(w)); }
  __device__ __forceinline__ float shadow2(V3 a, V3 b) const { return 1.0f / (1.0f + lambda(a) + lambda(b)); }
  // ndf(m) * shadow2(a, b); the first and last argument are the two directions, in the reference's order
  __device__ __forceinline__ float ndf_shadow(V3 a, V3 m, V3 b) const { return ndf(m) * shadow2(a, b); }
  // density of visible normals seen from w, floored like the reference's PDF()
  __device__ __forceinline__ float pdf(V3 w, V3 m) const {
    return pmax(shadow1(w) / abs_cos_n(w) * ndf(m) * absdot(w, m), kEpsilon);
  }
  // A visible normal for direction w: stretch w, pick a point of the projected hemisphere's disk, unstretch.
  __device__ __forceinline__ V3 sample_visible(V3 w, V2 u) const {
    V3 s = normalize(V3(ax * w.x, ay * w.y, w.z));
    if (s.z < 0.0f) s = -s;
    const V3 e1 = (s.z < 0.99999f) ? normalize(cross(V3(0, 0, 1), s)) : V3(1, 0, 0);
    const V3 e2 = cross(s, e1);
    V2 d = sample_disk_polar(u);
    const float half_chord = sqrtf(1.0f - sqr(d.x));
    d.y = plerp((1.0f + s.z) / 2, half_chord, d.y);
    const float dz = sqrtf(pmax(0.0f, 1.0f - length_squared(d)));
    const V3 m = d.x * e1 + d.y * e2 + dz * s;
    return normalize(V3(ax * m.x, ay * m.y, pmax(1e-6f, m.z)));
  }
};

enum BxdfKind { BX_DIFFUSE = 0, BX_CONDUCTOR, BX_REFRACTIVE, BX_REFR_DIELECTRIC, BX_DIFF_DIELECTRIC, BX_BSSRDF };

struct Bxdf {
  int kind;
  V3 albedo;
  float roughness, ior;
  V3 sigma_s;
  V3 wi;  // local
};
struct BsdfSample {
  V3 wo, f;
  float pdf;
  bool is_delta;
};

__device__ __forceinline__ bool bxdf_is_delta(const Bxdf& b) {
  switch (b.kind) {
    case BX_CONDUCTOR: case BX_REFRACTIVE: case BX_REFR_DIELECTRIC: return b.roughness < 1e-2f;
    default: return false;
  }
}

// psl::lerp(t, a, b) used by Ggx::sample_visible: see pmath.h plerp (vecmath.h:873 / psl math.h:119)
template <class Smp>
__device__ __forceinline__ bool bxdf_sample(const DScene& sc, const Bxdf& b, Smp& smp, BsdfSample& bs) {
  const V3 wi = b.wi;
  bs.is_delta = false;
  bs.pdf = 0;
  const float alpha = sqr(b.roughness);
  switch (b.kind) {
    case BX_DIFFUSE: {
      V3 wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
      if (cos_n(wi) < 0) wo = -wo;
      bs.wo = wo;
      bs.pdf = abs_cos_n(wo) / kPi;
      bs.f = b.albedo / kPi;
      return true;
    }
    case BX_BSSRDF: {
      V3 wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
      if (cos_n(wi) > 0) wo = -wo;
      bs.wo = wo;
      bs.pdf = abs_cos_n(wo) / kPi;
      bs.f = b.albedo / kPi;
      return true;
    }
    case BX_CONDUCTOR: {
      if (alpha < 1e-4f) {
        bs.wo = mirror_z(wi);
        bs.f = schlick(b.albedo, abs_cos_n(bs.wo)) / abs_cos_n(bs.wo);
        bs.pdf = 1.0f;
        bs.is_delta = true;
        return true;
      }
      Ggx d{alpha, alpha};
      V3 wm = d.sample_visible(wi, smp_get2d(sc, smp));
      V3 wo = mirror_about(wi, wm);
      if (!same_side(wi, wo)) return false;
      V3 fr = schlick(b.albedo, absdot(wi, wm));
      bs.wo = wo;
      bs.pdf = d.pdf(wi, wm) / (4 * absdot(wi, wm));
      bs.f = fr * (d.ndf_shadow(wo, wm, wi) / (4 * cos_n(wi) * cos_n(wo)));
      return true;
    }
    case BX_REFRACTIVE: {
      if (alpha < 1e-4f) {
        bs.wo = mirror_z(wi);
        bs.f = b.albedo;
        bs.pdf = abs_cos_n(bs.wo);
        bs.is_delta = true;
        return true;
      }
      Ggx d{alpha, alpha};
      V3 wm = d.sample_visible(wi, smp_get2d(sc, smp));
      V3 wo = mirror_about(wi, wm);
      if (!same_side(wi, wo)) return false;
      bs.wo = wo;
      bs.pdf = d.pdf(wi, wm) / (4 * absdot(wi, wm));
      bs.f = b.albedo * (d.ndf_shadow(wo, wm, wi) / (4 * cos_n(wi) * cos_n(wo)));
      return true;
    }
    case BX_REFR_DIELECTRIC: {
      float fr = fresnel_unpolarized(cos_n(wi), b.ior);
      if (alpha < 1e-4f) {
        if (smp_get1d(sc, smp) < fr) {
          bs.wo = mirror_z(wi);
          bs.f = b.albedo * (fr / abs_cos_n(bs.wo));
          bs.pdf = fr;
          bs.is_delta = true;
        } else {
          if (!refract_through(wi, V3(0, 0, 1), b.ior, bs.wo)) return false;
          bs.f = b.albedo * ((1 - fr) / abs_cos_n(bs.wo));
          bs.pdf = 1 - fr;
          bs.is_delta = true;
        }
        return true;
      }
      Ggx d{alpha, alpha};
      V3 wm = d.sample_visible(wi, smp_get2d(sc, smp));
      if (smp_get1d(sc, smp) < fr) {
        V3 wo = mirror_about(wi, wm);
        if (!same_side(wi, wo)) return false;
        bs.wo = wo;
        bs.pdf = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
        bs.f = b.albedo * (fr * d.ndf_shadow(wo, wm, wi) / (4 * cos_n(wi) * cos_n(wo)));
      } else {
        float eta = 1.0f;
        if (!refract_through(wi, wm, b.ior, bs.wo, &eta)) return false;
        V3 wo = bs.wo;
        float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta);
        bs.pdf = (1 - fr) * d.pdf(wi, wm) * absdot(wo, wm) / denom;
        bs.f = b.albedo * ((1 - fr) * d.ndf(wm) * d.shadow2(wi, wo) *
                           pabs(dot(wo, wm) * dot(wi, wm) / (denom * cos_n(wi) * cos_n(wo))));
      }
      return true;
    }
    case BX_DIFF_DIELECTRIC: {
      float fr = fresnel_unpolarized(cos_n(wi), b.ior);
      if (alpha < 1e-4f) {
        if (smp_get1d(sc, smp) < fr) {
          bs.wo = mirror_z(wi);
          bs.f = V3(fr);
          bs.pdf = fr * abs_cos_n(bs.wo);
          bs.is_delta = true;
        } else {
          bs.wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
          bs.f = b.albedo * ((1 - fr) / kPi);
          bs.pdf = (1 - fr) * abs_cos_n(bs.wo) / kPi;
        }
        return true;
      }
      Ggx d{alpha, alpha};
      V3 wm = d.sample_visible(wi, smp_get2d(sc, smp));
      if (smp_get1d(sc, smp) < fr) {
        V3 wo = mirror_about(wi, wm);
        if (!same_side(wi, wo)) return false;
        bs.wo = wo;
        bs.f = V3(fr * d.ndf_shadow(wi, wm, wo) / (4 * cos_n(wi) * cos_n(wo)));
        bs.pdf = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      } else {
        bs.wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
        bs.f = b.albedo * ((1 - fr) / kPi);
        bs.pdf = abs_cos_n(bs.wo) * (1 - fr) / kPi;
      }
      return true;
    }
  }
  return false;
}

__device__ __forceinline__ V3 bxdf_f(const Bxdf& b, V3 wo) {
  const V3 wi = b.wi;
  const float alpha = sqr(b.roughness);
  Ggx d{alpha, alpha};
  switch (b.kind) {
    case BX_DIFFUSE:
      if (!same_side(wi, wo)) return V3(0.0f);
      return b.albedo / kPi;
    case BX_BSSRDF: return b.albedo / kPi;
    case BX_CONDUCTOR: {
      if (!same_side(wi, wo)) return V3();
      V3 wm = normalize(wi + wo);
      if (is_zero(wm)) return V3();
      V3 fr = schlick(b.albedo, absdot(wi, wm));
      return fr * (d.ndf_shadow(wo, wm, wi) / (4 * abs_cos_n(wo) * abs_cos_n(wi)));
    }
    case BX_REFRACTIVE: {
      float cosO = cos_n(wo), cosI = cos_n(wi);
      if (!(cosI * cosO > 0)) return V3();
      V3 wm = up_facing(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      return b.albedo * (d.ndf_shadow(wi, wm, wo) / pabs(4 * cosI * cosO));
    }
    case BX_REFR_DIELECTRIC: {
      float cosO = cos_n(wo), cosI = cos_n(wi);
      bool reflect = cosI * cosO > 0;
      int eta = 1;  // the reference declares `auto eta = 1` (an int): ior truncates (bxdf.cpp:219-221)
      if (!reflect) eta = int(cosI > 0 ? b.ior : 1 / b.ior);
      V3 wm = up_facing(normalize(wo * float(eta) + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      float fr = fresnel_unpolarized(dot(wi, wm), b.ior);
      if (reflect) return b.albedo * (fr * d.ndf_shadow(wi, wm, wo) / pabs(4 * cosI * cosO));
      float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta) * cosI * cosO;
      return b.albedo * ((1 - fr) * d.ndf(wm) * d.shadow2(wi, wo) * pabs(dot(wo, wm) * dot(wi, wm) / denom));
    }
    case BX_DIFF_DIELECTRIC: {
      if (!same_side(wi, wo)) return V3();
      float cosO = cos_n(wo), cosI = cos_n(wi);
      V3 wm = up_facing(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      float fr = fresnel_unpolarized(dot(wi, wm), b.ior);
      V3 diffused = b.albedo * (1 - fr) / kPi;
      if (alpha < 1e-4f) return diffused;
      float reflected = fr * d.ndf_shadow(wo, wm, wi) / pabs(4 * cosI * cosO);
      return V3(reflected) + diffused;
    }
  }
  return V3();
}

__device__ __forceinline__ float bxdf_pdf(const Bxdf& b, V3 wo) {
  const V3 wi = b.wi;
  const float alpha = sqr(b.roughness);
  Ggx d{alpha, alpha};
  switch (b.kind) {
    case BX_DIFFUSE:
      if (!same_side(wi, wo)) return 0.0f;
      return abs_cos_n(wo) / kPi;
    case BX_BSSRDF: return abs_cos_n(wo) / kPi;
    case BX_CONDUCTOR: {
      if (!same_side(wi, wo)) return 0.0f;
      V3 wm = normalize(wi + wo);
      if (is_zero(wm)) return 0.0f;
      wm = up_facing(wm);
      return d.pdf(wi, wm) / (4 * absdot(wi, wm));
    }
    case BX_REFRACTIVE: {
      float cosO = cos_n(wo), cosI = cos_n(wi);
      if (!(cosI * cosO > 0)) return 0.0f;
      V3 wm = up_facing(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      return d.pdf(wi, wm) / (4 * absdot(wi, wm));
    }
    case BX_REFR_DIELECTRIC: {
      float cosO = cos_n(wo), cosI = cos_n(wi);
      bool reflect = cosI * cosO > 0;
      int eta = 1;
      if (!reflect) eta = int(cosI > 0 ? b.ior : 1 / b.ior);
      V3 wm = up_facing(normalize(wo * float(eta) + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      float fr = fresnel_unpolarized(dot(wi, wm), b.ior);
      if (reflect) return fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta);
      float dwm_dwo = absdot(wo, wm) / denom;
      return (1 - fr) * d.pdf(wi, wm) * dwm_dwo;
    }
    case BX_DIFF_DIELECTRIC: {
      if (!same_side(wi, wo)) return 0.0f;
      float cosO = cos_n(wo), cosI = cos_n(wi);
      V3 wm = up_facing(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      float fr = fresnel_unpolarized(dot(wi, wm), b.ior);
      float pt = (1 - fr) * abs_cos_n(wo) / kPi;
      if (alpha < 1e-4f) return pt;
      float pr = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      return pr + pt;
    }
  }
  return 0.0f;
}

}  // namespace pb2
