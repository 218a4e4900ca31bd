// bsdf.cuh -- local-frame BSDFs: sample / f / pdf / is_delta, and Material::sample_bxdf.
//
// Reference: scattering.h:8-152 (frame helpers, Reflect/Refract, FrDielectric, FrSchlick,
// TrowbridgeReitzDistribution); bxdf.cpp:11-324 (Diffuse, Conductor, Refractive,
// RefractiveDielectric, DiffusiveDielectric), :329-372 (BSSRDF lobe); material.h:30-121 and
// material.cpp:9-28 (material -> BxDF choice). Same expressions in the same order; transcendental
// functions (powf, cosf, sinf) come from the CUDA math library, so shading agrees with the
// reference to libm rounding, not bit for bit.
#pragma once
#include "sampling.cuh"

namespace pb2 {

__device__ __forceinline__ float CosTheta(V3 w) { return w.z; }
__device__ __forceinline__ float Cos2Theta(V3 w) { return sqr(w.z); }
__device__ __forceinline__ float AbsCosTheta(V3 w) { return pabs(w.z); }
__device__ __forceinline__ float Sin2Theta(V3 w) { return 1.0f - Cos2Theta(w); }
__device__ __forceinline__ float SinTheta(V3 w) { return sqrtf(Sin2Theta(w)); }
__device__ __forceinline__ float Tan2Theta(V3 w) { return Sin2Theta(w) / pmax(Cos2Theta(w), kEpsilon); }
__device__ __forceinline__ float CosPhi(V3 w) {
  float s = SinTheta(w);
  return (s == 0) ? 1 : pclamp(w.x / s, -1.0f, 1.0f);
}
__device__ __forceinline__ float SinPhi(V3 w) {
  float s = SinTheta(w);
  return (s == 0) ? 1 : pclamp(w.y / s, -1.0f, 1.0f);
}
__device__ __forceinline__ bool SameHemisphere(V3 a, V3 b) { return a.z * b.z > 0.0f; }
__device__ __forceinline__ V3 FaceNormal(V3 v) { return v.z < 0.0f ? -v : v; }
__device__ __forceinline__ V3 Reflect(V3 w) { return V3(-w.x, -w.y, w.z); }
__device__ __forceinline__ V3 Reflect(V3 wi, V3 n) { return 2.0f * dot(wi, n) * n - wi; }
__device__ __forceinline__ bool Refract(V3 wi, V3 n, float eta, V3& wt, float* etap = nullptr) {
  float cosThetaI = dot(n, wi);
  if (cosThetaI < 0) {
    eta = 1.0f / eta;
    cosThetaI = -cosThetaI;
    n = -n;
  }
  float sin2ThetaI = pmax(0.0f, 1.0f - sqr(cosThetaI));
  float sin2ThetaT = sin2ThetaI / sqr(eta);
  if (sin2ThetaT >= 1) return false;
  float cosThetaT = sqrtf(1.0f - sin2ThetaT);
  wt = -wi / eta + (cosThetaI / eta - cosThetaT) * n;
  if (etap) *etap = eta;
  return true;
}
__device__ __forceinline__ float FrDielectric(float cosThetaI, float eta) {
  if (cosThetaI < 0) {
    eta = 1 / eta;
    cosThetaI = -cosThetaI;
  }
  float sin2ThetaI = 1.0f - sqr(cosThetaI);
  float sin2ThetaT = sin2ThetaI / sqr(eta);
  if (sin2ThetaT >= 1.0f) return 1.0f;
  float cosThetaT = sqrtf(1.0f - sin2ThetaT);
  float rParl = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
  float rPerp = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
  return (sqr(rParl) + sqr(rPerp)) / 2.0f;
}
__device__ __forceinline__ V3 FrSchlick(V3 F0, float cosTheta) {
  return F0 + (V3(1.0f) - F0) * powf(1.0f - cosTheta, 5.0f);
}

struct TRDist {  // TrowbridgeReitzDistribution(alpha, alpha)
  float ax, ay;
  __device__ __forceinline__ float D(V3 wm) const {
    float tan2 = Tan2Theta(wm);
    float cos4 = sqr(Cos2Theta(wm));
    if (cos4 < 1e-6f) return 0.0f;
    float e = tan2 * (sqr(CosPhi(wm) / ax) + sqr(SinPhi(wm) / ay));
    return 1.0f / (kPi * ax * ay * cos4 * sqr(1 + e));
  }
  __device__ __forceinline__ float Lambda(V3 w) const {
    float tan2 = Tan2Theta(w);
    float alpha2 = sqr(CosPhi(w) * ax) + sqr(SinPhi(w) * ay);
    return (sqrtf(1.0f + alpha2 * tan2) - 1.0f) / 2.0f;
  }
  __device__ __forceinline__ float G1(V3 w) const { return 1.0f / (1.0f + Lambda(w)); }
  __device__ __forceinline__ float G(V3 wi, V3 wo) const { return 1.0f / (1.0f + Lambda(wi) + Lambda(wo)); }
  __device__ __forceinline__ float D_G(V3 wi, V3 wm, V3 wo) const { return D(wm) * G(wi, wo); }
  __device__ __forceinline__ float Dw(V3 w, V3 wm) const { return G1(w) / AbsCosTheta(w) * D(wm) * absdot(w, wm); }
  __device__ __forceinline__ float pdf(V3 w, V3 wm) const { return pmax(Dw(w, wm), kEpsilon); }
  __device__ __forceinline__ V3 SampleWm(V3 w, V2 u) const {
    V3 wh = normalize(V3(ax * w.x, ay * w.y, w.z));
    if (wh.z < 0.0f) wh = -wh;
    V3 T1 = (wh.z < 0.99999f) ? normalize(cross(V3(0, 0, 1), wh)) : V3(1, 0, 0);
    V3 T2 = cross(wh, T1);
    V2 p = sample_disk_polar(u);
    float h = sqrtf(1.0f - sqr(p.x));
    p.y = plerp((1.0f + wh.z) / 2, h, p.y);
    float pz = sqrtf(pmax(0.0f, 1.0f - length_squared(p)));
    V3 nh = p.x * T1 + p.y * T2 + pz * wh;
    return normalize(V3(ax * nh.x, ay * nh.y, pmax(1e-6f, nh.z)));
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

// psl::lerp(t, a, b) used by SampleWm: see pmath.h plerp (vecmath.h:873 / psl math.h:119)
template <class Smp>
__device__ __forceinline__ bool bxdf_sample(const DScene& sc, const Bxdf& b, Smp& smp, BsdfSample& bs) {
  const V3 wi = b.wi;
  bs.is_delta = false;
  bs.pdf = 0;
  const float alpha = sqr(b.roughness);
  switch (b.kind) {
    case BX_DIFFUSE: {
      V3 wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
      if (CosTheta(wi) < 0) wo = -wo;
      bs.wo = wo;
      bs.pdf = AbsCosTheta(wo) / kPi;
      bs.f = b.albedo / kPi;
      return true;
    }
    case BX_BSSRDF: {
      V3 wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
      if (CosTheta(wi) > 0) wo = -wo;
      bs.wo = wo;
      bs.pdf = AbsCosTheta(wo) / kPi;
      bs.f = b.albedo / kPi;
      return true;
    }
    case BX_CONDUCTOR: {
      if (alpha < 1e-4f) {
        bs.wo = Reflect(wi);
        bs.f = FrSchlick(b.albedo, AbsCosTheta(bs.wo)) / AbsCosTheta(bs.wo);
        bs.pdf = 1.0f;
        bs.is_delta = true;
        return true;
      }
      TRDist d{alpha, alpha};
      V3 wm = d.SampleWm(wi, smp_get2d(sc, smp));
      V3 wo = Reflect(wi, wm);
      if (!SameHemisphere(wi, wo)) return false;
      V3 fr = FrSchlick(b.albedo, absdot(wi, wm));
      bs.wo = wo;
      bs.pdf = d.pdf(wi, wm) / (4 * absdot(wi, wm));
      bs.f = fr * (d.D_G(wo, wm, wi) / (4 * CosTheta(wi) * CosTheta(wo)));
      return true;
    }
    case BX_REFRACTIVE: {
      if (alpha < 1e-4f) {
        bs.wo = Reflect(wi);
        bs.f = b.albedo;
        bs.pdf = AbsCosTheta(bs.wo);
        bs.is_delta = true;
        return true;
      }
      TRDist d{alpha, alpha};
      V3 wm = d.SampleWm(wi, smp_get2d(sc, smp));
      V3 wo = Reflect(wi, wm);
      if (!SameHemisphere(wi, wo)) return false;
      bs.wo = wo;
      bs.pdf = d.pdf(wi, wm) / (4 * absdot(wi, wm));
      bs.f = b.albedo * (d.D_G(wo, wm, wi) / (4 * CosTheta(wi) * CosTheta(wo)));
      return true;
    }
    case BX_REFR_DIELECTRIC: {
      float fr = FrDielectric(CosTheta(wi), b.ior);
      if (alpha < 1e-4f) {
        if (smp_get1d(sc, smp) < fr) {
          bs.wo = Reflect(wi);
          bs.f = b.albedo * (fr / AbsCosTheta(bs.wo));
          bs.pdf = fr;
          bs.is_delta = true;
        } else {
          if (!Refract(wi, V3(0, 0, 1), b.ior, bs.wo)) return false;
          bs.f = b.albedo * ((1 - fr) / AbsCosTheta(bs.wo));
          bs.pdf = 1 - fr;
          bs.is_delta = true;
        }
        return true;
      }
      TRDist d{alpha, alpha};
      V3 wm = d.SampleWm(wi, smp_get2d(sc, smp));
      if (smp_get1d(sc, smp) < fr) {
        V3 wo = Reflect(wi, wm);
        if (!SameHemisphere(wi, wo)) return false;
        bs.wo = wo;
        bs.pdf = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
        bs.f = b.albedo * (fr * d.D_G(wo, wm, wi) / (4 * CosTheta(wi) * CosTheta(wo)));
      } else {
        float eta = 1.0f;
        if (!Refract(wi, wm, b.ior, bs.wo, &eta)) return false;
        V3 wo = bs.wo;
        float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta);
        bs.pdf = (1 - fr) * d.pdf(wi, wm) * absdot(wo, wm) / denom;
        bs.f = b.albedo * ((1 - fr) * d.D(wm) * d.G(wi, wo) *
                           pabs(dot(wo, wm) * dot(wi, wm) / (denom * CosTheta(wi) * CosTheta(wo))));
      }
      return true;
    }
    case BX_DIFF_DIELECTRIC: {
      float fr = FrDielectric(CosTheta(wi), b.ior);
      if (alpha < 1e-4f) {
        if (smp_get1d(sc, smp) < fr) {
          bs.wo = Reflect(wi);
          bs.f = V3(fr);
          bs.pdf = fr * AbsCosTheta(bs.wo);
          bs.is_delta = true;
        } else {
          bs.wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
          bs.f = b.albedo * ((1 - fr) / kPi);
          bs.pdf = (1 - fr) * AbsCosTheta(bs.wo) / kPi;
        }
        return true;
      }
      TRDist d{alpha, alpha};
      V3 wm = d.SampleWm(wi, smp_get2d(sc, smp));
      if (smp_get1d(sc, smp) < fr) {
        V3 wo = Reflect(wi, wm);
        if (!SameHemisphere(wi, wo)) return false;
        bs.wo = wo;
        bs.f = V3(fr * d.D_G(wi, wm, wo) / (4 * CosTheta(wi) * CosTheta(wo)));
        bs.pdf = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      } else {
        bs.wo = cosine_weighted_hemisphere(smp_get2d(sc, smp));
        bs.f = b.albedo * ((1 - fr) / kPi);
        bs.pdf = AbsCosTheta(bs.wo) * (1 - fr) / kPi;
      }
      return true;
    }
  }
  return false;
}

__device__ __forceinline__ V3 bxdf_f(const Bxdf& b, V3 wo) {
  const V3 wi = b.wi;
  const float alpha = sqr(b.roughness);
  TRDist d{alpha, alpha};
  switch (b.kind) {
    case BX_DIFFUSE:
      if (!SameHemisphere(wi, wo)) return V3(0.0f);
      return b.albedo / kPi;
    case BX_BSSRDF: return b.albedo / kPi;
    case BX_CONDUCTOR: {
      if (!SameHemisphere(wi, wo)) return V3();
      V3 wm = normalize(wi + wo);
      if (is_zero(wm)) return V3();
      V3 fr = FrSchlick(b.albedo, absdot(wi, wm));
      return fr * (d.D_G(wo, wm, wi) / (4 * AbsCosTheta(wo) * AbsCosTheta(wi)));
    }
    case BX_REFRACTIVE: {
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      if (!(cosI * cosO > 0)) return V3();
      V3 wm = FaceNormal(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      return b.albedo * (d.D_G(wi, wm, wo) / pabs(4 * cosI * cosO));
    }
    case BX_REFR_DIELECTRIC: {
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      bool reflect = cosI * cosO > 0;
      int eta = 1;  // the reference declares `auto eta = 1` (an int): ior truncates (bxdf.cpp:219-221)
      if (!reflect) eta = int(cosI > 0 ? b.ior : 1 / b.ior);
      V3 wm = FaceNormal(normalize(wo * float(eta) + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      float fr = FrDielectric(dot(wi, wm), b.ior);
      if (reflect) return b.albedo * (fr * d.D_G(wi, wm, wo) / pabs(4 * cosI * cosO));
      float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta) * cosI * cosO;
      return b.albedo * ((1 - fr) * d.D(wm) * d.G(wi, wo) * pabs(dot(wo, wm) * dot(wi, wm) / denom));
    }
    case BX_DIFF_DIELECTRIC: {
      if (!SameHemisphere(wi, wo)) return V3();
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      V3 wm = FaceNormal(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return V3();
      float fr = FrDielectric(dot(wi, wm), b.ior);
      V3 diffused = b.albedo * (1 - fr) / kPi;
      if (alpha < 1e-4f) return diffused;
      float reflected = fr * d.D_G(wo, wm, wi) / pabs(4 * cosI * cosO);
      return V3(reflected) + diffused;
    }
  }
  return V3();
}

__device__ __forceinline__ float bxdf_pdf(const Bxdf& b, V3 wo) {
  const V3 wi = b.wi;
  const float alpha = sqr(b.roughness);
  TRDist d{alpha, alpha};
  switch (b.kind) {
    case BX_DIFFUSE:
      if (!SameHemisphere(wi, wo)) return 0.0f;
      return AbsCosTheta(wo) / kPi;
    case BX_BSSRDF: return AbsCosTheta(wo) / kPi;
    case BX_CONDUCTOR: {
      if (!SameHemisphere(wi, wo)) return 0.0f;
      V3 wm = normalize(wi + wo);
      if (is_zero(wm)) return 0.0f;
      wm = FaceNormal(wm);
      return d.pdf(wi, wm) / (4 * absdot(wi, wm));
    }
    case BX_REFRACTIVE: {
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      if (!(cosI * cosO > 0)) return 0.0f;
      V3 wm = FaceNormal(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      return d.pdf(wi, wm) / (4 * absdot(wi, wm));
    }
    case BX_REFR_DIELECTRIC: {
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      bool reflect = cosI * cosO > 0;
      int eta = 1;
      if (!reflect) eta = int(cosI > 0 ? b.ior : 1 / b.ior);
      V3 wm = FaceNormal(normalize(wo * float(eta) + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      float fr = FrDielectric(dot(wi, wm), b.ior);
      if (reflect) return fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      float denom = sqr(dot(wo, wm) + dot(wi, wm) / eta);
      float dwm_dwo = absdot(wo, wm) / denom;
      return (1 - fr) * d.pdf(wi, wm) * dwm_dwo;
    }
    case BX_DIFF_DIELECTRIC: {
      if (!SameHemisphere(wi, wo)) return 0.0f;
      float cosO = CosTheta(wo), cosI = CosTheta(wi);
      V3 wm = FaceNormal(normalize(wo + wi));
      if (dot(wm, wo) * cosO <= 0 || dot(wm, wi) * cosI <= 0) return 0.0f;
      float fr = FrDielectric(dot(wi, wm), b.ior);
      float pt = (1 - fr) * AbsCosTheta(wo) / kPi;
      if (alpha < 1e-4f) return pt;
      float pr = fr * d.pdf(wi, wm) / (4 * absdot(wi, wm));
      return pr + pt;
    }
  }
  return 0.0f;
}

}  // namespace pb2
