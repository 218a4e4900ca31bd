// trace_kernels.cu -- Accel::intersect / Accel::hit (reference src/pine/core/accel.h:16-24) over a
// caller-supplied ray set: the test hooks pb2_trace_closest / pb2_trace_any and the traversal
// micro-benchmark. One ray per thread, rays read as two 128-bit loads.
#include "kernels.cuh"
#include "surface.cuh"
#include "traverse.cuh"

namespace pb2 {

__device__ __forceinline__ Ray load_ray(const float* rays, size_t i) {
  const float4* p = reinterpret_cast<const float4*>(rays) + 2 * i;
  float4 a = __ldg(p), b = __ldg(p + 1);
  Ray r;
  r.o = V3(a.x, a.y, a.z);
  r.d = V3(a.w, b.x, b.y);
  r.tmin = b.z;
  r.tmax = b.w;
  return r;
}

template <int ACCEL>
__global__ void __launch_bounds__(128) k_trace_closest(DScene sc, const float* __restrict__ rays, size_t n,
                                                     int32_t* __restrict__ geom, int32_t* __restrict__ prim,
                                                     float* __restrict__ t, float* __restrict__ surf) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    Ray ray = load_ray(rays, i);
    Hit h;
    bool hit = ACCEL == 0 ? closest_production(sc, ray, h) : closest_ref_order(sc, ray, h);
    geom[i] = hit ? h.geom : -1;
    prim[i] = hit ? h.prim : -1;
    t[i] = ray.tmax;
    if (surf) {
      Surf s;
      s.p = V3(); s.n = V3(); s.uv = V2();
      if (hit) surface_info(sc, h.geom, h.prim, ray.o + ray.tmax * ray.d, s);
      float4* o = reinterpret_cast<float4*>(surf) + 2 * i;
      o[0] = make_float4(s.p.x, s.p.y, s.p.z, s.n.x);
      o[1] = make_float4(s.n.y, s.n.z, s.uv.x, s.uv.y);
    }
  }
}

template <int ACCEL>
__global__ void __launch_bounds__(128) k_trace_any(DScene sc, const float* __restrict__ rays, size_t n,
                                                 uint8_t* __restrict__ occ) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    Ray ray = load_ray(rays, i);
    occ[i] = (ACCEL == 0 ? any_wide(sc, ray) : any_ref_order(sc, ray)) ? 1 : 0;
  }
}

static int grid_for(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 16;  // persistent-style: a multiple of the SM count, grid-stride inside
  if (want > cap) want = cap;
  size_t g = ((want + kNumSMs - 1) / kNumSMs) * kNumSMs;
  return int(g ? g : kNumSMs);
}

void launch_trace_closest(const DScene& sc, const float* rays, size_t n, int accel, int32_t* geom,
                          int32_t* prim, float* t, float* surf, cudaStream_t st) {
  if (!n) return;
  int g = grid_for(n, 128);
  if (accel == 0) k_trace_closest<0><<<g, 128, 0, st>>>(sc, rays, n, geom, prim, t, surf);
  else k_trace_closest<1><<<g, 128, 0, st>>>(sc, rays, n, geom, prim, t, surf);
}
void launch_trace_any(const DScene& sc, const float* rays, size_t n, int accel, uint8_t* occ, cudaStream_t st) {
  if (!n) return;
  int g = grid_for(n, 128);
  if (accel == 0) k_trace_any<0><<<g, 128, 0, st>>>(sc, rays, n, occ);
  else k_trace_any<1><<<g, 128, 0, st>>>(sc, rays, n, occ);
}

}  // namespace pb2
