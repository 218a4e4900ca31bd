// trace_kernels.cu -- Accel::intersect / Accel::hit (reference src/pine/core/accel.h:16-24) over a
// caller-supplied ray set: the test hooks pb2_trace_closest / pb2_trace_any and the traversal
// micro-benchmark. Same persistent wide walk + fallback as the renderer (wide_persistent.cuh).
#include "kernels.cuh"
#include "surface.cuh"
#include "wide_persistent.cuh"

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

struct HookIO {
  const float* rays;
  int32_t* geom;
  int32_t* prim;
  float* t;
  float* surf;
  uint8_t* occ;
  __device__ __forceinline__ Ray load(int i) const { return load_ray(rays, size_t(i)); }
  __device__ __forceinline__ void store_closest(const DScene& sc, int i, const Ray& ray, bool hit, Hit h) const {
    geom[i] = hit ? h.geom : -1;
    prim[i] = hit ? h.prim : -1;
    t[i] = ray.tmax;
    if (surf) {
      Surf s;
      s.p = V3(); s.n = V3(); s.uv = V2();
      if (hit) surface_info(sc, h.geom, h.prim, ray.o + ray.tmax * ray.d, s);
      float4* o = reinterpret_cast<float4*>(surf) + 2 * size_t(i);
      o[0] = make_float4(s.p.x, s.p.y, s.p.z, s.n.x);
      o[1] = make_float4(s.n.y, s.n.z, s.uv.x, s.uv.y);
    }
  }
  __device__ __forceinline__ void store_any(int i, bool occluded) const { occ[i] = occluded ? 1 : 0; }
};

template <bool ANY>
__global__ void __launch_bounds__(kTraceBlock, PB2_TRACE_MINBLOCKS) k_hook_wide(DScene sc, int n, WideQueue wq, HookIO io) {
  wide_persistent<ANY>(sc, n, wq, io);
}
template <bool ANY>
__global__ void __launch_bounds__(kTraceBlock) k_hook_fallback(DScene sc, WideQueue wq, HookIO io) {
  wide_fallback<ANY>(sc, wq, io);
}
template <bool ANY>
__global__ void __launch_bounds__(kTraceBlock) k_hook_ref(DScene sc, int n, HookIO io) {
  ref_order_all<ANY>(sc, n, io);
}

template <bool ANY>
static void launch_hook(const DScene& sc, int n, int accel, const HookIO& io, int* scratch, cudaStream_t st) {
  const bool wide = accel == 0 && sc.wide_ready &&
                    (ANY ? sc.n_box_shapes <= kMaxBoxShapes : !sc.order_dependent);
  if (!wide) {
    int blocks = (n + kTraceBlock - 1) / kTraceBlock;
    k_hook_ref<ANY><<<blocks, kTraceBlock, 0, st>>>(sc, n, io);
    return;
  }
  // scratch: [0] next, [1] fb_count, [2..] fb_list
  cudaMemsetAsync(scratch, 0, 8, st);
  WideQueue wq{scratch, scratch + 1, scratch + 2};
  k_hook_wide<ANY><<<persistent_grid(), kTraceBlock, 0, st>>>(sc, n, wq, io);
  k_hook_fallback<ANY><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, st>>>(sc, wq, io);
}

void launch_trace_closest(const DScene& sc, const float* rays, size_t n, int accel, int32_t* geom,
                          int32_t* prim, float* t, float* surf, int* scratch, cudaStream_t st) {
  if (!n) return;
  HookIO io{rays, geom, prim, t, surf, nullptr};
  launch_hook<false>(sc, int(n), accel, io, scratch, st);
}
void launch_trace_any(const DScene& sc, const float* rays, size_t n, int accel, uint8_t* occ, int* scratch,
                      cudaStream_t st) {
  if (!n) return;
  HookIO io{rays, nullptr, nullptr, nullptr, nullptr, occ};
  launch_hook<true>(sc, int(n), accel, io, scratch, st);
}

}  // namespace pb2
