// trace_kernels.cu -- Accel::intersect / Accel::hit (reference src/pine/core/accel.h:16-24) over a
// caller-supplied ray set: the test hooks pb2_trace_closest / pb2_trace_any and the traversal
// micro-benchmark. Same persistent wide walk + fallback as the renderer (wide_persistent.cuh).
#include <mutex>

#include "context.h"
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

template <bool ANY, bool RARE>
__global__ void __launch_bounds__(kTraceBlock, PB2_TRACE_MINBLOCKS) k_hook_wide(DScene sc, int n, WideQueue wq, HookIO io) {
  wide_persistent<ANY, (RARE ? kShapesAll : kShapesCommon)>(sc, n, wq, io);
}
template <bool ANY, bool RARE>
__global__ void __launch_bounds__(kTraceBlock) k_hook_fallback(DScene sc, WideQueue wq, HookIO io) {
  wide_fallback<ANY, (RARE ? kShapesAll : kShapesCommon)>(sc, wq, io);
}
template <bool ANY, bool RARE>
__global__ void __launch_bounds__(kTraceBlock) k_hook_ref(DScene sc, int n, HookIO io) {
  ref_order_all<ANY, (RARE ? kShapesAll : kShapesCommon)>(sc, n, io);
}

template <bool ANY>
static void launch_hook(const DScene& sc, int n, int accel, const HookIO& io, int* scratch, cudaStream_t st) {
  const bool wide = accel == 0 && sc.wide_ready &&
                    (ANY ? sc.n_box_shapes <= kMaxBoxShapes : !sc.order_dependent);
  if (!wide) {
    int blocks = (n + kTraceBlock - 1) / kTraceBlock;
    if (sc.has_rare_shapes) k_hook_ref<ANY, true><<<blocks, kTraceBlock, 0, st>>>(sc, n, io);
    else k_hook_ref<ANY, false><<<blocks, kTraceBlock, 0, st>>>(sc, n, io);
    return;
  }
  // scratch: [0] next, [1] fb_count, [2..] fb_list
  cudaMemsetAsync(scratch, 0, 8, st);
  WideQueue wq{scratch, scratch + 1, scratch + 2};
  if (sc.has_rare_shapes) k_hook_wide<ANY, true><<<persistent_grid(), kTraceBlock, 0, st>>>(sc, n, wq, io);
  else k_hook_wide<ANY, false><<<persistent_grid(), kTraceBlock, 0, st>>>(sc, n, wq, io);
  if (sc.has_rare_shapes) k_hook_fallback<ANY, true><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, st>>>(sc, wq, io);
  else k_hook_fallback<ANY, false><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, st>>>(sc, wq, io);
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

// ---- pb2_trace_closest_device_noalloc: the same closest-hit launch with its scratch in static device memory ----
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). pb2_trace_closest_device
// reserves `trace_scratch` (claim counter, fallback count, fallback list: (n + 2) ints) with cudaMalloc when n
// grows; here those live in a file-scope array that every GPU loading the library holds until exit
// (kNoAllocScratchInts * 4 = 4 MiB + 256 B per device). All calls on one device share it: the two counters are
// cleared on the launch stream before every launch, the list needs no clearing, and calls that OVERLAP on the
// GPU (different streams, no ordering between them) race on it.
namespace pb2 {
constexpr int kNoAllocListOffset = 64;  // counters in the first 256-byte line, the fallback list on the next
constexpr size_t kNoAllocScratchInts = kNoAllocListOffset + size_t(PB2_NOALLOC_MAX_RAYS);
static __device__ __align__(256) int g_noalloc_scratch[kNoAllocScratchInts];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: each device's address is
// looked up once (cudaGetSymbolAddress with that device current) and cached.
constexpr int kNoAllocMaxDevices = 64;
static int* g_noalloc_ptr[kNoAllocMaxDevices];
static std::mutex g_noalloc_mutex;

static int noalloc_scratch_of(int device, int** out) {
  if (device < 0 || device >= kNoAllocMaxDevices) {
    set_error("pb2_trace_closest_device_noalloc: device index " + std::to_string(device) + " outside [0, " +
              std::to_string(kNoAllocMaxDevices) + ")");
    return PB2_E_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    int* ptr = nullptr;
    PB2_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch));
    if (!ptr) { set_error("cudaGetSymbolAddress returned a null scratch address"); return PB2_E_CUDA; }
    g_noalloc_ptr[device] = ptr;
  }
  *out = g_noalloc_ptr[device];
  return PB2_OK;
}
}  // namespace pb2

using namespace pb2;

extern "C" int pb2_trace_closest_device_noalloc(pb2_context* ctx, const float* rays_dev, size_t n, int accel,
                                                int32_t* geom_dev, int32_t* prim_dev, float* t_dev,
                                                void* cuda_stream) {
  if (!ctx) { set_error("null context"); return PB2_E_INVALID; }
  if (ctx->device < 0) { set_error("host-only context: no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  if (!ctx->scene_ready) { set_error("pb2_upload_scene has not succeeded on this context"); return PB2_E_STATE; }
  if (n > size_t(PB2_NOALLOC_MAX_RAYS)) {
    set_error("pb2_trace_closest_device_noalloc: n = " + std::to_string(n) + " rays exceeds PB2_NOALLOC_MAX_RAYS = " +
              std::to_string(PB2_NOALLOC_MAX_RAYS) + " (the static scratch is fixed; split the batch or use "
              "pb2_trace_closest_device)");
    return PB2_E_INVALID;
  }
  if (n && (!rays_dev || !geom_dev || !prim_dev || !t_dev)) { set_error("null ray or output pointer"); return PB2_E_INVALID; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  int* scratch = nullptr;
  int rc = noalloc_scratch_of(ctx->device, &scratch);  // n == 0: set-up only (first call per device, outside capture)
  if (rc || n == 0) return rc;
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  const DScene& sc = ctx->scene;
  HookIO io{rays_dev, geom_dev, prim_dev, t_dev, nullptr, nullptr};
  const bool wide = accel == 0 && sc.wide_ready && !sc.order_dependent;
  if (!wide) {  // reference-order walk: no scratch at all
    const int blocks = (int(n) + kTraceBlock - 1) / kTraceBlock;
    if (sc.has_rare_shapes) k_hook_ref<false, true><<<blocks, kTraceBlock, 0, st>>>(sc, int(n), io);
    else k_hook_ref<false, false><<<blocks, kTraceBlock, 0, st>>>(sc, int(n), io);
    ctx->counters.kernel_launches++;
  } else {
    PB2_CUDA_CHECK(cudaMemsetAsync(scratch, 0, 2 * sizeof(int), st));  // claim counter and fallback count start at zero
    WideQueue wq{scratch, scratch + 1, scratch + kNoAllocListOffset};
    if (sc.has_rare_shapes) k_hook_wide<false, true><<<persistent_grid(), kTraceBlock, 0, st>>>(sc, int(n), wq, io);
    else k_hook_wide<false, false><<<persistent_grid(), kTraceBlock, 0, st>>>(sc, int(n), wq, io);
    if (sc.has_rare_shapes) k_hook_fallback<false, true><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, st>>>(sc, wq, io);
    else k_hook_fallback<false, false><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, st>>>(sc, wq, io);
    ctx->counters.kernel_launches += 2;
  }
  PB2_CUDA_CHECK(cudaGetLastError());
  return PB2_OK;
}
