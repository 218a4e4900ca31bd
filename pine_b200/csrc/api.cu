// api.cu -- context lifetime, sampler tables, and the traversal test hooks of include/pine_b200.h.
#include <dlfcn.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "context.h"
#include "kernels.cuh"

using namespace pb2;

namespace pb2 {
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
}  // namespace pb2

extern "C" const char* pb2_last_error(void) { return g_error.c_str(); }

extern "C" int pb2_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" int pb2_create(int device, pb2_context** out) {
  if (!out) { set_error("null out pointer"); return PB2_E_INVALID; }
  *out = nullptr;
  auto* ctx = new pb2_context();
  ctx->device = device;
  if (device >= 0) {
    int n = pb2_device_count();
    if (device >= n) {
      delete ctx;
      set_error("CUDA device " + std::to_string(device) + " not available (" + std::to_string(n) +
                " visible); this backend has no CPU fallback");
      return PB2_E_CUDA;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev2);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev3);
    if (e != cudaSuccess) {
      set_error(std::string("pb2_create: ") + cudaGetErrorString(e));
      delete ctx;
      return PB2_E_CUDA;
    }
  }
  *out = ctx;
  return PB2_OK;
}

extern "C" void pb2_destroy(pb2_context* ctx) {
  if (!ctx) return;
  if (ctx->device >= 0) {
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    free_wavefront(ctx);
    ctx->host.unpin();
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->ev2) cudaEventDestroy(ctx->ev2);
    if (ctx->ev3) cudaEventDestroy(ctx->ev3);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
  }
  ctx->host.unmap();
  delete ctx;
}

static int log2_exact(int v) {
  for (int i = 0; i < 9; i++)
    if ((1 << i) == v) return i;
  return -1;
}

extern "C" int pb2_upload_sampler_tables(pb2_context* ctx, int spp_pow2, const uint8_t* sobol,
                                         const uint8_t* scramble, const uint8_t* rank) {
  if (!ctx || !sobol || !scramble || !rank) { set_error("null argument"); return PB2_E_INVALID; }
  int k = log2_exact(spp_pow2);
  if (k < 0) { set_error("sampler tables exist for spp in {1,2,4,...,256} only"); return PB2_E_INVALID; }
  ctx->tables[k].sobol.assign(sobol, sobol + 256 * 256);
  ctx->tables[k].scramble.assign(scramble, scramble + 128 * 128 * 8);
  ctx->tables[k].rank.assign(rank, rank + 128 * 128 * 8);
  ctx->has_tables[k] = true;
  if (ctx->tables_spp == spp_pow2) ctx->tables_spp = 0;  // force re-upload
  return PB2_OK;
}

// Reads one spp variant from the table blob pine_b200/tools/extract_tables.cpp writes ("PB2BLUE1", u32 count,
// then per variant u32 spp, u32 pad, sobol[65536], scramble[131072], rank[131072]) and registers it.
// path NULL/"" = pine_b200/data/bluesobol_tables.bin next to this library.
extern "C" int pb2_load_sampler_tables_file(pb2_context* ctx, const char* path, int spp_pow2) {
  if (!ctx) { set_error("null argument"); return PB2_E_INVALID; }
  int k = log2_exact(spp_pow2);
  if (k < 0) { set_error("sampler tables exist for spp in {1,2,4,...,256} only"); return PB2_E_INVALID; }
  if (ctx->has_tables[k]) return PB2_OK;
  std::string file = path ? path : "";
  if (file.empty()) {
    Dl_info info;
    if (!dladdr(reinterpret_cast<const void*>(&pb2_load_sampler_tables_file), &info) || !info.dli_fname) {
      set_error("cannot locate libpine_b200.so to find data/bluesobol_tables.bin");
      return PB2_E_STATE;
    }
    file = info.dli_fname;
    size_t slash = file.rfind('/');
    file = (slash == std::string::npos ? std::string(".") : file.substr(0, slash)) + "/../data/bluesobol_tables.bin";
  }
  FILE* f = std::fopen(file.c_str(), "rb");
  if (!f) { set_error("cannot open sampler table file " + file); return PB2_E_INVALID; }
  char magic[8];
  uint32_t n = 0, pad = 0;
  bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "PB2BLUE1", 8) == 0 &&
            std::fread(&n, 4, 1, f) == 1 && std::fread(&pad, 4, 1, f) == 1;
  int rc = PB2_E_INVALID;
  std::vector<uint8_t> buf(65536 + 2 * 131072);
  for (uint32_t i = 0; ok && i < n; i++) {
    uint32_t spp = 0;
    if (std::fread(&spp, 4, 1, f) != 1 || std::fread(&pad, 4, 1, f) != 1 ||
        std::fread(buf.data(), 1, buf.size(), f) != buf.size()) break;
    if (int(spp) == spp_pow2) {
      rc = pb2_upload_sampler_tables(ctx, spp_pow2, buf.data(), buf.data() + 65536, buf.data() + 65536 + 131072);
      break;
    }
  }
  std::fclose(f);
  if (rc == PB2_E_INVALID) set_error("sampler table file " + file + " is malformed or lacks spp=" + std::to_string(spp_pow2));
  return rc;
}

// Makes the tables of `spp` the active device tables.
int pb2::activate_tables(pb2_context* ctx, int spp) {
  int k = log2_exact(spp);
  if (k < 0 || !ctx->has_tables[k]) {
    set_error("no sampler tables uploaded for spp=" + std::to_string(spp));
    return PB2_E_STATE;
  }
  if (ctx->tables_spp != spp) {
    PB2_CUDA_CHECK(ctx->tab_sobol.upload(ctx->tables[k].sobol));
    PB2_CUDA_CHECK(ctx->tab_scramble.upload(ctx->tables[k].scramble));
    PB2_CUDA_CHECK(ctx->tab_rank.upload(ctx->tables[k].rank));
    ctx->tables_spp = spp;
  }
  ctx->scene.sobol = ctx->tab_sobol.as<uint8_t>();
  ctx->scene.scramble = ctx->tab_scramble.as<uint8_t>();
  ctx->scene.rank = ctx->tab_rank.as<uint8_t>();
  return PB2_OK;
}

static int require_device_scene(pb2_context* ctx) {
  if (!ctx) { set_error("null context"); return PB2_E_INVALID; }
  if (ctx->device < 0) { set_error("host-only context: no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  if (!ctx->scene_ready) { set_error("pb2_upload_scene has not succeeded on this context"); return PB2_E_STATE; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  return PB2_OK;
}

extern "C" int pb2_trace_closest_device(pb2_context* ctx, const float* rays_dev, size_t n, int accel,
                                        int32_t* geom_dev, int32_t* prim_dev, float* t_dev,
                                        void* cuda_stream) {
  int rc = require_device_scene(ctx);
  if (rc) return rc;
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  PB2_CUDA_CHECK(ctx->trace_scratch.reserve((n + 2) * 4));
  launch_trace_closest(ctx->scene, rays_dev, n, accel, geom_dev, prim_dev, t_dev, nullptr,
                       ctx->trace_scratch.as<int>(), st);
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  return PB2_OK;
}

extern "C" int pb2_trace_closest(pb2_context* ctx, const float* rays, size_t n, int accel,
                                 int32_t* geom, int32_t* prim, float* t, float* surf) {
  int rc = require_device_scene(ctx);
  if (rc) return rc;
  if (n == 0) return PB2_OK;
  PB2_CUDA_CHECK(ctx->rays.reserve(n * 32));
  PB2_CUDA_CHECK(ctx->out_geom.reserve(n * 4));
  PB2_CUDA_CHECK(ctx->out_prim.reserve(n * 4));
  PB2_CUDA_CHECK(ctx->out_t.reserve(n * 4));
  if (surf) PB2_CUDA_CHECK(ctx->out_surf.reserve(n * 32));
  PB2_CUDA_CHECK(ctx->trace_scratch.reserve((n + 2) * 4));
  cudaStream_t st = ctx->stream;
  PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->rays.ptr, rays, n * 32, cudaMemcpyHostToDevice, st));
  launch_trace_closest(ctx->scene, ctx->rays.as<float>(), n, accel, ctx->out_geom.as<int32_t>(),
                       ctx->out_prim.as<int32_t>(), ctx->out_t.as<float>(),
                       surf ? ctx->out_surf.as<float>() : nullptr, ctx->trace_scratch.as<int>(), st);
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaMemcpyAsync(geom, ctx->out_geom.ptr, n * 4, cudaMemcpyDeviceToHost, st));
  PB2_CUDA_CHECK(cudaMemcpyAsync(prim, ctx->out_prim.ptr, n * 4, cudaMemcpyDeviceToHost, st));
  PB2_CUDA_CHECK(cudaMemcpyAsync(t, ctx->out_t.ptr, n * 4, cudaMemcpyDeviceToHost, st));
  if (surf) PB2_CUDA_CHECK(cudaMemcpyAsync(surf, ctx->out_surf.ptr, n * 32, cudaMemcpyDeviceToHost, st));
  PB2_CUDA_CHECK(cudaStreamSynchronize(st));
  return PB2_OK;
}

extern "C" int pb2_trace_any(pb2_context* ctx, const float* rays, size_t n, int accel,
                             uint8_t* occluded) {
  int rc = require_device_scene(ctx);
  if (rc) return rc;
  if (n == 0) return PB2_OK;
  PB2_CUDA_CHECK(ctx->rays.reserve(n * 32));
  PB2_CUDA_CHECK(ctx->out_occ.reserve(n));
  PB2_CUDA_CHECK(ctx->trace_scratch.reserve((n + 2) * 4));
  cudaStream_t st = ctx->stream;
  PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->rays.ptr, rays, n * 32, cudaMemcpyHostToDevice, st));
  launch_trace_any(ctx->scene, ctx->rays.as<float>(), n, accel, ctx->out_occ.as<uint8_t>(),
                   ctx->trace_scratch.as<int>(), st);
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaMemcpyAsync(occluded, ctx->out_occ.ptr, n, cudaMemcpyDeviceToHost, st));
  PB2_CUDA_CHECK(cudaStreamSynchronize(st));
  return PB2_OK;
}

extern "C" int pb2_set_ray_record(pb2_context* ctx, pb2_ray_record* rec) {
  if (!ctx) { set_error("null argument"); return PB2_E_INVALID; }
  ctx->record = rec;
  if (rec) rec->n_closest = rec->n_shadow = 0;
  return PB2_OK;
}

// Debug hook of -DPB2_COUNT builds (scripts/gpu_variants.sh): wide-walk work counters since the last
// call: {closest node tests, closest primitive tests, any-hit node tests, any-hit primitive tests,
// closest rays, any-hit rays, closest fallbacks by reason: box shape touched, tie/stack during the walk,
// tie with the winner, winner's leaf gate failed} = 10 values. In the product build the kernels do not count and this returns zeros.
extern "C" int pb2_debug_walk_stats(pb2_context* ctx, uint64_t* out10) {
  if (!ctx || !out10) { set_error("null argument"); return PB2_E_INVALID; }
  for (int i = 0; i < 10; i++) out10[i] = 0;
  if (!ctx->walk_stats.ptr) return PB2_OK;
  PB2_CUDA_CHECK(cudaMemcpy(out10, ctx->walk_stats.ptr, 80, cudaMemcpyDeviceToHost));
  PB2_CUDA_CHECK(cudaMemset(ctx->walk_stats.ptr, 0, 80));
  return PB2_OK;
}

// ---- read-bandwidth microbenchmark ("L2 peak to be measured with a resident-buffer read") ------
namespace pb2 {
__global__ void __launch_bounds__(256) k_read_bw(const uint4* __restrict__ buf, size_t n_vec, int iters, uint4* sink) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, stride = size_t(gridDim.x) * blockDim.x;
  for (int it = 0; it < iters; it++) {
    size_t i = tid;
    for (; i + 3 * stride < n_vec; i += 4 * stride) {  // four independent 128-bit loads in flight per thread
      const uint4 a = __ldcg(buf + i), b = __ldcg(buf + i + stride), c = __ldcg(buf + i + 2 * stride),
                  d = __ldcg(buf + i + 3 * stride);  // ld.global.cg: cached in L2 only, so L1 hits cannot inflate it
      acc.x ^= a.x ^ b.x ^ c.x ^ d.x; acc.y ^= a.y ^ b.y ^ c.y ^ d.y;
      acc.z ^= a.z ^ b.z ^ c.z ^ d.z; acc.w ^= a.w ^ b.w ^ c.w ^ d.w;
    }
    for (; i < n_vec; i += stride) { const uint4 a = __ldcg(buf + i); acc.x ^= a.x; acc.y ^= a.y; acc.z ^= a.z; acc.w ^= a.w; }
  }
  if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9e3779b9u) sink[tid % 32] = acc;  // keeps the loads alive
}
}  // namespace pb2

// Sustained read bandwidth of a `bytes`-sized device buffer read `iters` times back to back by a full grid
// (after one untimed pass): bytes well below the 126 MB L2 measure the L2, bytes far above it the HBM.
extern "C" int pb2_measure_read_gbs(pb2_context* ctx, size_t bytes, int iters, double* gbs) {
  if (!ctx || !gbs || bytes < 4096 || iters <= 0) { set_error("invalid argument"); return PB2_E_INVALID; }
  if (ctx->device < 0) { set_error("host-only context: no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  DBuf buf, sink;
  PB2_CUDA_CHECK(buf.reserve(bytes));
  PB2_CUDA_CHECK(sink.reserve(32 * 16));
  PB2_CUDA_CHECK(cudaMemsetAsync(buf.ptr, 1, bytes, ctx->stream));
  const size_t n_vec = bytes / 16;
  const int grid = kNumSMs * 8;
  k_read_bw<<<grid, 256, 0, ctx->stream>>>(buf.as<uint4>(), n_vec, 1, sink.as<uint4>());
  PB2_CUDA_CHECK(cudaEventRecord(ctx->ev2, ctx->stream));
  k_read_bw<<<grid, 256, 0, ctx->stream>>>(buf.as<uint4>(), n_vec, iters, sink.as<uint4>());
  PB2_CUDA_CHECK(cudaEventRecord(ctx->ev3, ctx->stream));
  ctx->counters.kernel_launches += 2;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  float ms = 0;
  PB2_CUDA_CHECK(cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3));
  *gbs = double(n_vec) * 16.0 * iters / (double(ms) * 1e-3) / 1e9;
  return PB2_OK;
}

extern "C" int pb2_get_counters(pb2_context* ctx, pb2_counters* out) {
  if (!ctx || !out) { set_error("null argument"); return PB2_E_INVALID; }
  *out = ctx->counters;
  return PB2_OK;
}
