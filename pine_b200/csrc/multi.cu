// multi.cu -- one PathIntegrator::render over every GPU of the box, behind the C ABI (include/pine_b200.h,
// pb2_multi_*). BASELINE.json north_star: "the 8xB200 box is used by replicating the scene and sharding pixel
// tiles or sample passes across GPUs, with a single NCCL reduce of the float film over NVLink".
//
// One host thread calls in (as the reference's Integrator::render(Scene&) is called, integrator.h:20-25);
// inside, the BVH is built ONCE on the host, every GPU pulls 1/N of each large scene array through PCIe and
// the replicas are completed by in-place ncclAllGather over NVLink; a render runs one short-lived host thread
// per GPU (each drives its own context and streams), then ONE ncclReduce(sum) of the float film to the first
// GPU and one device->host copy. NCCL is bound at run time (dlopen "libnccl.so.2": whichever NCCL the process
// already has, else the system's) so that libpine_b200.so has no link-time dependency on it and a single-GPU
// caller never loads it.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <thread>
#include <atomic>
#include <memory>
#include <cstdlib>
#include <vector>

#include "context.h"
#include "kernels.cuh"

using namespace pb2;

namespace {

// the subset of nccl.h this file uses (NCCL 2.x ABI: ncclFloat32 = 7, ncclUint8 = 1, ncclSum = 0)
typedef struct ncclComm* ncclComm_t;
struct Nccl {
  void* lib = nullptr;
  int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool load(std::string& err) {
    if (lib) return true;
    // PB2_NCCL_LIB names a file; otherwise "libnccl.so.2" resolves to the NCCL the process has ALREADY loaded (same
    // SONAME: e.g. the one bundled with PyTorch, which pine_b200/integrator.py pre-loads so that a later `import
    // torch` finds its own version) or, failing that, to the system one.
    const char* forced = std::getenv("PB2_NCCL_LIB");
    if (forced && *forced) lib = dlopen(forced, RTLD_NOW | RTLD_LOCAL);
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      if (lib) break;
      lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
    }
    if (!lib) { err = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror(); return false; }
    auto sym = [&](const char* n) { return dlsym(lib, n); };
    CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
    CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
    GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
    GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
    Reduce = reinterpret_cast<decltype(Reduce)>(sym("ncclReduce"));
    AllGather = reinterpret_cast<decltype(AllGather)>(sym("ncclAllGather"));
    GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
    if (!CommInitAll || !CommDestroy || !GroupStart || !GroupEnd || !Reduce || !AllGather || !GetErrorString) {
      err = "libnccl.so.2 lacks a required symbol";
      return false;
    }
    return true;
  }
};
constexpr int kNcclUint8 = 1, kNcclFloat32 = 7, kNcclSum = 0;

}  // namespace

struct pb2_multi {
  std::vector<pb2_context*> ctx;  // ctx[0] owns the host scene; the others borrow it (host_view)
  std::vector<int> devices;
  Nccl nccl;
  std::vector<ncclComm_t> comms;
  bool scene_ready = false;
  double last_render_ms = 0, last_reduce_ms = 0, last_stage_ms = 0;
  struct Acc { uint64_t closest = 0, shadow = 0, fb_closest = 0, fb_shadow = 0; double ms = 0; };
  std::vector<Acc> acc;  // per GPU, summed over the render calls of the last pb2_multi_render
  // pb2_multi_set_progress_callback: every device reports its own fraction, the caller sees the mean
  struct Progress { pb2_multi* m; int i; };
  std::vector<Progress> progress_slot;
  std::unique_ptr<std::atomic<float>[]> progress_of;
  pb2_progress_fn progress_fn = nullptr;
  void* progress_user = nullptr;
};

static void device_progress(float fraction, void* user) {
  auto* slot = static_cast<pb2_multi::Progress*>(user);
  pb2_multi* m = slot->m;
  m->progress_of[slot->i].store(fraction, std::memory_order_relaxed);
  float sum = 0;
  for (size_t k = 0; k < m->ctx.size(); k++) sum += m->progress_of[k].load(std::memory_order_relaxed);
  if (m->progress_fn) m->progress_fn(sum / float(m->ctx.size()), m->progress_user);
}

#define PB2_NCCL_CHECK(m, expr)                                                                       \
  do {                                                                                                \
    int r_ = (expr);                                                                                  \
    if (r_ != 0) {                                                                                    \
      set_error(std::string(#expr) + ": " + (m)->nccl.GetErrorString(r_));                            \
      return PB2_E_CUDA;                                                                              \
    }                                                                                                 \
  } while (0)

extern "C" int pb2_create_multi(const int* devices, int n_devices, pb2_multi** out) {
  if (!out) { set_error("null out pointer"); return PB2_E_INVALID; }
  *out = nullptr;
  std::vector<int> devs;
  if (!devices || n_devices <= 0) {
    const int n = pb2_device_count();
    for (int i = 0; i < n; i++) devs.push_back(i);
  } else {
    devs.assign(devices, devices + n_devices);
  }
  if (devs.empty()) { set_error("no CUDA device visible; this backend has no CPU fallback"); return PB2_E_CUDA; }
  for (size_t i = 0; i < devs.size(); i++)
    for (size_t j = 0; j < i; j++)
      if (devs[i] == devs[j]) { set_error("pb2_create_multi: a device is listed twice"); return PB2_E_INVALID; }
  auto* m = new pb2_multi();
  m->devices = devs;
  for (int d : devs) {
    pb2_context* c = nullptr;
    int rc = pb2_create(d, &c);
    if (rc) { pb2_destroy_multi(m); return rc; }
    m->ctx.push_back(c);
  }
  if (devs.size() > 1) {
    std::string err;
    if (!m->nccl.load(err)) { set_error(err); pb2_destroy_multi(m); return PB2_E_CUDA; }
    m->comms.assign(devs.size(), nullptr);
    int r = m->nccl.CommInitAll(m->comms.data(), int(devs.size()), devs.data());
    if (r != 0) {
      set_error(std::string("ncclCommInitAll: ") + m->nccl.GetErrorString(r));
      m->comms.clear();
      pb2_destroy_multi(m);
      return PB2_E_CUDA;
    }
  }
  *out = m;
  return PB2_OK;
}

extern "C" void pb2_destroy_multi(pb2_multi* m) {
  if (!m) return;
  for (ncclComm_t c : m->comms)
    if (c) m->nccl.CommDestroy(c);
  // borrowers first: their host_view points into ctx[0]
  for (size_t i = m->ctx.size(); i-- > 0;) pb2_destroy(m->ctx[i]);
  delete m;
}

extern "C" int pb2_multi_size(pb2_multi* m) { return m ? int(m->ctx.size()) : 0; }
extern "C" pb2_context* pb2_multi_context(pb2_multi* m, int i) {
  return (m && i >= 0 && size_t(i) < m->ctx.size()) ? m->ctx[i] : nullptr;
}

extern "C" int pb2_multi_upload_sampler_tables(pb2_multi* m, int spp_pow2, const uint8_t* sobol,
                                               const uint8_t* scramble, const uint8_t* rank) {
  if (!m) { set_error("null argument"); return PB2_E_INVALID; }
  for (pb2_context* c : m->ctx) {
    int rc = pb2_upload_sampler_tables(c, spp_pow2, sobol, scramble, rank);
    if (rc) return rc;
  }
  return PB2_OK;
}
extern "C" int pb2_multi_load_sampler_tables_file(pb2_multi* m, const char* path, int spp_pow2) {
  if (!m) { set_error("null argument"); return PB2_E_INVALID; }
  for (pb2_context* c : m->ctx) {
    int rc = pb2_load_sampler_tables_file(c, path, spp_pow2);
    if (rc) return rc;
  }
  return PB2_OK;
}

// Stages the host scene of ctx[0] into every replica: GPU i copies part i of n of every large array from the
// pinned host copy (small arrays whole), then one in-place all-gather per array over NVLink.
static int stage_replicas(pb2_multi* m) {
  const int n = int(m->ctx.size());
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::vector<pb2_device_array>> arrays(n, std::vector<pb2_device_array>(64));
  std::vector<int> n_arrays(n, 0);
  for (int i = 0; i < n; i++) {
    int rc = pb2_restage_scene_part(m->ctx[i], nullptr, i, n, arrays[i].data(), 64, &n_arrays[i]);
    if (rc) return rc;
  }
  if (n > 1) {
    for (int a = 0; a < n_arrays[0]; a++) {
      PB2_NCCL_CHECK(m, m->nccl.GroupStart());
      for (int i = 0; i < n; i++) {
        const pb2_device_array& A = arrays[i][a];
        char* base = static_cast<char*>(A.device_ptr);
        int r = m->nccl.AllGather(base + size_t(i) * A.part_bytes, base, A.part_bytes, kNcclUint8, m->comms[i],
                                  m->ctx[i]->stream);
        if (r != 0) { m->nccl.GroupEnd(); set_error(std::string("ncclAllGather: ") + m->nccl.GetErrorString(r)); return PB2_E_CUDA; }
      }
      PB2_NCCL_CHECK(m, m->nccl.GroupEnd());
    }
  }
  for (int i = 0; i < n; i++) {
    PB2_CUDA_CHECK(cudaSetDevice(m->devices[i]));
    PB2_CUDA_CHECK(cudaStreamSynchronize(m->ctx[i]->stream));
    m->ctx[i]->scene_ready = true;
  }
  m->last_stage_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return PB2_OK;
}

extern "C" int pb2_multi_upload_scene(pb2_multi* m, const pb2_scene_desc* sd) {
  if (!m || !sd) { set_error("null argument"); return PB2_E_INVALID; }
  m->scene_ready = false;
  if (m->ctx.size() == 1) {
    int rc = pb2_upload_scene(m->ctx[0], sd);
    m->scene_ready = rc == PB2_OK;
    return rc;
  }
  pb2_context* c0 = m->ctx[0];
  c0->defer_staging = true;
  int rc = pb2_upload_scene(c0, sd);  // validation + host BVH build + layout + page-lock, ONCE
  c0->defer_staging = false;
  if (rc) return rc;
  for (size_t i = 1; i < m->ctx.size(); i++) {
    pb2_context* c = m->ctx[i];
    c->scene_ready = false;
    c->scene = c0->scene;  // tree roots, counts, camera, flags; device pointers are rebound by the staging
    c->only_area_lights = c0->only_area_lights;
    c->material_kinds = c0->material_kinds;
    c->needs_rng_in_shading = c0->needs_rng_in_shading;
    c->has_node_programs = c0->has_node_programs;
    c->n_pixels = c0->n_pixels;
    c->counters.bvh_build_ms = c0->counters.bvh_build_ms;
    c->counters.scene_bytes = c0->counters.scene_bytes;
    c->counters.bvh2_nodes = c0->counters.bvh2_nodes;
    c->counters.wide_nodes = c0->counters.wide_nodes;
    c->counters.n_primitives = c0->counters.n_primitives;
    c->counters.build_threads = c0->counters.build_threads;
    c->host_view = &c0->host;
    c->host_scene_ready = true;
  }
  rc = stage_replicas(m);
  m->scene_ready = rc == PB2_OK;
  return rc;
}

extern "C" int pb2_multi_restage_scene(pb2_multi* m) {
  if (!m || !m->scene_ready) { set_error("pb2_multi_upload_scene has not succeeded"); return PB2_E_STATE; }
  if (m->ctx.size() == 1) {
    int rc = pb2_restage_scene(m->ctx[0], nullptr);
    if (rc) return rc;
    PB2_CUDA_CHECK(cudaStreamSynchronize(m->ctx[0]->stream));
    return PB2_OK;
  }
  return stage_replicas(m);
}

// Renders the shares, reduces the film to device 0 (film of every context: zero where it rendered nothing,
// so shares ADD to the full film -- bit-identically for row bands) and leaves it there.
static int render_and_reduce(pb2_multi* m, const pb2_render_params* p, int shard_mode, float** film_dev0) {
  const int n = int(m->ctx.size());
  const int H = m->ctx[0]->scene.camera.height, W = m->ctx[0]->scene.camera.width;
  struct Call { pb2_render_params p; };
  std::vector<std::vector<Call>> plan(n);
  if (shard_mode == PB2_SHARD_SAMPLES) {
    const int total = p->sample_end - p->sample_begin, base = total / n, extra = total % n;
    int b = p->sample_begin;
    for (int i = 0; i < n; i++) {
      const int cnt = base + (i < extra ? 1 : 0);
      if (cnt > 0) {
        Call c{*p};
        c.p.sample_begin = b;
        c.p.sample_end = b + cnt;
        plan[i].push_back(c);
      }
      b += cnt;
    }
  } else if (shard_mode == PB2_SHARD_ROWS) {
    const int n_bands = (p->row_end - p->row_begin + PB2_SHARD_ROW_BAND - 1) / PB2_SHARD_ROW_BAND;
    for (int i = 0; i < n && i < n_bands; i++) {
      Call c{*p};
      c.p.row_band = PB2_SHARD_ROW_BAND;
      c.p.row_parts = n;
      c.p.row_part = i;
      plan[i].push_back(c);
    }
  } else {
    set_error("unknown shard mode");
    return PB2_E_INVALID;
  }
  std::vector<int> rcs(n, PB2_OK);
  std::vector<std::string> errs(n);
  std::vector<float*> films(n, nullptr);
  m->acc.assign(n, pb2_multi::Acc());
  auto work = [&](int i) {
    pb2_context* c = m->ctx[i];
    if (plan[i].empty()) {  // more GPUs than shares: contribute a zero film
      pb2_render_params z = *p;
      z.row_end = z.row_begin;
      rcs[i] = pb2_render_device(c, &z, nullptr, &films[i]);
    }
    for (const Call& call : plan[i]) {
      rcs[i] = pb2_render_device(c, &call.p, nullptr, &films[i]);
      if (rcs[i]) break;
      pb2_multi::Acc& a = m->acc[i];
      a.closest += c->counters.last_closest_rays; a.shadow += c->counters.last_shadow_rays;
      a.fb_closest += c->counters.last_closest_fallbacks; a.fb_shadow += c->counters.last_shadow_fallbacks;
      a.ms += c->counters.last_render_ms;
    }
    if (rcs[i]) errs[i] = pb2_last_error();
  };
  auto t0 = std::chrono::steady_clock::now();
  {
    std::vector<std::thread> pool;
    for (int i = 1; i < n; i++) pool.emplace_back(work, i);
    work(0);
    for (auto& t : pool) t.join();
  }
  for (int i = 0; i < n; i++)
    if (rcs[i]) { set_error("GPU " + std::to_string(m->devices[i]) + ": " + errs[i]); return rcs[i]; }
  m->last_render_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  auto t1 = std::chrono::steady_clock::now();
  if (n > 1) {
    PB2_NCCL_CHECK(m, m->nccl.GroupStart());
    for (int i = 0; i < n; i++) {
      int r = m->nccl.Reduce(films[i], films[i], size_t(W) * H * 4, kNcclFloat32, kNcclSum, 0, m->comms[i], m->ctx[i]->stream);
      if (r != 0) { m->nccl.GroupEnd(); set_error(std::string("ncclReduce: ") + m->nccl.GetErrorString(r)); return PB2_E_CUDA; }
    }
    PB2_NCCL_CHECK(m, m->nccl.GroupEnd());
    for (int i = n - 1; i >= 0; i--) {
      PB2_CUDA_CHECK(cudaSetDevice(m->devices[i]));
      PB2_CUDA_CHECK(cudaStreamSynchronize(m->ctx[i]->stream));
    }
  }
  m->last_reduce_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
  *film_dev0 = films[0];
  return PB2_OK;
}

extern "C" int pb2_multi_render(pb2_multi* m, const pb2_render_params* p, int shard_mode, float* film_rgba_host) {
  if (!m || !p || !film_rgba_host) { set_error("null argument"); return PB2_E_INVALID; }
  if (!m->scene_ready) { set_error("pb2_multi_upload_scene has not succeeded"); return PB2_E_STATE; }
  if (m->progress_of)
    for (size_t i = 0; i < m->ctx.size(); i++) m->progress_of[i].store(0.0f);
  if (m->ctx.size() == 1) return pb2_render(m->ctx[0], p, film_rgba_host);
  float* dev = nullptr;
  int rc = render_and_reduce(m, p, shard_mode, &dev);
  if (rc) return rc;
  pb2_context* c0 = m->ctx[0];
  PB2_CUDA_CHECK(cudaSetDevice(m->devices[0]));
  const size_t bytes = size_t(c0->scene.camera.width) * c0->scene.camera.height * 16;
  PB2_CUDA_CHECK(cudaMemcpyAsync(film_rgba_host, dev, bytes, cudaMemcpyDeviceToHost, c0->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(c0->stream));
  return PB2_OK;
}

extern "C" int pb2_multi_set_progress_callback(pb2_multi* m, pb2_progress_fn fn, void* user) {
  if (!m) { set_error("null argument"); return PB2_E_INVALID; }
  m->progress_fn = fn;
  m->progress_user = user;
  if (!m->progress_of) {
    m->progress_of.reset(new std::atomic<float>[m->ctx.size()]);
    m->progress_slot.resize(m->ctx.size());
    for (size_t i = 0; i < m->ctx.size(); i++) m->progress_slot[i] = pb2_multi::Progress{m, int(i)};
  }
  for (size_t i = 0; i < m->ctx.size(); i++) {
    m->progress_of[i].store(0.0f);
    int rc = pb2_set_progress_callback(m->ctx[i], fn ? device_progress : nullptr, fn ? &m->progress_slot[i] : nullptr);
    if (rc) return rc;
  }
  return PB2_OK;
}

extern "C" int pb2_multi_film_finalize(pb2_multi* m, int tonemapper, int flip_y, uint8_t* rgba8_host,
                                       float* tonemapped_rgba_host) {
  if (!m || !m->scene_ready) { set_error("pb2_multi_upload_scene has not succeeded"); return PB2_E_STATE; }
  // after pb2_multi_render the first device holds the reduced film
  return pb2_film_finalize_device(m->ctx[0], tonemapper, flip_y, rgba8_host, tonemapped_rgba_host);
}

extern "C" int pb2_multi_get_counters(pb2_multi* m, pb2_counters* out, double* reduce_ms, double* stage_ms) {
  if (!m || !out) { set_error("null argument"); return PB2_E_INVALID; }
  *out = m->ctx[0]->counters;
  for (size_t i = 1; i < m->ctx.size(); i++) {
    const pb2_counters& c = m->ctx[i]->counters;
    out->closest_rays += c.closest_rays; out->shadow_rays += c.shadow_rays;
    out->camera_samples += c.camera_samples; out->kernel_launches += c.kernel_launches;
  }
  if (m->ctx.size() > 1 && !m->acc.empty()) {  // last_*: the whole last multi-GPU render, device time = slowest GPU
    out->last_closest_rays = out->last_shadow_rays = out->last_closest_fallbacks = out->last_shadow_fallbacks = 0;
    out->last_render_ms = 0;
    for (const pb2_multi::Acc& a : m->acc) {
      out->last_closest_rays += a.closest; out->last_shadow_rays += a.shadow;
      out->last_closest_fallbacks += a.fb_closest; out->last_shadow_fallbacks += a.fb_shadow;
      out->last_render_ms = std::max(out->last_render_ms, a.ms);
    }
  }
  if (reduce_ms) *reduce_ms = m->last_reduce_ms;
  if (stage_ms) *stage_ms = m->last_stage_ms;
  return PB2_OK;
}
