// oracle/shim_harness.cpp -- TEST INFRASTRUCTURE ONLY.
//
// The drop-in test's C entry points: render the live pine::Scene of an oracle/ref_harness.cpp handle through
// pine::CudaPathIntegrator (integration/cuda_path_integrator.cpp, the Pine-side shim a maintainer adds),
// i.e. through `Integrator::render(Scene&)` (src/pine/core/integrator.h:20-25) into the product's C ABI.
// Built into oracle/_ref/libpine_shim.so, which links libpine_ref.so (Pine's own classes) and
// libpine_b200.so (the product); libpine_ref.so itself -- the checker -- depends on neither.
#include <pine/core/scene.h>
#include <pine/core/integrator.h>
#include <pine/core/lightsampler.h>
#include <pine/core/sampler.h>
#include <pine/impl/accel/bvh.h>

#include "../integration/cuda_path_integrator.h"

#include <chrono>
#include <cstring>

using namespace pine;

extern "C" {
void* ref_scene_ptr(void* h);  // oracle/ref_harness.cpp

// The drop-in: the SAME pine::Scene rendered by pine::CudaPathIntegrator (integration/), constructed
// with PathIntegrator's own argument list. Returns wall seconds; stats = {closest rays, shadow rays,
// kernel launches, device ms, GPUs used}; device < 0 = every visible GPU. Aborts through SEVERE like the reference on unsupported content.
double ref_render_cuda_kind(void* h, int sampler_kind, int spp, int depth, int device, float* film_rgba,
                            double* stats) {
  Scene& scene = *(Scene*)ref_scene_ptr(h);
  auto sampler = sampler_kind == 0   ? Sampler(BlueSobolSampler(spp))
                 : sampler_kind == 1 ? Sampler(SobolSampler(spp))
                                     : sampler_kind == 3 ? Sampler(HaltonSampler(spp)) : Sampler(UniformSampler(spp));
  auto integrator = CudaPathIntegrator(Accel(BVH()), sampler, LightSampler(UniformLightSampler()), depth);
  if (device >= 0) integrator.set_device(device);  // device < 0: every visible GPU (the shim's default)
  auto t0 = std::chrono::steady_clock::now();
  Integrator& base = integrator;  // through the virtual interface, as the PRL binding calls it
  base.render(scene);
  auto t1 = std::chrono::steady_clock::now();
  if (film_rgba) {
    auto& film = scene.camera.film();
    std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  }
  if (stats) {
    stats[0] = double(integrator.last_closest_rays());
    stats[1] = double(integrator.last_shadow_rays());
    stats[2] = double(integrator.kernel_launches());
    stats[3] = integrator.last_render_ms();
    stats[4] = double(integrator.n_devices());
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
// render through the shim, then the shim's device epilogue: rgba8 = w * h * 4 bytes, y-flipped like
// save_film_as_image (film.cpp:12-15)
double ref_render_cuda_rgba8(void* h, int spp, int depth, int device, float* film_rgba, unsigned char* rgba8) {
  Scene& scene = *(Scene*)ref_scene_ptr(h);
  auto integrator = CudaPathIntegrator(Accel(BVH()), Sampler(BlueSobolSampler(spp)), LightSampler(UniformLightSampler()), depth);
  if (device >= 0) integrator.set_device(device);
  auto t0 = std::chrono::steady_clock::now();
  integrator.render(scene);
  auto& film = scene.camera.film();
  integrator.finalize_rgba8(film, rgba8, true);
  auto t1 = std::chrono::steady_clock::now();
  if (film_rgba) std::memcpy(film_rgba, film.data(), sizeof(float) * 4 * size_t(film.width()) * film.height());
  return std::chrono::duration<double>(t1 - t0).count();
}
double ref_render_cuda(void* h, int spp, int depth, int device, float* film_rgba, double* stats) {
  return ref_render_cuda_kind(h, 0, spp, depth, device, film_rgba, stats);
}
}  // extern "C"
