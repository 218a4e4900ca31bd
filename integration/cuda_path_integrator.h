// integration/cuda_path_integrator.h -- the Pine-side shim a maintainer adds to the reference tree.
//
// `pine::CudaPathIntegrator` is a `pine::Integrator` (src/pine/core/integrator.h:20-25) with the
// constructor signatures of `pine::PathIntegrator` (src/pine/impl/integrator/path.h:7) whose
// `render(Scene&)` flattens the live `pine::Scene` into the POD records of include/pine_b200.h,
// calls the C ABI of libpine_b200.so (hand-written sm_100a kernels), and writes the float film
// back into `scene.camera.film()` -- the same observable effect as path.cpp:26-41.
//
// It compiles against the UNMODIFIED reference headers (this file and the .cpp are the only new
// sources on the Pine side) and is registered under the PRL name by swapping the lambda at
// src/pine/core/program_context.cpp:79-81. In this repository it is compiled
// into the test harness oracle/_ref/libpine_ref.so (the only place the reference headers exist)
// and exercised by tests/test_gpu_dropin.py: the SAME pine::Scene object is rendered by the
// reference's PathIntegrator and by this class, and the films are compared.
#pragma once
#include <pine/core/integrator.h>

struct pb2_multi;

namespace pine {

class CudaPathIntegrator : public Integrator {
 public:
  // Same arguments as PathIntegrator (path.h:7). `accel` and `light_sampler` are accepted for
  // signature compatibility: the CUDA backend builds its own BVH (bit-identical closest hits to
  // pine::BVH) and implements UniformLightSampler (the only LightSampler the reference has,
  // lightsampler.h:34).
  CudaPathIntegrator(Accel accel, Sampler sampler, LightSampler light_sampler, int max_path_length);
  // The PRL-level constructor (program_context.cpp:79-81).
  CudaPathIntegrator(Sampler sampler, int max_path_length);
  ~CudaPathIntegrator() override;
  CudaPathIntegrator(CudaPathIntegrator&&) noexcept;
  CudaPathIntegrator(const CudaPathIntegrator&) = delete;

  void render(Scene& scene) override;

  // Film::save's epilogue on the GPU (save_film_as_image, film.cpp:12-15: Film::finalize + gamma 1/2.2, x256,
  // clamp, optional y-flip, fileio.cpp:42-54) applied to the film the last render() left on the device, so only
  // 4 bytes per pixel cross PCIe. rgba8: width * height * 4 bytes. The tone mapper is read from `film` (it is
  // private there: a one-pixel copy of the Film is finalized on the host and compared with Uncharted2's value).
  void finalize_rgba8(const Film& film, unsigned char* rgba8, bool flip_y = true);

  // Which CUDA devices to use -- not part of the reference interface. Default: EVERY visible device (the scene
  // is replicated, interleaved 16-row bands of the film are dealt to the GPUs and ONE ncclReduce assembles the
  // film: bit-identical to a single-GPU render). set_device(d) pins the render to one device.
  void set_device(int device) { set_devices(&device, 1); }
  void set_devices(const int* devices, int n) {
    devices_.clear();
    for (int i = 0; i < n; i++) devices_.push_back(devices[i]);
  }
  // PB2_SHARD_ROWS (default, bit-identical to one GPU) or PB2_SHARD_SAMPLES (include/pine_b200.h).
  void set_shard_mode(int mode) { shard_mode_ = mode; }
  int n_devices() const { return n_devices_; }
  // Counters of the last render (rays traced, kernel launches, device ms); see pb2_counters.
  uint64_t last_closest_rays() const { return last_closest_; }
  uint64_t last_shadow_rays() const { return last_shadow_; }
  uint64_t kernel_launches() const { return launches_; }
  double last_render_ms() const { return last_ms_; }

 private:
  Sampler sampler_;
  int max_path_length_;
  psl::vector<int> devices_;  // empty: all visible
  int shard_mode_ = 1;        // PB2_SHARD_ROWS
  int n_devices_ = 0;
  pb2_multi* multi_ = nullptr;
  uint64_t last_closest_ = 0, last_shadow_ = 0, launches_ = 0;
  double last_ms_ = 0;
};

}  // namespace pine
