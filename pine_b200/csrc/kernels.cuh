// kernels.cuh -- host-callable launchers implemented in the .cu files.
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>

#include "device_scene.h"

struct pb2_context;
struct pb2_render_params;

namespace pb2 {

int activate_tables(pb2_context* ctx, int spp);
void free_wavefront(pb2_context* ctx);

// trace_kernels.cu
// scratch: device ints, at least 2 + n (queue cursor, fallback count, fallback list)
void launch_trace_closest(const DScene& sc, const float* rays, size_t n, int accel, int32_t* geom,
                          int32_t* prim, float* t, float* surf, int* scratch, cudaStream_t st);
void launch_trace_any(const DScene& sc, const float* rays, size_t n, int accel, uint8_t* occ, int* scratch,
                      cudaStream_t st);

constexpr int kNumSMs = 148;  // B200
constexpr int kTraceBlock = 128;
constexpr int kFallbackBlocksPerSM = 8;  // reference-order re-trace: one ray per thread, latency bound. ncu on C1
                                         // (profiles/r2/s17): at 6 blocks x 72 registers the kernel ran 24 of 64 warps
                                         // per SM; 8 blocks at <= 64 registers (launch bounds) fill the register file
#ifndef PB2_TRACE_MINBLOCKS
#define PB2_TRACE_MINBLOCKS 8  // __launch_bounds__ min blocks per SM of the persistent traversal kernels
#endif
constexpr int kTraceBlocksPerSM = PB2_TRACE_MINBLOCKS;  // persistent traversal grid = kNumSMs * this
inline int persistent_grid() {
  static const int per_sm = [] {
    const char* v = std::getenv("PB2_TRACE_BLOCKS");  // tuning knob: resident blocks per SM
    const int n = v ? std::atoi(v) : kTraceBlocksPerSM;
    return n > 0 ? n : kTraceBlocksPerSM;
  }();
  return kNumSMs * per_sm;
}

}  // namespace pb2
