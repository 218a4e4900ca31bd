// kernels.cuh -- host-callable launchers implemented in the .cu files.
#pragma once
#include <cuda_runtime.h>

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
constexpr int kTraceBlocksPerSM = 6;  // persistent traversal grid = kNumSMs * kTraceBlocksPerSM
inline int persistent_grid() { return kNumSMs * kTraceBlocksPerSM; }

}  // namespace pb2
