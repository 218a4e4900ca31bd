// context.h -- host-side state behind the opaque pb2_context of include/pine_b200.h.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <string>
#include <vector>

#include "bvh_build.h"
#include "bvh_layout.h"
#include "device_scene.h"

namespace pb2 {

void set_error(const std::string& msg);
#define PB2_CUDA_CHECK(expr)                                                              \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      pb2::set_error(std::string(#expr) + ": " + cudaGetErrorString(e_));                 \
      return PB2_E_CUDA;                                                                  \
    }                                                                                     \
  } while (0)

// Owning device allocation that grows on demand.
struct DBuf {
  void* ptr = nullptr;
  size_t cap = 0;
  ~DBuf() { release(); }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    cap = 0;
  }
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc(&ptr, bytes ? bytes : 16);
    if (e == cudaSuccess) cap = bytes ? bytes : 16;
    return e;
  }
  template <class T>
  cudaError_t upload(const std::vector<T>& v) {
    cudaError_t e = reserve(v.size() * sizeof(T));
    if (e != cudaSuccess || v.empty()) return e;
    return cudaMemcpy(ptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
  }
  template <class T>
  T* as() const { return reinterpret_cast<T*>(ptr); }
};

struct SamplerTables {
  std::vector<uint8_t> sobol, scramble, rank;  // host copies per spp variant
};

struct WavefrontBuffers;  // render.cu
}  // namespace pb2
struct pb2_context;
namespace pb2 {

// Host copy of everything pb2_upload_scene lays out for the device. The vectors stay alive (and
// page-locked, cudaHostRegister) for as long as the scene is current, so that the host->device
// staging can be repeated from pinned memory without rebuilding (pb2_restage_scene).
struct HostScene {
  LayoutOut lay;
  std::vector<DTopPrim> top_prims;
  std::vector<DMesh> meshes;
  std::vector<DGeom> geoms;
  std::vector<int32_t> lights;
  std::vector<pb2_light> light_recs;
  raw_vector<float> verts, normals, uvs;  // (raw_vector: no zero-fill of pools the upload overwrites completely)
  raw_vector<uint32_t> tris;
  std::vector<pb2_material> materials;
  std::vector<uint32_t> texels;
  std::vector<DTexture> textures;
  std::vector<float> srgb_lut;
  std::vector<pb2_node_instr> node_instrs;
  std::vector<pb2_aabb> aabbs;
  std::vector<pb2_obb> obbs;
  std::vector<pb2_sphere> spheres;
  std::vector<pb2_triangle> triangles;
  std::vector<pb2_cone> cones;
  std::vector<pb2_rect> rects;
  std::vector<pb2_disk> disks;
  std::vector<pb2_plane> planes;
  std::vector<pb2_cylinder> cylinders;
  // staging list, in upload order. `dst` names the destination by member so that the same list stages the
  // scene into ANY context: the replicas of a pb2_multi share the host scene their first context built.
  struct Item { DBuf pb2_context::* dst; const void* src; size_t bytes; };
  std::vector<Item> items;
  std::vector<void*> registered;    // ranges currently page-locked
  // pb2_import_host_scene: the arrays live in a private file mapping (written by another process's
  // pb2_export_host_scene) instead of the vectors above
  void* map_base = nullptr;
  size_t map_size = 0;
  void unmap();
  bool pinned = false;
  void pin() {  // page-lock the large staging arrays (once): repeated staging then runs at PCIe speed
    if (pinned) return;
    pinned = true;
    for (const Item& it : items)
      if (it.bytes >= (1u << 16) && cudaHostRegister(const_cast<void*>(it.src), it.bytes, cudaHostRegisterDefault) == cudaSuccess)
        registered.push_back(const_cast<void*>(it.src));
      else
        cudaGetLastError();
  }
  void unpin() {
    for (void* p : registered) cudaHostUnregister(p);
    registered.clear();
    pinned = false;
  }
};

}  // namespace pb2

struct pb2_context {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
  bool scene_ready = false;       // device copy valid
  bool has_node_programs = false;  // some material parameter is a node program: the GENERIC shade kernels run
  bool host_scene_ready = false;  // host trees valid (also on a host-only context, device = -1)
  bool defer_staging = false;     // pb2_upload_scene stops after the host build (pb2_multi stages the replicas)

  // host copies kept for the dump API
  pb2::Tree2 top_tree;
  std::vector<pb2::Tree2> mesh_trees;

  pb2::HostScene host;
  const pb2::HostScene* host_view = nullptr;  // whose staging list restage() copies: &host, or the scene's owner (pb2_multi)
  pb2::DScene scene{};  // device pointers
  pb2::DBuf nodes2, nodes8, leaves, wide_prims, box_shapes, leaf_prims, top_prims, meshes, geoms, materials, lights, light_recs;
  pb2::DBuf aabbs, obbs, spheres, triangles, cones, rects, disks, planes, cylinders;
  pb2::DBuf verts, tris, normals, uvs, texels, textures, srgb_lut, node_instrs;
  pb2::DBuf tab_sobol, tab_scramble, tab_rank;
  pb2::DBuf halton_primes, halton_sums, halton_perms;  // HaltonSampler tables, made on first use
  int tables_spp = 0;
  pb2::SamplerTables tables[9];  // index = log2(spp)
  bool has_tables[9] = {};
  bool only_area_lights = true;       // no delta light, no environment light: the specialised shade kernels apply
  unsigned material_kinds = 0;        // bit k: some geometry uses pb2_material_kind k
  bool needs_rng_in_shading = false;  // Uber / Subsurface present (they draw randf(), material.cpp:9-28)
  int n_pixels = 0;

  // scratch for trace hooks / render
  pb2::DBuf rays, out_geom, out_prim, out_t, out_surf, out_occ, trace_scratch;
  pb2::DBuf film, film_accum, walk_stats;
  bool l2_window_set = false;  // an access-policy window is installed on the render stream
  size_t film_pixels = 0;  // pixels the film buffer was last cleared for (PB2_RENDER_KEEP_FILM)
  pb2::WavefrontBuffers* wf = nullptr;

  pb2_ray_record* record = nullptr;  // test hook (pb2_set_ray_record)
  // pb2_set_progress_callback: called from cudaLaunchHostFunc behind every wave of a render
  pb2_progress_fn progress_fn = nullptr;
  void* progress_user = nullptr;
  std::atomic<int> progress_done{0};
  int progress_total = 1;

  pb2_counters counters{};
};
