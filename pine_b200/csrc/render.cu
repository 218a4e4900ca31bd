// render.cu -- the wavefront path tracer: PathIntegrator::render / ::radiance
// (reference src/pine/impl/integrator/path.cpp:26-124) as sm_100a kernels.
//
// One wave = a block of (pixel, sample) slots. Per path depth k:
//   k_trace   closest hit for every live ray                     (Accel::intersect, path.cpp:48)
//   k_shade   surface, emission/termination, material -> BxDF, light sample + shadow ray,
//             BSDF sample + next ray, vertex record k            (path.cpp:74-120)
//   k_shadow  any-hit for the shadow rays; occluded => zero the vertex's light term (path.cpp:100)
// and after the last depth
//   k_fold    the recursion's unwind: X = min(beta*(nee + (X*f)*s), 8) deepest vertex first
//             (path.cpp:119-121 -- the clamp makes the estimator non-linear in the suffix, so it
//             cannot be a running throughput product)
//   k_resolve per pixel L += sample radiance in sample order (path.cpp:36), one writer per pixel.
// Queues are compacted with warp ballots (one atomic per warp); counts live in device memory and
// every kernel is a grid-stride loop over a grid that is a multiple of the SM count, so a whole
// render is a stream of launches with no host synchronisation in between.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bsdf.cuh"
#include "context.h"
#include "kernels.cuh"
#include "surface.cuh"
#include "wide_persistent.cuh"

namespace pb2 {

struct WaveView {
  // ray queues (double buffered): o.xyz + tmax, d.xyz + slot
  float4* ray_o[2];
  float4* ray_d[2];
  float4* hit;            // per queue entry: t, geom, prim, -
  float4* sh_o;           // shadow queue: o.xyz + tmax
  float4* sh_d;           // d.xyz + (slot | depth << 24)
  // per slot
  float* pv_pdf;          // Vertex::pdf
  int* pv_flags;          // length | diffuse_length << 8 | is_delta << 16
  int* dim;               // BlueSobolSampler::dimension
  float4* term;           // terminal radiance X.rgb, number of vertex records
  float4* vtx;            // [depth][slot][2]: (nee.rgb, s) (f.rgb, beta code)
  // per pixel of this render's row range
  uint2* rng;             // 2 x u64 state as 4 x u32
  float4* lacc;           // running L
  int* counts;            // [0..D] live rays per depth, [D+1..2D+1] shadow rays per depth, then per
                          // depth {closest cursor, closest fallbacks, shadow cursor, shadow fallbacks}
  int* fb_list;           // fallback list shared by the traversal kernels (used one at a time)
  uint8_t* occ;           // NULL unless a ray record is being taken
  int* lists;             // shade lists: [list][n_slots] queue indices, one list per shading class
  int list_of_class[8];   // class (0 = terminal, k = pb2_material_kind k) -> list number, -1 = absent
  int n_slots;            // capacity of this wave
  int pixels;             // pixels per sample layer in this wave
  int pixel0;             // first pixel (linear index in the row range) of this wave
  int width, row_begin;
  int row_band, row_parts, row_part;  // interleaved row bands (pb2_render_params); row_parts <= 1: contiguous rows
  int s_batch, sample0;   // samples [sample0, sample0 + s_batch) of every pixel in the wave
  int max_depth;          // PathIntegrator::max_path_length
  int accel;
};

// Wavefront state. A render is cut into waves; up to kMaxLanes waves over DISJOINT pixel ranges are in
// flight at once, each on its own stream with its own queue set ("lane"), so that the drain tail of one
// lane's persistent traversal launch, its few-ray reference-order re-trace and its memory-streaming shade
// kernels are filled by the other lane's work instead of leaving SMs idle (pb2_render_device).
constexpr int kMaxLanes = 4;
constexpr int kDefaultLanes = 4;  // measured on B200 (profiles/r2): C5 458 / 502 / 511 / 516 Msamples/s with 1 / 2 / 3 / 4
                                  // lanes, C3 1123 / 1197 / - / 1213, C1 822 / 870 / - / 866
constexpr long kMinLaneSlots = 1L << 19;  // a lane must still fill the GPU: below this, one lane
struct LaneBuffers {
  DBuf ray_o[2], ray_d[2], hit, sh_o, sh_d, pv_pdf, pv_flags, dim, term, vtx, fb_list, occ, lists;
  cudaStream_t stream = nullptr;  // lanes >= 1 only; lane 0 runs on the caller's stream
  cudaEvent_t done = nullptr;
};
struct WavefrontBuffers {
  LaneBuffers lane[kMaxLanes];
  DBuf rng, lacc, counts;           // per pixel of the row range / per wave: shared by the lanes
  cudaEvent_t fork = nullptr;
  std::vector<int> host_counts;
  std::vector<cudaEvent_t> events;  // PB2_RENDER_TIME_KERNELS: one event between consecutive launches
  ~WavefrontBuffers() {
    for (cudaEvent_t e : events) cudaEventDestroy(e);
    for (LaneBuffers& l : lane) {
      if (l.stream) cudaStreamDestroy(l.stream);
      if (l.done) cudaEventDestroy(l.done);
    }
    if (fork) cudaEventDestroy(fork);
  }
};

void free_wavefront(pb2_context* ctx) {
  delete ctx->wf;
  ctx->wf = nullptr;
}

// ---- helpers ------------------------------------------------------------------------------------------
// Film row of the render's `local`-th row: contiguous from row_begin, or the bands of `band` rows this call owns
// out of every `parts` (band k of the range belongs to part k % parts).
__device__ __forceinline__ int film_row(int row_begin, int band, int parts, int part, int local) {
  if (parts <= 1) return row_begin + local;
  return row_begin + ((local / band) * parts + part) * band + local % band;
}
__device__ __forceinline__ int warp_compact(bool keep, int* counter) {
  unsigned mask = __ballot_sync(0xffffffffu, keep);
  int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == 0 && mask) base = atomicAdd(counter, __popc(mask));
  base = __shfl_sync(0xffffffffu, base, 0);
  return base + __popc(mask & ((1u << lane) - 1));
}
// Warp-uniform grid-stride loop: every lane of a warp runs the same number of iterations.
#define PB2_WARP_LOOP(i, n)                                                                       \
  for (int base_ = (blockIdx.x * blockDim.x + threadIdx.x) & ~31, i = base_ + (threadIdx.x & 31); \
       base_ < (n); base_ += gridDim.x * blockDim.x, i = base_ + (threadIdx.x & 31))

__device__ __forceinline__ Rng load_rng(const uint2* p, int pixel) {
  uint2 a = p[2 * pixel], b = p[2 * pixel + 1];
  Rng g;
  g.s0 = uint64_t(a.x) | (uint64_t(a.y) << 32);
  g.s1 = uint64_t(b.x) | (uint64_t(b.y) << 32);
  return g;
}
__device__ __forceinline__ void store_rng(uint2* p, int pixel, const Rng& g) {
  p[2 * pixel] = make_uint2(uint32_t(g.s0), uint32_t(g.s0 >> 32));
  p[2 * pixel + 1] = make_uint2(uint32_t(g.s1), uint32_t(g.s1 >> 32));
}

// ---- raygen (path.cpp:32-35) -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_raygen(DScene sc, WaveView w, int seed_sample, int first_wave_of_pixel) {
  PB2_WARP_LOOP(i, w.pixels) {
    if (i >= w.pixels) continue;
    int pix = w.pixel0 + i;  // linear in the row range
    int px = pix % w.width, py = film_row(w.row_begin, w.row_band, w.row_parts, w.row_part, pix / w.width);
    Rng g = first_wave_of_pixel ? rng_seed(murmur_pixel(px, py, seed_sample)) : load_rng(w.rng, pix);
    for (int s = 0; s < w.s_batch; s++) {
      // gen_ray((p + rand2f()) / size, rand2f()) (path.cpp:35): g++ evaluates the arguments right
      // to left, so the lens sample is drawn BEFORE the pixel jitter.
      float r2 = rng_nextf(g), r3 = rng_nextf(g), r0 = rng_nextf(g), r1 = rng_nextf(g);
      V2 p_film((float(px) + r0) / float(sc.camera.width), (float(py) + r1) / float(sc.camera.height));
      Ray r = camera_ray(sc.camera, p_film, V2(r2, r3));
      int slot = s * w.pixels + i;
      w.ray_o[0][slot] = make_float4(r.o.x, r.o.y, r.o.z, r.tmax);
      w.ray_d[0][slot] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(slot));
      w.pv_pdf[slot] = 0.0f;
      w.pv_flags[slot] = 1 << 16;  // Vertex::first_vertex(): length 0, diffuse_length 0, is_delta
      w.dim[slot] = 0;
      w.term[slot] = make_float4(0, 0, 0, __int_as_float(0));
    }
    store_rng(w.rng, pix, g);
    if (i == 0) w.counts[0] = w.pixels * w.s_batch;
  }
}

// ---- closest hit (Accel::intersect, path.cpp:48) ------------------------------------------------------------
struct TraceIO {
  const float4* ro;
  const float4* rd;
  float4* hit;
  __device__ __forceinline__ Ray load(int i) const {
    float4 a = ro[i], b = rd[i];
    Ray ray;
    ray.o = V3(a.x, a.y, a.z);
    ray.d = V3(b.x, b.y, b.z);
    ray.tmin = 0.0f;
    ray.tmax = a.w;
    return ray;
  }
  __device__ __forceinline__ void store_closest(const DScene&, int i, const Ray& ray, bool ok, Hit h) const {
    hit[i] = make_float4(ray.tmax, __int_as_float(ok ? h.geom : -1), __int_as_float(h.prim), 0.0f);
  }
  __device__ __forceinline__ void store_any(int, bool) const {}
};
// ---- shadow rays (RTIntegrator::hit, path.cpp:100): occluded => zero the vertex's light term ----------------
struct ShadowIO {
  const float4* so;
  const float4* sd;
  float4* vtx;  // records of this depth: [slot][2]
  uint8_t* occ; // test hook (pb2_set_ray_record): occlusion per shadow-queue entry, else NULL
  __device__ __forceinline__ Ray load(int i) const {
    float4 a = so[i], b = sd[i];
    Ray ray;
    ray.o = V3(a.x, a.y, a.z);
    ray.d = V3(b.x, b.y, b.z);
    ray.tmin = 0.0f;
    ray.tmax = a.w;
    return ray;
  }
  __device__ __forceinline__ void store_closest(const DScene&, int, const Ray&, bool, Hit) const {}
  __device__ __forceinline__ void store_any(int i, bool occluded) const {
    if (occ) occ[i] = occluded ? 1 : 0;
    if (!occluded) return;
    const int slot = __float_as_int(sd[i].w);
    float4* v = vtx + size_t(slot) * 2;
    v[0].x = 0.0f; v[0].y = 0.0f; v[0].z = 0.0f;
  }
};
__device__ __forceinline__ WideQueue wave_queue(const WaveView& w, int depth, int which) {
  int* c = w.counts + 2 * (w.max_depth + 1) + 4 * depth + 2 * which;
  return WideQueue{c, c + 1, w.fb_list};
}
__device__ __forceinline__ TraceIO trace_io(const WaveView& w, int depth) {
  return TraceIO{w.ray_o[depth & 1], w.ray_d[depth & 1], w.hit};
}
__device__ __forceinline__ ShadowIO shadow_io(const WaveView& w, int depth) {
  return ShadowIO{w.sh_o, w.sh_d, w.vtx + size_t(depth) * w.n_slots * 2, w.occ};
}
// SHAPES: which analytic shapes are compiled in (shapes.cuh kShapes*); the reference-order kernels below know two
// levels only (false = kShapesCommon, true = kShapesAll: the scene holds a Plane or a Cylinder)
template <int SHAPES>
__global__ void __launch_bounds__(kTraceBlock, PB2_TRACE_MINBLOCKS) k_trace(DScene sc, WaveView w, int depth) {
  wide_persistent<false, SHAPES>(sc, w.counts[depth], wave_queue(w, depth, 0), trace_io(w, depth));
}
template <int SHAPES>
__global__ void __launch_bounds__(kTraceBlock, kFallbackBlocksPerSM) k_trace_fallback(DScene sc, WaveView w, int depth) {
  wide_fallback<false, SHAPES>(sc, wave_queue(w, depth, 0), trace_io(w, depth));
}
template <bool RARE>
__global__ void __launch_bounds__(kTraceBlock) k_trace_ref(DScene sc, WaveView w, int depth) {
  ref_order_all<false, (RARE ? kShapesAll : kShapesCommon)>(sc, w.counts[depth], trace_io(w, depth));
}
template <int SHAPES>
__global__ void __launch_bounds__(kTraceBlock, PB2_TRACE_MINBLOCKS) k_shadow(DScene sc, WaveView w, int depth) {
  wide_persistent<true, SHAPES>(sc, w.counts[w.max_depth + 1 + depth], wave_queue(w, depth, 1), shadow_io(w, depth));
}
template <bool RARE>
__global__ void __launch_bounds__(kTraceBlock, kFallbackBlocksPerSM) k_shadow_fallback(DScene sc, WaveView w, int depth) {
  wide_fallback<true, (RARE ? kShapesAll : kShapesCommon)>(sc, wave_queue(w, depth, 1), shadow_io(w, depth));
}
template <bool RARE>
__global__ void __launch_bounds__(kTraceBlock) k_shadow_ref(DScene sc, WaveView w, int depth) {
  ref_order_all<true, (RARE ? kShapesAll : kShapesCommon)>(sc, w.counts[w.max_depth + 1 + depth], shadow_io(w, depth));
}

// sky_color(direction) (color.cpp:100-103): sqr(lerp(d.y / 2 + 0.7, warm, cool)), lerp = a (1 - t) + b t
__device__ __forceinline__ V3 sky_color(V3 d) {
  const float t = d.y / 2 + 0.7f;
  const V3 c = V3(1.0f, 0.8f, 0.6f) * (1.0f - t) + V3(0.6f, 0.8f, 1.0f) * t;
  return c * c;
}

// ---- material -> BxDF (material.h:30-121, material.cpp:9-28) -------------------------------------------------
__device__ __forceinline__ bool with_probability(float prob, Rng& g) {  // sampler.h:317-324
  if (prob == 0) return false;
  if (prob == 1) return true;
  return rng_nextf(g) < prob;
}
// NodeImage(NodeUV(), image)(ctx) (node.cpp:20-23, image.cpp:7-18): nearest texel with wrap, gamma-decoded
__device__ __forceinline__ V3 image_node(const DScene& sc, int texture1, V2 uv) {
  const DTexture t = sc.textures[texture1 - 1];
  const float fx = uv.x - floorf(uv.x), fy = uv.y - floorf(uv.y);  // psl::fract (psl/math.h:151-154)
  const int x = min(int(float(t.width) * fx), t.width - 1), y = min(int(float(t.height) * fy), t.height - 1);
  const uint32_t c = __ldg(sc.texels + t.offset + size_t(x) + size_t(y) * size_t(t.width));
  return V3(__ldg(sc.srgb_lut + (c & 0xffu)), __ldg(sc.srgb_lut + ((c >> 8) & 0xffu)),
            __ldg(sc.srgb_lut + ((c >> 16) & 0xffu)));
}
// Mnode<T>::operator()(NodeEvalCtx{p, n, uv}) for a graph flattened to a postfix program (include/pine_b200.h
// pb2_node_op; node.h:105-250, node.cpp:8-27). Every stack value is a vec3; a float node's value sits in all
// three components, so NodeBinary<float, op> is NodeBinary<vec3, op> component by component. `start` is 1-based.
__device__ __noinline__ V3 eval_node_program(const DScene& sc, int start, V3 p, V3 n, V2 uv) {
  V3 st[PB2_NODE_MAX_STACK];
  int sp = 0;
  for (const float4* pc = reinterpret_cast<const float4*>(sc.node_instrs) + (start - 1);; pc++) {
    const float4 in = __ldg(pc);
    const V3 a(in.y, in.z, in.w);
    switch (__float_as_int(in.x)) {
      case PB2_NODE_END: return st[sp - 1];
      case PB2_NODE_CONST: st[sp++] = a; break;
      case PB2_NODE_POSITION: st[sp++] = p; break;
      case PB2_NODE_NORMAL: st[sp++] = n; break;
      case PB2_NODE_UV: st[sp++] = V3(uv.x, uv.y, 0.0f); break;
      case PB2_NODE_ADD: sp--; st[sp - 1] = st[sp - 1] + st[sp]; break;
      case PB2_NODE_SUB: sp--; st[sp - 1] = st[sp - 1] - st[sp]; break;
      case PB2_NODE_MUL: sp--; st[sp - 1] = st[sp - 1] * st[sp]; break;
      case PB2_NODE_DIV: sp--; st[sp - 1] = V3(st[sp - 1].x / st[sp].x, st[sp - 1].y / st[sp].y, st[sp - 1].z / st[sp].z); break;
      case PB2_NODE_POW: sp--; st[sp - 1] = V3(powf(st[sp - 1].x, st[sp].x), powf(st[sp - 1].y, st[sp].y), powf(st[sp - 1].z, st[sp].z)); break;
      case PB2_NODE_NEG: st[sp - 1] = -st[sp - 1]; break;
      case PB2_NODE_ABS: st[sp - 1] = V3(pabs(st[sp - 1].x), pabs(st[sp - 1].y), pabs(st[sp - 1].z)); break;
      case PB2_NODE_SQR: st[sp - 1] = st[sp - 1] * st[sp - 1]; break;
      case PB2_NODE_SQRT: st[sp - 1] = V3(sqrtf(st[sp - 1].x), sqrtf(st[sp - 1].y), sqrtf(st[sp - 1].z)); break;
      case PB2_NODE_FRACT: st[sp - 1] = V3(st[sp - 1].x - floorf(st[sp - 1].x), st[sp - 1].y - floorf(st[sp - 1].y), st[sp - 1].z - floorf(st[sp - 1].z)); break;
      case PB2_NODE_COMPONENT: { const float v = st[sp - 1][int(a.x)]; st[sp - 1] = V3(v); break; }
      case PB2_NODE_VEC3: sp -= 2; st[sp - 1] = V3(st[sp - 1].x, st[sp].x, st[sp + 1].x); break;
      case PB2_NODE_CHECKERBOARD: {
        const V3 q = st[sp - 1];
        const V3 x = V3(q.x - floorf(q.x), q.y - floorf(q.y), q.z - floorf(q.z)) - V3(a.x);
        st[sp - 1] = V3(x.x * x.y * x.z > 0 ? 1.0f : 0.0f);
        break;
      }
      case PB2_NODE_IMAGE: st[sp - 1] = image_node(sc, int(a.x), V2(st[sp - 1].x, st[sp - 1].y)); break;
      case PB2_NODE_IMAGEF: st[sp - 1] = V3(image_node(sc, int(a.x), V2(st[sp - 1].x, st[sp - 1].y)).x); break;
      default: return V3(0.0f);  // rejected at pb2_upload_scene
    }
  }
}

// NODES: the scene has node programs (only the GENERIC shade kernels are compiled with it)
template <int MAT, bool NODES, class Smp>
__device__ __forceinline__ Bxdf material_bxdf(const DScene& sc, const pb2_material& m, V3 wi_world, V3 p, V3 n, V2 uv,
                                              float min_roughness, bool diffused, Smp& smp, Rng& g) {
  Bxdf b;
  b.albedo = V3(m.color[0], m.color[1], m.color[2]);
  b.roughness = m.roughness;
  float metallic = m.metallic, transmission = m.transmission;
  if (MAT == PB2_MAT_UBER) {
    if (m.color_texture) b.albedo = image_node(sc, m.color_texture, uv);
    if (m.roughness_texture) b.roughness = image_node(sc, m.roughness_texture, uv).y;
    if (m.metallic_texture) metallic = image_node(sc, m.metallic_texture, uv).z;
  }
  b.ior = m.ior;
  if (NODES) {
    if (m.color_node) b.albedo = eval_node_program(sc, m.color_node, p, n, uv);
    if (m.roughness_node) b.roughness = eval_node_program(sc, m.roughness_node, p, n, uv).x;
    if ((MAT == PB2_MAT_GLOSSY || MAT == PB2_MAT_GLASS) && m.ior_node) b.ior = eval_node_program(sc, m.ior_node, p, n, uv).x;
    if (MAT == PB2_MAT_UBER) {
      if (m.metallic_node) metallic = eval_node_program(sc, m.metallic_node, p, n, uv).x;
      if (m.transmission_node) transmission = eval_node_program(sc, m.transmission_node, p, n, uv).x;
    }
  }
  b.sigma_s = V3(m.sigma_s[0], m.sigma_s[1], m.sigma_s[2]);
  b.kind = BX_DIFFUSE;
  if (MAT == PB2_MAT_METAL) { b.kind = BX_CONDUCTOR; b.roughness = pmax(b.roughness, min_roughness); }
  if (MAT == PB2_MAT_GLOSSY) { b.kind = BX_DIFF_DIELECTRIC; b.roughness = pmax(b.roughness, min_roughness); }
  if (MAT == PB2_MAT_GLASS) { b.kind = BX_REFR_DIELECTRIC; b.roughness = pmax(b.roughness, min_roughness); }
  if (MAT == PB2_MAT_UBER) {
    if (with_probability(metallic, g)) b.kind = BX_CONDUCTOR;
    else if (with_probability(transmission, g)) b.kind = BX_REFR_DIELECTRIC;
    else b.kind = BX_DIFF_DIELECTRIC;
  }
  if (MAT == PB2_MAT_SUBSURFACE) {
    b.ior = 1.4f;  // SubsurfaceMaterial::ior (material.h:103)
    float fr = fresnel_unpolarized(dot(wi_world, n), b.ior);
    if (smp_get1d(sc, smp) < fr) { b.kind = BX_REFRACTIVE; b.roughness = pmax(b.roughness, min_roughness); }
    else if (diffused) b.kind = BX_DIFFUSE;
    else b.kind = BX_BSSRDF;
  }
  return b;
}

// BSSRDF::sample_p (bxdf.cpp:329-357): random walk inside the hit shape only. For analytic shapes
// the reference's Shape::intersect leaves the exit interaction untouched (p = n = 0); reproduced.
template <class Smp>
__device__ inline bool bssrdf_walk(const DScene& sc, const Bxdf& b, int geom, V3 wi_world, Surf& it,
                                   Smp& smp, Rng& g, V3& beta, V3& wi_out) {
  V3 p = it.p;
  V3 w = -wi_world;
  if (!refract_through(wi_world, it.n, b.ior, w)) return false;
  int channel = int(rng_nextf(g) * 3);
  beta = V3(0.0f);
  beta[channel] = 3;
  float sigma_t_inv = 1 / b.sigma_s[channel];
  DGeom gm = sc.geoms[geom];
  for (int i = 0; i < 4096; i++) {
    Ray ray;
    if (i == 0) {  // spawn_ray(p, n, w) (ray.h:39-46)
      ray.o = offset_ray_origin(p, it.n);
      ray.d = w; ray.tmin = 0; ray.tmax = kFloatMax * (1.0f - 1e-3f);
    } else {
      ray.o = p; ray.d = w; ray.tmin = 0; ray.tmax = kFloatMax;
    }
    Surf exit_it;
    exit_it.p = V3(); exit_it.n = V3(); exit_it.uv = V2();
    bool hit;
    if (gm.kind == PB2_SHAPE_MESH) {
      const DMesh& m = sc.meshes[gm.index];
      Octant q = make_octant(ray);
      int prim = -1;
      hit = walk_ref_order<false>(sc, m.tree, ray, q, [&](int tri) {
        V3 v0, v1, v2;
        mesh_tri(sc, m, tri, v0, v1, v2);
        bool ok = tri_intersect(ray, v0, v1, v2);
        if (ok) prim = tri;
        return ok;
      });
      if (hit) surface_info(sc, geom, prim, ray.o + ray.tmax * ray.d, exit_it);
    } else {
      hit = shape_intersect<kShapesCommon>(sc, gm.kind, gm.index, ray);  // Subsurface on a Plane / Cylinder is refused at upload
    }
    if (!hit) return false;
    float t = -logf(1 - smp_get1d(sc, smp)) * sigma_t_inv;
    if (ray.tmax < t) {
      wi_out = -w;
      it.p = exit_it.p;
      it.n = exit_it.n;
      return true;
    }
    p = ray.o + t * ray.d;
    w = uniform_sphere(smp_get2d(sc, smp));
  }
  return false;
}

// ---- shade ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_vertex(const WaveView& w, int depth, int slot, V3 nee, float s, V3 f, int beta_code) {
  float4* v = w.vtx + (size_t(depth) * w.n_slots + slot) * 2;
  v[0] = make_float4(nee.x, nee.y, nee.z, s);
  v[1] = make_float4(f.x, f.y, f.z, __int_as_float(beta_code));
}

// The shade queue is bucketed by shading class first (k_classify): class 0 = the path ends at this
// vertex (miss, emitter, or depth cut: path.cpp:56-66,89), class k = pb2_material_kind k. Each class
// then runs its own kernel, compiled with only that material's BxDF code: one kernel for all
// materials was 315 KB of SASS running at 8.5 of 32 lanes and stalled on instruction fetch
// (profiles/r1), because neighbouring rays hit different materials.
__device__ __forceinline__ int* list_count(const WaveView& w, int depth, int cls) {
  return w.counts + 6 * (w.max_depth + 1) + 8 * depth + cls;
}
// One block buckets a tile of 256 x kClassifyItems rays: positions inside the tile come from shared-memory
// counters (one shared atomic per warp and class), and the tile's ranges in the global lists from ONE global
// atomic per class and tile. With one global atomic per warp the kernel ran at the rate the L2 serialises
// same-address atomics (0.37 ms for 16.7 M rays on C3, 10% of the step), not at memory speed.
constexpr int kClassifyItems = 8;
__global__ void __launch_bounds__(256) k_classify(DScene sc, WaveView w, int depth) {
  __shared__ int s_count[8], s_base[8];
  const int n = w.counts[depth];
  const unsigned lane = threadIdx.x & 31u;
  const int tile_items = 256 * kClassifyItems;
  for (int tile = blockIdx.x * tile_items; tile < n; tile += gridDim.x * tile_items) {  // block-uniform
    if (threadIdx.x < 8) s_count[threadIdx.x] = 0;
    __syncthreads();
    int cls_of[kClassifyItems], local[kClassifyItems];
#pragma unroll
    for (int k = 0; k < kClassifyItems; k++) {
      const int i = tile + k * 256 + int(threadIdx.x);
      int cls = -1;
      if (i < n) {
        const int geom = __float_as_int(w.hit[i].y);
        cls = 0;
        if (geom >= 0 && depth + 1 < w.max_depth) {  // Vertex::length == depth on this queue
          const int kind = sc.materials[sc.geoms[geom].material].kind;
          if (kind != PB2_MAT_EMISSIVE) cls = kind;
        }
      }
      const unsigned peers = __match_any_sync(0xffffffffu, cls);
      const int leader = __ffs(peers) - 1;
      int base = 0;
      if (cls >= 0 && int(lane) == leader) base = atomicAdd(&s_count[cls], __popc(peers));
      base = __shfl_sync(0xffffffffu, base, leader);
      cls_of[k] = cls;
      local[k] = base + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    if (threadIdx.x < 8 && s_count[threadIdx.x])
      s_base[threadIdx.x] = atomicAdd(list_count(w, depth, threadIdx.x), s_count[threadIdx.x]);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kClassifyItems; k++)
      if (cls_of[k] >= 0)
        w.lists[size_t(w.list_of_class[cls_of[k]]) * w.n_slots + s_base[cls_of[k]] + local[k]] =
            tile + k * 256 + int(threadIdx.x);
    __syncthreads();
  }
}

// class 0: emitted radiance seen by this vertex, or nothing (path.cpp:56-66)
__global__ void __launch_bounds__(256) k_shade_terminal(DScene sc, WaveView w, int depth) {
  const int n = *list_count(w, depth, 0);
  const int cur = depth & 1;
  const int* list = w.lists + size_t(w.list_of_class[0]) * w.n_slots;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const int i = list[j];
    const float4 b = w.ray_d[cur][i], hrec = w.hit[i];
    const int slot = __float_as_int(b.w);
    const int geom = __float_as_int(hrec.y);
    V3 X(0.0f);
    if (geom < 0 && sc.env_kind) {
      // Lo += Tr * env_light->color(ray.d); light_pdf = env_light->pdf(ray.d) unless pv.is_delta (path.cpp:76-81)
      X = V3(1.0f) * (V3(sc.env_color[0], sc.env_color[1], sc.env_color[2]) * sky_color(V3(b.x, b.y, b.z)));
      const bool pv_delta = (w.pv_flags[slot] >> 16) & 1;
      if (!pv_delta && depth > 0) {
        const float mis = balance_heuristic(w.pv_pdf[slot], 1 / (4 * kPi));
        float4* v = w.vtx + (size_t(depth - 1) * w.n_slots + slot) * 2;
        v[0].w = v[0].w * mis;
      }
    }
    if (geom >= 0) {
      const DGeom g = sc.geoms[geom];
      const pb2_material m = sc.materials[g.material];
      if (m.kind == PB2_MAT_EMISSIVE) {
        const float4 a = w.ray_o[cur][i];
        const V3 ro(a.x, a.y, a.z), rd(b.x, b.y, b.z);
        const V3 wi = -rd;
        const float t = hrec.x;
        Surf it;
        surface_info(sc, geom, __float_as_int(hrec.z), ro + t * rd, it);
        // Lo += Tr * le; light_pdf = light_sampler.pdf(ray, it) unless the previous vertex was delta
        if (!(dot(wi, it.n) < 0.0f)) X = V3(1.0f) * V3(m.color[0], m.color[1], m.color[2]);
        const bool pv_delta = (w.pv_flags[slot] >> 16) & 1;
        if (!pv_delta && depth > 0) {
          float light_pdf = shape_pdf(sc, geom, ro, rd, t, it.n) / float(sc.n_lights);
          float bs_pdf = w.pv_pdf[slot];
          float mis = balance_heuristic(bs_pdf, light_pdf);
          float4* v = w.vtx + (size_t(depth - 1) * w.n_slots + slot) * 2;
          v[0].w = v[0].w * mis;  // s = cosine / bs.pdf * mis
        }
      }
    }
    w.term[slot] = make_float4(X.x, X.y, X.z, __int_as_float(depth));
  }
}

// class MAT: NEE + BSDF sample at a surface vertex (path.cpp:74-120)
template <int MAT, bool GENERIC>
#ifndef PB2_SHADE_MINBLOCKS
#define PB2_SHADE_MINBLOCKS 1
#endif
__global__ void __launch_bounds__(128, PB2_SHADE_MINBLOCKS) k_shade_mat(DScene sc, WaveView w, int depth) {
  const int n = *list_count(w, depth, MAT);
  const int cur = depth & 1, nxt = cur ^ 1;
  int* next_count = w.counts + depth + 1;
  int* shadow_count = w.counts + w.max_depth + 1 + depth;
  const int* list = w.lists + size_t(w.list_of_class[MAT]) * w.n_slots;
  // Output positions: warps add their counts to shared counters, the block takes ONE range per queue from
  // the global counters (k_classify's reasoning: one global atomic per warp and queue ran into the L2's
  // same-address atomic rate). The loop is block-uniform so the barriers are safe.
  __shared__ int s_count[2], s_base[2];
  for (int tile = blockIdx.x * blockDim.x; tile < n; tile += gridDim.x * blockDim.x) {
    const int j = tile + int(threadIdx.x);
    const bool valid = j < n;
    if (threadIdx.x < 2) s_count[threadIdx.x] = 0;
    __syncthreads();
    bool emit_next = false, emit_shadow = false;
    Ray next_ray, shadow_ray;
    int slot = 0;
    if (valid) {
      const int i = list[j];
      float4 a = w.ray_o[cur][i], b = w.ray_d[cur][i], hrec = w.hit[i];
      V3 ro(a.x, a.y, a.z), rd(b.x, b.y, b.z);
      slot = __float_as_int(b.w);
      const float t = hrec.x;
      const int geom = __float_as_int(hrec.y), prim = __float_as_int(hrec.z);
      const int flags = w.pv_flags[slot];
      const int pv_length = flags & 0xff, pv_diffuse = (flags >> 8) & 0xff;
      const V3 wi = -rd;
      Surf it;
      surface_info(sc, geom, prim, ro + t * rd, it);
      const DGeom g = sc.geoms[geom];
      const pb2_material m = sc.materials[g.material];
      int pix = w.pixel0 + slot % w.pixels;
      int px = pix % w.width, py = film_row(w.row_begin, w.row_band, w.row_parts, w.row_part, pix / w.width);
      PixelSamplerT<GENERIC> smp{px, py, w.sample0 + slot / w.pixels, w.dim[slot]};
      Rng g_rng;
      constexpr bool uses_rng = MAT == PB2_MAT_UBER || MAT == PB2_MAT_SUBSURFACE;
      if (uses_rng) g_rng = load_rng(w.rng, pix);
      const bool diffused = pv_diffuse > 0;
      Bxdf bx = material_bxdf<MAT, GENERIC>(sc, m, wi, it.p, it.n, it.uv, diffused ? 0.6f : 0.0f, diffused, smp, g_rng);
      M3 l2w = coordinate_system(it.n);
      M3 w2l = transpose(l2w);
      bx.wi = w2l * wi;
      int beta_code = 0;
      if constexpr (MAT == PB2_MAT_SUBSURFACE) {
        if (bx.kind == BX_BSSRDF) {
          V3 beta, wi2;
          if (bssrdf_walk(sc, bx, geom, wi, it, smp, g_rng, beta, wi2)) {
            beta_code = beta.x != 0 ? 1 : (beta.y != 0 ? 2 : 3);
            l2w = coordinate_system(it.n);
            w2l = transpose(l2w);
            bx.wi = w2l * wi2;
          }
        }
      }
      if (uses_rng) store_rng(w.rng, pix, g_rng);
      // next-event estimation (path.cpp:98-113); draws get2d then get1d (g++ evaluates the
      // arguments of light_sampler.sample(p, get1d(), get2d()) right to left, lightsampler.h:27)
      V3 nee(0.0f);
      if (!bxdf_is_delta(bx)) {
        V2 u2 = smp_get2d(sc, smp);
        float u1 = smp_get1d(sc, smp);
        if (sc.n_lights > 0) {
          u1 *= sc.n_lights;
          int index = int(u1);
          int lgeom = sc.lights[index];
          if (!GENERIC || lgeom >= 0) {  // AreaLight::sample (light.cpp:55-69)
            ShapeSample ss;
            if (shape_sample(sc, lgeom, it.p, u2, u1 - index, ss)) {
              const pb2_material lm = sc.materials[sc.geoms[lgeom].material];
              V3 le = dot(-ss.w, ss.n) < 0.0f ? V3(0.0f) : V3(lm.color[0], lm.color[1], lm.color[2]);
              if (!is_zero(le)) {
                float ls_pdf = ss.pdf / sc.n_lights;
                float cosine = absdot(ss.w, it.n);
                V3 wo = w2l * ss.w;
                V3 f = bxdf_f(bx, wo);
                float mis = balance_heuristic(ls_pdf, bxdf_pdf(bx, wo));
                nee = le * V3(1.0f) * cosine * f / ls_pdf * mis;
                // a light sample the BxDF gives no weight (below the surface of an opaque lobe) adds
                // nothing whether or not it is occluded: no shadow ray for it
                if (!is_zero(nee)) {
                  shadow_ray = spawn_ray(it.p, it.n, ss.w, ss.distance);
                  emit_shadow = true;
                }
              }
            }
          } else {
            // PointLight / SpotLight / DirectionalLight / Sky ::sample (light.cpp:11-53,71-84)
            const pb2_light L = sc.light_recs[index];
            V3 lwo, le;
            float ldist, lpdf;
            bool ok = true;
            if (L.kind == PB2_LIGHT_DIRECTIONAL) {
              ldist = 1e+10f;
              lwo = V3(L.direction[0], L.direction[1], L.direction[2]);
              lpdf = 1.0f;
              le = V3(L.color[0], L.color[1], L.color[2]);
            } else if (L.kind == PB2_LIGHT_SKY) {
              lwo = uniform_sphere(u2);
              lpdf = 1 / (4 * kPi);
              ldist = kFloatMax;
              le = V3(L.color[0], L.color[1], L.color[2]) * sky_color(lwo);
            } else {
              lwo = normalize(V3(L.position[0], L.position[1], L.position[2]) - it.p, ldist);
              le = V3(L.color[0], L.color[1], L.color[2]);
              if (L.kind == PB2_LIGHT_SPOT) {
                const float c = -dot(lwo, V3(L.direction[0], L.direction[1], L.direction[2]));
                if (c > L.falloff_cos) {}
                else if (c > L.cutoff_cos) le = le * (c - L.cutoff_cos) / (L.falloff_cos - L.cutoff_cos);
                else ok = false;
              }
              lpdf = sqr(ldist);
            }
            if (ok) {
              const float ls_pdf = lpdf / sc.n_lights;
              const float cosine = absdot(lwo, it.n);
              const V3 wo = w2l * lwo;
              const V3 f = bxdf_f(bx, wo);
              if (L.kind != PB2_LIGHT_SKY) nee = le * V3(1.0f) * cosine * f / ls_pdf;  // is_delta: no MIS (path.cpp:104-106)
              else nee = le * V3(1.0f) * cosine * f / ls_pdf * balance_heuristic(ls_pdf, bxdf_pdf(bx, wo));
              if (!is_zero(nee)) {
                shadow_ray = spawn_ray(it.p, it.n, lwo, ldist);
                emit_shadow = true;
              }
            }
          }
        }
      }
      BsdfSample bs;
      if (bxdf_sample(sc, bx, smp, bs)) {
        V3 wo_world = l2w * bs.wo;
        float cosine = absdot(wo_world, it.n);
        store_vertex(w, depth, slot, nee, cosine / bs.pdf * 1.0f, bs.f, beta_code);
        next_ray = spawn_ray(it.p, it.n, wo_world);
        emit_next = true;
        w.pv_pdf[slot] = bs.pdf;
        w.pv_flags[slot] = (pv_length + 1) | ((pv_diffuse + (bs.is_delta ? 0 : 1)) << 8) | (int(bs.is_delta) << 16);
      } else {
        store_vertex(w, depth, slot, nee, 0.0f, V3(0.0f), beta_code);
      }
      w.dim[slot] = smp.dim;
      w.term[slot] = make_float4(0, 0, 0, __int_as_float(depth + 1));
    }
    const int pos = warp_compact(emit_next, &s_count[0]);      // position inside the block's range
    const int spos = warp_compact(emit_shadow, &s_count[1]);
    __syncthreads();
    if (threadIdx.x == 0 && s_count[0]) s_base[0] = atomicAdd(next_count, s_count[0]);
    if (threadIdx.x == 32 && s_count[1]) s_base[1] = atomicAdd(shadow_count, s_count[1]);
    __syncthreads();
    if (emit_next) {
      const int at = s_base[0] + pos;
      w.ray_o[nxt][at] = make_float4(next_ray.o.x, next_ray.o.y, next_ray.o.z, next_ray.tmax);
      w.ray_d[nxt][at] = make_float4(next_ray.d.x, next_ray.d.y, next_ray.d.z, __int_as_float(slot));
    }
    if (emit_shadow) {
      const int at = s_base[1] + spos;
      w.sh_o[at] = make_float4(shadow_ray.o.x, shadow_ray.o.y, shadow_ray.o.z, shadow_ray.tmax);
      w.sh_d[at] = make_float4(shadow_ray.d.x, shadow_ray.d.y, shadow_ray.d.z, __int_as_float(slot));
    }
    __syncthreads();  // s_base / s_count are reused by the next tile
  }
}

// ---- unwind + per-pixel accumulation -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fold_resolve(WaveView w) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w.pixels; i += gridDim.x * blockDim.x) {
    int pix = w.pixel0 + i;
    float4 acc = w.lacc[pix];
    V3 L(acc.x, acc.y, acc.z);
    for (int s = 0; s < w.s_batch; s++) {
      int slot = s * w.pixels + i;
      float4 tm = w.term[slot];
      V3 X(tm.x, tm.y, tm.z);
      int nv = __float_as_int(tm.w);
      for (int k = nv - 1; k >= 0; k--) {
        const float4* v = w.vtx + (size_t(k) * w.n_slots + slot) * 2;
        float4 v0 = v[0], v1 = v[1];
        V3 lo = V3(v0.x, v0.y, v0.z) + X * V3(v1.x, v1.y, v1.z) * v0.w;
        int code = __float_as_int(v1.w);
        V3 beta = code == 0 ? V3(1.0f) : V3(code == 1 ? 3.0f : 0.0f, code == 2 ? 3.0f : 0.0f, code == 3 ? 3.0f : 0.0f);
        V3 c = V3(1.0f) * beta * lo;
        X = V3(pmin(c.x, 8.0f), pmin(c.y, 8.0f), pmin(c.z, 8.0f));
      }
      L += X;
    }
    w.lacc[pix] = make_float4(L.x, L.y, L.z, 0.0f);
  }
}

__global__ void __launch_bounds__(256) k_film(const float4* lacc, float4* film, int pixels, int width, int row_begin,
                                            int band, int parts, int part, int spp, float alpha) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < pixels; i += gridDim.x * blockDim.x) {
    float4 a = lacc[i];
    const int row = film_row(row_begin, band, parts, part, i / width);
    film[size_t(row) * width + i % width] = make_float4(a.x / float(spp), a.y / float(spp), a.z / float(spp), alpha);
  }
}

}  // namespace pb2

// ======================================================================================================
// host driver
// ======================================================================================================
using namespace pb2;

namespace {

int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : dflt;
}
int sm_grid(int work_items, int block, int max_waves_per_sm) {
  long want = (long(work_items) + block - 1) / block;
  long cap = long(kNumSMs) * max_waves_per_sm;
  if (want > cap) want = cap;
  long g = ((want + kNumSMs - 1) / kNumSMs) * kNumSMs;
  return int(g > 0 ? g : kNumSMs);
}

// Host tables of HaltonSampler: the first 1000 primes and their running sums (primes.cpp), the digit
// permutations compute_radical_inverse_permutations(RNG()) makes (lowdiscrepancy.cpp:5-17: identity
// shuffled with `other = i + rng.next32u(count - i)`, lowdiscrepancy.h:53-60), and multInverse
// (sampler.cpp:20-41,61-62).
struct HaltonHost {
  std::vector<int32_t> primes, sums;
  std::vector<uint16_t> perms;
  int inv0 = 0, inv1 = 0;
};
int64_t pmod(int64_t a, int64_t n) { int64_t r = a % n; return r < 0 ? r + n : r; }
void extended_gcd(uint64_t a, uint64_t b, int64_t& gcd, int64_t& x, int64_t& y) {
  const int d = int(a / b), r = int(a - uint64_t(d) * b);
  if (r == 0) { gcd = int64_t(b); x = 0; y = 1; return; }
  int64_t nx, ny;
  extended_gcd(b, uint64_t(r), gcd, nx, ny);
  x = ny;
  y = nx - d * ny;
}
const HaltonHost& halton_host() {
  static const HaltonHost H = [] {
    HaltonHost h;
    for (int n = 2; int(h.primes.size()) < kHaltonPrimes; n++) {
      bool prime = true;
      for (int d = 2; d * d <= n && prime; d++) prime = n % d != 0;
      if (prime) h.primes.push_back(n);
    }
    int32_t sum = 0;
    for (int p : h.primes) { h.sums.push_back(sum); sum += p; }
    h.perms.resize(size_t(sum));
    Rng g = rng_seed(0);
    uint16_t* p = h.perms.data();
    for (int i = 0; i < kHaltonPrimes; i++) {
      const int count = h.primes[i];
      for (int j = 0; j < count; j++) p[j] = uint16_t(j);
      for (int j = 0; j < count; j++) {
        const uint64_t u = rng_next(g);
        const uint32_t other = uint32_t(j) + uint32_t(u ^ (u >> 32)) % uint32_t(count - j);
        std::swap(p[j], p[other]);
      }
      p += count;
    }
    int64_t gcd, x, y;
    extended_gcd(kHaltonScale2, kHaltonScale3, gcd, x, y);
    h.inv0 = int(pmod(x, kHaltonScale3));
    extended_gcd(kHaltonScale3, kHaltonScale2, gcd, x, y);
    h.inv1 = int(pmod(x, kHaltonScale2));
    return h;
  }();
  return H;
}

int floor_log2(int v) {  // psl::log2i == hsb (psl/math.h:81,262)
  int r = 0;
  while (v > 1) { v >>= 1; r++; }
  return r;
}
// Binds the sampler of the next render to the device scene: table set for BlueSobolSampler,
// log2_spp / nbase4_digits for SobolSampler (sampler.h:128-130, SobolSampler::init sampler.cpp:82-85).
// cudaLaunchHostFunc target behind every wave: one more wave of the render is complete on the device.
void CUDART_CB wave_completed(void* arg) {
  pb2_context* ctx = static_cast<pb2_context*>(arg);
  const pb2_progress_fn fn = ctx->progress_fn;
  const int done = ++ctx->progress_done;
  if (fn) fn(float(done) / float(ctx->progress_total > 0 ? ctx->progress_total : 1), ctx->progress_user);
}

int select_sampler(pb2_context* ctx, int kind, int spp, int seed) {
  DScene& sc = ctx->scene;
  sc.sampler_kind = kind;
  sc.sampler_seed = seed;
  sc.log2_spp = floor_log2(spp);
  int res = std::max(sc.camera.width, sc.camera.height), pow2 = 1;
  while (pow2 < res) pow2 <<= 1;  // psl::roundup2
  sc.nbase4_digits = floor_log2(pow2) + (sc.log2_spp + 1) / 2;
  if (kind == PB2_SAMPLER_BLUE) return activate_tables(ctx, spp);
  if (kind == PB2_SAMPLER_HALTON) {
    const HaltonHost& H = halton_host();
    if (!ctx->halton_perms.ptr) {
      PB2_CUDA_CHECK(ctx->halton_primes.upload(H.primes));
      PB2_CUDA_CHECK(ctx->halton_sums.upload(H.sums));
      PB2_CUDA_CHECK(ctx->halton_perms.upload(H.perms));
    }
    sc.halton_primes = ctx->halton_primes.as<int32_t>();
    sc.halton_sums = ctx->halton_sums.as<int32_t>();
    sc.halton_perms = ctx->halton_perms.as<uint16_t>();
    sc.halton_inv0 = H.inv0;
    sc.halton_inv1 = H.inv1;
  }
  return PB2_OK;
}

int check_params(pb2_context* ctx, const pb2_render_params* p) {
  if (!ctx || !p) { set_error("null argument"); return PB2_E_INVALID; }
  if (ctx->device < 0) { set_error("host-only context: no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  if (!ctx->scene_ready) { set_error("pb2_upload_scene has not succeeded on this context"); return PB2_E_STATE; }
  if (p->max_path_length <= 0) {  // path.cpp:12-13
    set_error("`PathIntegrator` expect `max_path_length` to be positive, get " + std::to_string(p->max_path_length));
    return PB2_E_INVALID;
  }
  if (p->max_path_length > 250) { set_error("max_path_length > 250 is not supported"); return PB2_E_UNSUPPORTED; }
  const int h = ctx->scene.camera.height;
  if (p->sampler == PB2_SAMPLER_BLUE) {
    if (p->spp <= 0 || p->spp > 256 || (p->spp & (p->spp - 1))) {
      set_error("spp must be BlueSobolSampler::spp(): a power of two <= 256");
      return PB2_E_INVALID;
    }
  } else if (p->sampler == PB2_SAMPLER_SOBOL || p->sampler == PB2_SAMPLER_UNIFORM || p->sampler == PB2_SAMPLER_HALTON) {
    if (p->spp <= 0) {  // sampler.h:16-17
      set_error("`UniformSampler` should have positive samples per pixel");
      return PB2_E_INVALID;
    }
    if (p->spp > 65536) { set_error("spp > 65536 is not supported"); return PB2_E_UNSUPPORTED; }
  } else {
    set_error("unknown sampler kind " + std::to_string(p->sampler));
    return PB2_E_UNSUPPORTED;
  }
  if (p->row_begin < 0 || p->row_end > h || p->row_begin > p->row_end || p->sample_begin < 0 ||
      p->sample_end > p->spp || p->sample_begin > p->sample_end) {
    set_error("row/sample range out of bounds");
    return PB2_E_INVALID;
  }
  if (p->row_parts > 1 && (p->row_band <= 0 || p->row_part < 0 || p->row_part >= p->row_parts)) {
    set_error("row_band must be positive and row_part in [0, row_parts)");
    return PB2_E_INVALID;
  }
  return PB2_OK;
}

}  // namespace

extern "C" int pb2_render_device(pb2_context* ctx, const pb2_render_params* p, void* cuda_stream,
                                 float** film_rgba_device) {
  int rc = check_params(ctx, p);
  if (rc) return rc;
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  rc = select_sampler(ctx, p->sampler, p->spp, p->sampler_seed);
  if (rc) return rc;
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  const DScene& sc = ctx->scene;
  const int W = sc.camera.width, H = sc.camera.height;
  const int D = p->max_path_length;
  int rows = p->row_end - p->row_begin;
  if (p->row_parts > 1) {  // the rows of the bands this call owns
    int mine = 0;
    for (int k = p->row_part; k * p->row_band < rows; k += p->row_parts) mine += std::min(p->row_band, rows - k * p->row_band);
    rows = mine;
  }
  const long P = long(rows) * W;
  const int n_samples = p->sample_end - p->sample_begin;

  PB2_CUDA_CHECK(ctx->film.reserve(size_t(W) * H * 16));
  if (!(p->flags & PB2_RENDER_KEEP_FILM) || ctx->film_pixels != size_t(W) * H)
    PB2_CUDA_CHECK(cudaMemsetAsync(ctx->film.ptr, 0, size_t(W) * H * 16, st));
  ctx->film_pixels = size_t(W) * H;
  if (film_rgba_device) *film_rgba_device = ctx->film.as<float>();
  ctx->counters.last_render_ms = 0;
  if (P == 0 || n_samples == 0) { PB2_CUDA_CHECK(cudaStreamSynchronize(st)); return PB2_OK; }

  // ---- wave plan ----
  // A wave = rows_per_wave film rows x s_batch samples <= max_slots (pixel, sample) slots. Waves over the
  // same rows run one after the other (they share the pixels' RNG streams and accumulators); waves over
  // different rows are independent, and n_lanes of them are in flight at once (see WavefrontBuffers).
  pb2_ray_record* rec = ctx->record;
  const bool timed = (p->flags & PB2_RENDER_TIME_KERNELS) != 0;
  const long max_slots = std::max(long(W), long(env_int("PB2_MAX_SLOTS", 1 << 24)));
  int n_lanes = rec ? 1 : std::min(kMaxLanes, std::max(1, env_int("PB2_LANES", kDefaultLanes)));
  if (P * n_samples < long(n_lanes) * env_int("PB2_MIN_LANE_SLOTS", int(kMinLaneSlots)) || rows < n_lanes) n_lanes = 1;
  int n_pixel_waves = int((rows + max_slots / W - 1) / (max_slots / W));
  n_pixel_waves = ((n_pixel_waves + n_lanes - 1) / n_lanes) * n_lanes;
  const int rows_per_wave = (rows + n_pixel_waves - 1) / n_pixel_waves;
  n_pixel_waves = (rows + rows_per_wave - 1) / rows_per_wave;
  const long pixels_per_wave = long(rows_per_wave) * W;
  int s_batch = int(std::max(1L, std::min(long(n_samples), max_slots / pixels_per_wave)));
  if (ctx->needs_rng_in_shading) s_batch = 1;  // randf() draws in shading share the per-pixel RNG stream
  const long n_slots = pixels_per_wave * s_batch;
  const int n_vtx = std::max(D - 1, 1);
  const int n_sample_waves = (n_samples + s_batch - 1) / s_batch;
  const int n_waves = n_pixel_waves * n_sample_waves;
  const int cstride = 14 * (D + 1);  // + 8 shade-list counters per depth
  if (rec && n_waves != 1) { set_error("ray record: the render must fit one wave"); return PB2_E_INVALID; }

  if (!ctx->wf) ctx->wf = new WavefrontBuffers();
  WavefrontBuffers& B = *ctx->wf;
  // shade lists: class 0 always, plus one per material kind some geometry uses
  int n_lists = 0, list_of_class[8];
  for (int c = 0; c < 8; c++) list_of_class[c] = -1;
  list_of_class[0] = n_lists++;
  for (int k = 1; k < 8; k++)
    if (ctx->material_kinds & (1u << k)) list_of_class[k] = n_lists++;
  PB2_CUDA_CHECK(B.rng.reserve(size_t(P) * 16));
  PB2_CUDA_CHECK(B.lacc.reserve(size_t(P) * 16));
  PB2_CUDA_CHECK(cudaMemsetAsync(B.lacc.ptr, 0, size_t(P) * 16, st));
  PB2_CUDA_CHECK(B.counts.reserve(size_t(n_waves) * cstride * 4));
  PB2_CUDA_CHECK(cudaMemsetAsync(B.counts.ptr, 0, size_t(n_waves) * cstride * 4, st));
  if (!B.fork) PB2_CUDA_CHECK(cudaEventCreateWithFlags(&B.fork, cudaEventDisableTiming));

  const bool ref_order = p->accel != 0 || !sc.wide_ready;
  const bool closest_wide_ok = !ref_order && !sc.order_dependent;
  const bool any_wide_ok = !ref_order && sc.n_box_shapes <= kMaxBoxShapes;
  const int pgrid = persistent_grid();

  WaveView lane_view[kMaxLanes];
  cudaStream_t lane_stream[kMaxLanes];
  for (int l = 0; l < n_lanes; l++) {
    LaneBuffers& L = B.lane[l];
    for (int k = 0; k < 2; k++) {
      PB2_CUDA_CHECK(L.ray_o[k].reserve(n_slots * 16));
      PB2_CUDA_CHECK(L.ray_d[k].reserve(n_slots * 16));
    }
    PB2_CUDA_CHECK(L.hit.reserve(n_slots * 16));
    PB2_CUDA_CHECK(L.sh_o.reserve(n_slots * 16));
    PB2_CUDA_CHECK(L.sh_d.reserve(n_slots * 16));
    PB2_CUDA_CHECK(L.pv_pdf.reserve(n_slots * 4));
    PB2_CUDA_CHECK(L.pv_flags.reserve(n_slots * 4));
    PB2_CUDA_CHECK(L.dim.reserve(n_slots * 4));
    PB2_CUDA_CHECK(L.term.reserve(n_slots * 16));
    PB2_CUDA_CHECK(L.vtx.reserve(size_t(n_slots) * n_vtx * 32));
    PB2_CUDA_CHECK(L.fb_list.reserve(size_t(n_slots) * 4));
    PB2_CUDA_CHECK(L.lists.reserve(size_t(n_lists) * n_slots * 4));
    if (rec) PB2_CUDA_CHECK(L.occ.reserve(size_t(n_slots)));
    if (l > 0 && !L.stream) PB2_CUDA_CHECK(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
    if (!L.done) PB2_CUDA_CHECK(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
    // the per-kernel timing pass keeps the wave plan but runs the lanes one after the other on the caller's
    // stream, so that the event pairs around a launch measure that launch alone
    lane_stream[l] = (l == 0 || timed) ? st : L.stream;
    WaveView& w = lane_view[l];
    w = WaveView{};
    for (int c = 0; c < 8; c++) w.list_of_class[c] = list_of_class[c];
    w.lists = L.lists.as<int>();
    w.occ = rec ? L.occ.as<uint8_t>() : nullptr;
    for (int k = 0; k < 2; k++) { w.ray_o[k] = L.ray_o[k].as<float4>(); w.ray_d[k] = L.ray_d[k].as<float4>(); }
    w.hit = L.hit.as<float4>(); w.sh_o = L.sh_o.as<float4>(); w.sh_d = L.sh_d.as<float4>();
    w.pv_pdf = L.pv_pdf.as<float>(); w.pv_flags = L.pv_flags.as<int>(); w.dim = L.dim.as<int>();
    w.term = L.term.as<float4>(); w.vtx = L.vtx.as<float4>();
    w.rng = B.rng.as<uint2>(); w.lacc = B.lacc.as<float4>(); w.fb_list = L.fb_list.as<int>();
    w.n_slots = int(n_slots); w.width = W; w.row_begin = p->row_begin; w.max_depth = D;
    w.row_band = p->row_band; w.row_parts = p->row_parts; w.row_part = p->row_part; w.accel = p->accel;
  }

  uint64_t launches = 0;
  // optional per-kernel timing: an event after every launch, classified on the host afterwards
  std::vector<int> ev_class;  // 0 trace, 1 shadow, 2 shade, 3 other
  size_t ev_used = 0;
  auto mark = [&](int cls) {
    if (!timed) return;
    if (ev_used == B.events.size()) { cudaEvent_t e; cudaEventCreate(&e); B.events.push_back(e); }
    cudaEventRecord(B.events[ev_used++], st);
    ev_class.push_back(cls);
  };
  // ray record (test hook): synchronous copies of the queues after the traversal kernels ran
  std::vector<float4> rec_a, rec_b, rec_c;
  std::vector<uint8_t> rec_occ;
  auto record = [&](const WaveView& w, bool shadow, int depth) -> int {
    int n = 0;
    const int* cnt = w.counts + (shadow ? D + 1 + depth : depth);
    if (cudaStreamSynchronize(st) != cudaSuccess ||
        cudaMemcpy(&n, cnt, 4, cudaMemcpyDeviceToHost) != cudaSuccess) { set_error("ray record: copy failed"); return PB2_E_CUDA; }
    uint64_t& have = shadow ? rec->n_shadow : rec->n_closest;
    if (have + uint64_t(n) > rec->capacity) { set_error("ray record: capacity exceeded"); return PB2_E_INVALID; }
    if (!n) return PB2_OK;
    rec_a.resize(n); rec_b.resize(n);
    const float4* src_o = shadow ? w.sh_o : w.ray_o[depth & 1];
    const float4* src_d = shadow ? w.sh_d : w.ray_d[depth & 1];
    cudaMemcpy(rec_a.data(), src_o, size_t(n) * 16, cudaMemcpyDeviceToHost);
    cudaMemcpy(rec_b.data(), src_d, size_t(n) * 16, cudaMemcpyDeviceToHost);
    float* rays = (shadow ? rec->shadow_rays : rec->closest_rays) + have * 8;
    for (int i = 0; i < n; i++) {
      float* r = rays + size_t(i) * 8;
      r[0] = rec_a[i].x; r[1] = rec_a[i].y; r[2] = rec_a[i].z;
      r[3] = rec_b[i].x; r[4] = rec_b[i].y; r[5] = rec_b[i].z;
      r[6] = 0.0f; r[7] = rec_a[i].w;
    }
    if (shadow) {
      if (cudaMemcpy(rec->shadow_occluded + have, w.occ, size_t(n), cudaMemcpyDeviceToHost) != cudaSuccess) {
        set_error("ray record: copy failed"); return PB2_E_CUDA;
      }
    } else {
      rec_c.resize(n);
      if (cudaMemcpy(rec_c.data(), w.hit, size_t(n) * 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
        set_error("ray record: copy failed"); return PB2_E_CUDA;
      }
      for (int i = 0; i < n; i++) {
        rec->closest_t[have + i] = rec_c[i].x;
        std::memcpy(&rec->closest_geom[have + i], &rec_c[i].y, 4);
        std::memcpy(&rec->closest_prim[have + i], &rec_c[i].z, 4);
        if (rec->closest_geom[have + i] < 0) rec->closest_prim[have + i] = -1;
      }
    }
    have += uint64_t(n);
    return PB2_OK;
  };
  // L2 persistence window over the top of the wide BVH (nodes8 is breadth first, so its prefix is the
  // top levels): node fetches that hit keep their lines, everything else in the window's stream
  // (ray queues, hit records, vertex records) is marked streaming. PB2_L2_PERSIST_MB = 0 turns it off.
  {
    static const int persist_mb = env_int("PB2_L2_PERSIST_MB", 0);
    size_t bytes = std::min(size_t(persist_mb) << 20, ctx->nodes8.cap);
    cudaStreamAttrValue attr{};
    if (bytes && sc.wide_ready) {
      int max_window = 0, max_persist = 0;
      cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
      cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device);
      bytes = std::min(bytes, std::min(size_t(max_window), size_t(max_persist)));
      if (bytes) {
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes);
        attr.accessPolicyWindow.base_ptr = const_cast<DNode8*>(sc.nodes8);
        attr.accessPolicyWindow.num_bytes = bytes;
        attr.accessPolicyWindow.hitRatio = 1.0f;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      }
    }
    if (bytes || ctx->l2_window_set) {
      for (int l = 0; l < n_lanes; l++)
        cudaStreamSetAttribute(lane_stream[l], cudaStreamAttributeAccessPolicyWindow, &attr);
      ctx->l2_window_set = bytes != 0;
      cudaGetLastError();  // best effort: a device without the feature just ignores it
    }
  }
  PB2_CUDA_CHECK(cudaEventRecord(ctx->ev0, st));
  mark(-1);
  // One wave on one lane: raygen, then per depth trace / classify / shade / shadow, then the unwind.
  // The first instantiated shape mask (shapes.cuh) that contains the scene's geometry kinds: the code compiled into
  // the persistent loop decides its register allocation. The closest-hit kernel skips the Rect + Mesh form (2.5%
  // slower in kernel on C5's deep trees than the no-box form: profiles/r2/s28_s31_*, s34_*).
  auto shapes_level = [&](bool closest) {
    for (int m : {kShapesRect, kShapesRectMesh, kShapesFlat, kShapesCbox, kShapesNoBox, kShapesCommon})
      if ((sc.shape_mask & ~m) == 0 && !(m == kShapesRectMesh && closest)) return m;
    return kShapesAll;
  };
  int trace_shapes = shapes_level(true), shadow_shapes = shapes_level(false);
  {  // tuning knobs: force another instantiated mask (must still contain the scene's kinds)
    const int ft = env_int("PB2_TRACE_SHAPES", -1), fs = env_int("PB2_SHADOW_SHAPES", -1);
    if (ft >= 0 && (sc.shape_mask & ~ft) == 0) trace_shapes = ft;
    if (fs >= 0 && (sc.shape_mask & ~fs) == 0) shadow_shapes = fs;
  }
  auto enqueue_wave = [&](int l, int pw, int sw) -> int {
    WaveView& w = lane_view[l];
    cudaStream_t ls = lane_stream[l];
    w.pixel0 = int(long(pw) * pixels_per_wave);
    w.pixels = int(std::min(pixels_per_wave, P - w.pixel0));
    w.sample0 = p->sample_begin + sw * s_batch;
    w.s_batch = std::min(s_batch, p->sample_end - w.sample0);
    w.counts = B.counts.as<int>() + size_t(pw * n_sample_waves + sw) * cstride;
    const int live = w.pixels * w.s_batch;
    k_raygen<<<sm_grid(w.pixels, 256, 8), 256, 0, ls>>>(sc, w, p->sample_begin, sw == 0);
    launches++; mark(3);
    for (int depth = 0; depth < D; depth++) {
      if (closest_wide_ok) {
        switch (trace_shapes) {
#define PB2_TRACE(M) case M: k_trace<M><<<pgrid, kTraceBlock, 0, ls>>>(sc, w, depth); break;
          PB2_TRACE(kShapesRect) PB2_TRACE(kShapesRectMesh) PB2_TRACE(kShapesFlat) PB2_TRACE(kShapesNoBox) PB2_TRACE(kShapesCbox) PB2_TRACE(kShapesCommon)
          default: k_trace<kShapesAll><<<pgrid, kTraceBlock, 0, ls>>>(sc, w, depth);
#undef PB2_TRACE
        }
        // the reference-order re-trace: the README Cornell box sends a fifth of its rays here (its two Boxes), so
        // it gets its own mask; everything else the common one
        if (sc.has_rare_shapes) k_trace_fallback<kShapesAll><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, ls>>>(sc, w, depth);
        else if ((sc.shape_mask & ~kShapesCbox) == 0) k_trace_fallback<kShapesCbox><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, ls>>>(sc, w, depth);
        else k_trace_fallback<kShapesCommon><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, ls>>>(sc, w, depth);
        launches += 2;
      } else {
        if (sc.has_rare_shapes) k_trace_ref<true><<<sm_grid(live, kTraceBlock, 16), kTraceBlock, 0, ls>>>(sc, w, depth);
        else k_trace_ref<false><<<sm_grid(live, kTraceBlock, 16), kTraceBlock, 0, ls>>>(sc, w, depth);
        launches++;
      }
      mark(0);
      if (rec) { int rc2 = record(w, false, depth); if (rc2) return rc2; }
      k_classify<<<sm_grid(live, 256, 8), 256, 0, ls>>>(sc, w, depth);
      k_shade_terminal<<<sm_grid(live, 256, 8), 256, 0, ls>>>(sc, w, depth);
      launches += 2;
      if (depth + 1 < D) {
        const int sg = sm_grid(live, 128, 16);
        const unsigned mk = ctx->material_kinds;
        const bool blue = sc.sampler_kind == PB2_SAMPLER_BLUE && ctx->only_area_lights && !ctx->has_node_programs;
#define PB2_SHADE(MAT)                                                      \
  if (mk & (1u << MAT)) {                                                   \
    if (blue) k_shade_mat<MAT, false><<<sg, 128, 0, ls>>>(sc, w, depth);    \
    else k_shade_mat<MAT, true><<<sg, 128, 0, ls>>>(sc, w, depth);          \
    launches++;                                                             \
  }
        PB2_SHADE(PB2_MAT_DIFFUSE) PB2_SHADE(PB2_MAT_METAL) PB2_SHADE(PB2_MAT_GLOSSY)
        PB2_SHADE(PB2_MAT_GLASS) PB2_SHADE(PB2_MAT_UBER) PB2_SHADE(PB2_MAT_SUBSURFACE)
#undef PB2_SHADE
      }
      mark(2);
      if (depth + 1 < D) {  // the last vertex never samples a light (path.cpp:89)
        if (any_wide_ok) {
          switch (shadow_shapes) {
#define PB2_SHADOW(M) case M: k_shadow<M><<<pgrid, kTraceBlock, 0, ls>>>(sc, w, depth); break;
            PB2_SHADOW(kShapesRect) PB2_SHADOW(kShapesRectMesh) PB2_SHADOW(kShapesFlat) PB2_SHADOW(kShapesNoBox) PB2_SHADOW(kShapesCbox) PB2_SHADOW(kShapesCommon)
            default: k_shadow<kShapesAll><<<pgrid, kTraceBlock, 0, ls>>>(sc, w, depth);
#undef PB2_SHADOW
          }
          if (sc.has_rare_shapes) k_shadow_fallback<true><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, ls>>>(sc, w, depth);
          else k_shadow_fallback<false><<<kNumSMs * kFallbackBlocksPerSM, kTraceBlock, 0, ls>>>(sc, w, depth);
          launches += 2;
        } else {
          if (sc.has_rare_shapes) k_shadow_ref<true><<<sm_grid(live, kTraceBlock, 16), kTraceBlock, 0, ls>>>(sc, w, depth);
          else k_shadow_ref<false><<<sm_grid(live, kTraceBlock, 16), kTraceBlock, 0, ls>>>(sc, w, depth);
          launches++;
        }
        mark(1);
        if (rec) { int rc2 = record(w, true, depth); if (rc2) return rc2; }
      }
    }
    k_fold_resolve<<<sm_grid(w.pixels, 256, 8), 256, 0, ls>>>(w);
    launches++; mark(3);
    return PB2_OK;
  };
  // fork: the lanes' own streams start after the clears above
  bool forked = false;
  for (int l = 1; l < n_lanes; l++)
    if (lane_stream[l] != st) {
      if (!forked) { PB2_CUDA_CHECK(cudaEventRecord(B.fork, st)); forked = true; }
      PB2_CUDA_CHECK(cudaStreamWaitEvent(lane_stream[l], B.fork, 0));
    }
  // pixel wave pw belongs to lane pw % n_lanes; launches are interleaved wave by wave so that every
  // lane's stream stays fed whatever the driver's launch-queue depth
  ctx->progress_done = 0;
  ctx->progress_total = n_waves;
  for (int pw0 = 0; pw0 < n_pixel_waves; pw0 += n_lanes)
    for (int sw = 0; sw < n_sample_waves; sw++)
      for (int l = 0; l < n_lanes && pw0 + l < n_pixel_waves; l++) {
        if ((rc = enqueue_wave(l, pw0 + l, sw))) return rc;
        if (ctx->progress_fn) PB2_CUDA_CHECK(cudaLaunchHostFunc(lane_stream[l], wave_completed, ctx));
      }
  // join
  for (int l = 1; l < n_lanes; l++)
    if (lane_stream[l] != st) {
      PB2_CUDA_CHECK(cudaEventRecord(B.lane[l].done, lane_stream[l]));
      PB2_CUDA_CHECK(cudaStreamWaitEvent(st, B.lane[l].done, 0));
    }
  const float alpha = float(n_samples) / float(p->spp);
  k_film<<<sm_grid(int(P), 256, 8), 256, 0, st>>>(B.lacc.as<float4>(), ctx->film.as<float4>(), int(P), W,
                                                  p->row_begin, p->row_band, p->row_parts, p->row_part, p->spp, alpha);
  launches++; mark(3);
  PB2_CUDA_CHECK(cudaEventRecord(ctx->ev1, st));
  PB2_CUDA_CHECK(cudaGetLastError());
  B.host_counts.resize(size_t(n_waves) * cstride);
  PB2_CUDA_CHECK(cudaMemcpyAsync(B.host_counts.data(), B.counts.ptr, B.host_counts.size() * 4,
                                 cudaMemcpyDeviceToHost, st));
  PB2_CUDA_CHECK(cudaStreamSynchronize(st));
  float ms = 0;
  PB2_CUDA_CHECK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->counters.last_render_ms = ms;
  double cls_ms[4] = {0, 0, 0, 0};
  uint64_t trace_launches = 0;
  for (size_t i = 1; i < ev_used; i++) {
    float e = 0;
    cudaEventElapsedTime(&e, B.events[i - 1], B.events[i]);
    cls_ms[ev_class[i]] += e;
    trace_launches += ev_class[i] == 0;
  }
  ctx->counters.last_trace_ms = cls_ms[0];
  ctx->counters.last_shadow_ms = cls_ms[1];
  ctx->counters.last_shade_ms = cls_ms[2];
  ctx->counters.last_other_ms = cls_ms[3];
  ctx->counters.last_trace_launches = trace_launches;
  uint64_t closest = 0, shadow = 0, fb_closest = 0, fb_shadow = 0;
  for (int wv = 0; wv < n_waves; wv++)
    for (int k = 0; k <= D; k++) {
      closest += uint64_t(B.host_counts[size_t(wv) * cstride + k]);
      shadow += uint64_t(B.host_counts[size_t(wv) * cstride + D + 1 + k]);
      fb_closest += uint64_t(B.host_counts[size_t(wv) * cstride + 2 * (D + 1) + 4 * k + 1]);
      fb_shadow += uint64_t(B.host_counts[size_t(wv) * cstride + 2 * (D + 1) + 4 * k + 3]);
    }
  ctx->counters.last_closest_fallbacks = fb_closest;
  ctx->counters.last_shadow_fallbacks = fb_shadow;
  ctx->counters.closest_rays += closest;
  ctx->counters.shadow_rays += shadow;
  ctx->counters.last_closest_rays = closest;
  ctx->counters.last_shadow_rays = shadow;
  ctx->counters.camera_samples += uint64_t(P) * n_samples;
  ctx->counters.kernel_launches += launches;
  return PB2_OK;
}

// Test hook: the uploaded scene's node program starting at `start` (1-based) evaluated at n points.
namespace pb2 {
__global__ void k_eval_node(DScene sc, int start, int n, const float* __restrict__ ctx8, float* __restrict__ out3) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* c = ctx8 + size_t(i) * 8;
  const V3 v = eval_node_program(sc, start, V3(c[0], c[1], c[2]), V3(c[3], c[4], c[5]), V2(c[6], c[7]));
  out3[size_t(i) * 3] = v.x; out3[size_t(i) * 3 + 1] = v.y; out3[size_t(i) * 3 + 2] = v.z;
}
}  // namespace pb2
extern "C" int pb2_test_eval_node(pb2_context* ctx, int start, size_t n, const float* ctx8_host, float* out3_host) {
  if (!ctx || !ctx8_host || !out3_host) { set_error("null argument"); return PB2_E_INVALID; }
  if (ctx->device < 0) { set_error("host-only context: no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  if (!ctx->scene_ready) { set_error("pb2_upload_scene has not succeeded on this context"); return PB2_E_STATE; }
  const HostScene& H = ctx->host_view ? *ctx->host_view : ctx->host;
  bool is_start = start >= 1 && size_t(start) <= H.node_instrs.size() &&
                  (start == 1 || H.node_instrs[size_t(start) - 2].op == PB2_NODE_END);
  if (!is_start) { set_error("pb2_test_eval_node: not the start of a program of the uploaded scene"); return PB2_E_INVALID; }
  if (n == 0) return PB2_OK;
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  PB2_CUDA_CHECK(ctx->rays.reserve(n * 32));
  PB2_CUDA_CHECK(ctx->out_surf.reserve(n * 12));
  PB2_CUDA_CHECK(cudaMemcpyAsync(ctx->rays.ptr, ctx8_host, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  k_eval_node<<<unsigned((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->scene, start, int(n), ctx->rays.as<float>(), ctx->out_surf.as<float>());
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaMemcpyAsync(out3_host, ctx->out_surf.ptr, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  return PB2_OK;
}

extern "C" int pb2_set_progress_callback(pb2_context* ctx, pb2_progress_fn fn, void* user) {
  if (!ctx) { set_error("null context"); return PB2_E_INVALID; }
  ctx->progress_fn = fn;
  ctx->progress_user = user;
  return PB2_OK;
}

extern "C" int pb2_render(pb2_context* ctx, const pb2_render_params* p, float* film_rgba_host) {
  if (!film_rgba_host) { set_error("null film pointer"); return PB2_E_INVALID; }
  float* dev = nullptr;
  int rc = pb2_render_device(ctx, p, nullptr, &dev);
  if (rc) return rc;
  const size_t bytes = size_t(ctx->scene.camera.width) * ctx->scene.camera.height * 16;
  PB2_CUDA_CHECK(cudaMemcpyAsync(film_rgba_host, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  return PB2_OK;
}

// ---- film epilogue: Film::finalize + to_uint8 (film.cpp:21-27, color.cpp:6-23, fileio.cpp:42-54) --------
namespace pb2 {
__device__ __forceinline__ V3 uncharted2_map(V3 x) {
  const V3 A(0.15f), B(0.50f), C(0.10f), D(0.20f), E(0.02f), F(0.30f);
  return (x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F) - E / F;
}
__global__ void k_finalize(const float4* film, int w, int h, int tonemapper, int flip_y, float4* mapped,
                           uchar4* rgba8) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w * h; i += gridDim.x * blockDim.x) {
    float4 px = film[i];
    V3 v = V3(px.x, px.y, px.z) * 1.0f;  // finalize(scale = 1)
    V3 c;
    if (tonemapper == 0) {
      c = uncharted2_map(v * 2.0f) * V3(1.0f) / uncharted2_map(V3(11.2f));
    } else {
      const float a = 2.51f, b = 0.03f, cc = 2.43f, d = 0.59f, e = 0.14f;
      V3 r = (v * (a * v + V3(b))) / (v * (cc * v + V3(d)) + V3(e));
      c = V3(pclamp(r.x, 0.0f, 1.0f), pclamp(r.y, 0.0f, 1.0f), pclamp(r.z, 0.0f, 1.0f));
    }
    if (mapped) mapped[i] = make_float4(c.x, c.y, c.z, 1.0f);
    if (rgba8) {
      int x = i % w, y = i / w;
      int o = flip_y ? (h - 1 - y) * w + x : i;
      auto q = [](float f) { return (unsigned char)pclamp(powf(f, 1.0f / 2.2f) * 256.0f, 0.0f, 255.0f); };
      rgba8[o] = make_uchar4(q(c.x), q(c.y), q(c.z), q(1.0f));
    }
  }
}

// ---- known-answer kernels ---------------------------------------------------------------------------------
__global__ void k_test_sampler(DScene sc, int px, int py, int sample_index, int n_rand, float* out_rand,
                               const char* pattern, int n_pattern, int n_samples, float* out_draws) {
  if (blockIdx.x || threadIdx.x) return;
  Rng g = rng_seed(murmur_pixel(px, py, sample_index));
  for (int i = 0; i < n_rand; i++) out_rand[i] = rng_nextf(g);
  PixelSampler s{px, py, sample_index, 0};
  for (int k = 0; k < n_samples; k++) {
    for (int c = 0; c < n_pattern; c++) {
      if (pattern[c] == '1') {
        *out_draws++ = smp_get1d(sc, s);
      } else {
        V2 u = smp_get2d(sc, s);
        *out_draws++ = u.x;
        *out_draws++ = u.y;
      }
    }
    s.dim = 0;  // start_next_sample (sampler.h:178-181)
    s.index++;
  }
}
__global__ void k_test_primary(DScene sc, int si, float4* rays) {
  const int W = sc.camera.width, H = sc.camera.height;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < W * H; i += gridDim.x * blockDim.x) {
    int px = i % W, py = i / W;
    Rng g = rng_seed(murmur_pixel(px, py, 0));
    Ray r;
    for (int s = 0; s <= si; s++) {
      float r2 = rng_nextf(g), r3 = rng_nextf(g), r0 = rng_nextf(g), r1 = rng_nextf(g);
      r = camera_ray(sc.camera, V2((float(px) + r0) / float(W), (float(py) + r1) / float(H)), V2(r2, r3));
    }
    rays[2 * i] = make_float4(r.o.x, r.o.y, r.o.z, r.d.x);
    rays[2 * i + 1] = make_float4(r.d.y, r.d.z, r.tmin, r.tmax);
  }
}
}  // namespace pb2

extern "C" int pb2_film_finalize(pb2_context* ctx, const float* film_rgba_host, int flip_y,
                                 uint8_t* rgba8_host, float* tonemapped_rgba_host) {
  if (!ctx || ctx->device < 0 || !ctx->scene_ready) { set_error("no device scene"); return PB2_E_STATE; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  const int W = ctx->scene.camera.width, H = ctx->scene.camera.height;
  const size_t n = size_t(W) * H;
  DBuf in, mapped, q;
  PB2_CUDA_CHECK(in.reserve(n * 16));
  PB2_CUDA_CHECK(mapped.reserve(n * 16));
  PB2_CUDA_CHECK(q.reserve(n * 4));
  PB2_CUDA_CHECK(cudaMemcpyAsync(in.ptr, film_rgba_host, n * 16, cudaMemcpyHostToDevice, ctx->stream));
  k_finalize<<<kNumSMs * 4, 256, 0, ctx->stream>>>(in.as<float4>(), W, H, ctx->scene.camera.tonemapper, flip_y,
                                                   mapped.as<float4>(), q.as<uchar4>());
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  if (rgba8_host) PB2_CUDA_CHECK(cudaMemcpyAsync(rgba8_host, q.ptr, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (tonemapped_rgba_host)
    PB2_CUDA_CHECK(cudaMemcpyAsync(tonemapped_rgba_host, mapped.ptr, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  return PB2_OK;
}

// The same epilogue on the film the last render LEFT ON THE DEVICE (after the multi-GPU reduce: the first
// device's film): nothing is uploaded, and 4 bytes per pixel come back instead of 16.
extern "C" int pb2_film_finalize_device(pb2_context* ctx, int tonemapper, int flip_y, uint8_t* rgba8_host,
                                        float* tonemapped_rgba_host) {
  if (!ctx || ctx->device < 0 || !ctx->scene_ready) { set_error("no device scene"); return PB2_E_STATE; }
  const int W = ctx->scene.camera.width, H = ctx->scene.camera.height;
  const size_t n = size_t(W) * H;
  if (!ctx->film.ptr || ctx->film_pixels != n) { set_error("no film on the device: render first"); return PB2_E_STATE; }
  if (tonemapper < 0) tonemapper = ctx->scene.camera.tonemapper;
  if (tonemapper > 1) { set_error("unknown tone mapper"); return PB2_E_INVALID; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  const size_t q_bytes = ((n * 4 + 15) / 16) * 16;  // RGBA8 first, then (optionally) the tone-mapped float film
  PB2_CUDA_CHECK(ctx->film_accum.reserve(q_bytes + (tonemapped_rgba_host ? n * 16 : 0)));
  uchar4* q = ctx->film_accum.as<uchar4>();
  float4* mapped = tonemapped_rgba_host ? reinterpret_cast<float4*>(static_cast<char*>(ctx->film_accum.ptr) + q_bytes) : nullptr;
  k_finalize<<<kNumSMs * 4, 256, 0, ctx->stream>>>(ctx->film.as<float4>(), W, H, tonemapper, flip_y, mapped, q);
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  if (rgba8_host) PB2_CUDA_CHECK(cudaMemcpyAsync(rgba8_host, q, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (tonemapped_rgba_host)
    PB2_CUDA_CHECK(cudaMemcpyAsync(tonemapped_rgba_host, mapped, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  return PB2_OK;
}

extern "C" int pb2_test_sampler(pb2_context* ctx, int spp, int px, int py, int sample_index, int n_rand,
                                float* out_rand, const char* pattern, int n_samples, float* out_draws) {
  if (!ctx || ctx->device < 0) { set_error("no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  int rc = select_sampler(ctx, PB2_SAMPLER_BLUE, spp, 0);
  if (rc) return rc;
  int n_pattern = int(strlen(pattern)), width = 0;
  for (int i = 0; i < n_pattern; i++) width += pattern[i] == '1' ? 1 : 2;
  DBuf d_rand, d_draws, d_pat;
  PB2_CUDA_CHECK(d_rand.reserve(size_t(n_rand + 1) * 4));
  PB2_CUDA_CHECK(d_draws.reserve(size_t(width) * n_samples * 4 + 4));
  PB2_CUDA_CHECK(d_pat.reserve(n_pattern + 1));
  PB2_CUDA_CHECK(cudaMemcpy(d_pat.ptr, pattern, n_pattern, cudaMemcpyHostToDevice));
  k_test_sampler<<<1, 1, 0, ctx->stream>>>(ctx->scene, px, py, sample_index, n_rand, d_rand.as<float>(),
                                           d_pat.as<char>(), n_pattern, n_samples, d_draws.as<float>());
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (n_rand) PB2_CUDA_CHECK(cudaMemcpy(out_rand, d_rand.ptr, size_t(n_rand) * 4, cudaMemcpyDeviceToHost));
  if (width * n_samples)
    PB2_CUDA_CHECK(cudaMemcpy(out_draws, d_draws.ptr, size_t(width) * n_samples * 4, cudaMemcpyDeviceToHost));
  return PB2_OK;
}

extern "C" int pb2_test_sampler_kind(pb2_context* ctx, int sampler, int spp, int sampler_seed, int px, int py,
                                     int sample_index, const char* pattern, int n_samples, float* out_draws) {
  if (!ctx || ctx->device < 0) { set_error("no CUDA device, and there is no CPU fallback"); return PB2_E_CUDA; }
  if (!ctx->scene_ready) { set_error("pb2_upload_scene has not succeeded on this context"); return PB2_E_STATE; }
  pb2_render_params chk{};
  chk.spp = spp; chk.max_path_length = 1; chk.row_end = ctx->scene.camera.height; chk.sample_end = spp;
  chk.sampler = sampler;
  int rc = check_params(ctx, &chk);
  if (rc) return rc;
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  rc = select_sampler(ctx, sampler, spp, sampler_seed);
  if (rc) return rc;
  int n_pattern = int(strlen(pattern)), width = 0;
  for (int i = 0; i < n_pattern; i++) width += pattern[i] == '1' ? 1 : 2;
  DBuf d_rand, d_draws, d_pat;
  PB2_CUDA_CHECK(d_rand.reserve(4));
  PB2_CUDA_CHECK(d_draws.reserve(size_t(width) * n_samples * 4 + 4));
  PB2_CUDA_CHECK(d_pat.reserve(n_pattern + 1));
  PB2_CUDA_CHECK(cudaMemcpy(d_pat.ptr, pattern, n_pattern, cudaMemcpyHostToDevice));
  k_test_sampler<<<1, 1, 0, ctx->stream>>>(ctx->scene, px, py, sample_index, 0, d_rand.as<float>(),
                                           d_pat.as<char>(), n_pattern, n_samples, d_draws.as<float>());
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (width * n_samples)
    PB2_CUDA_CHECK(cudaMemcpy(out_draws, d_draws.ptr, size_t(width) * n_samples * 4, cudaMemcpyDeviceToHost));
  return PB2_OK;
}

// Host evaluation of the SobolSampler / UniformSampler draw functions above (the same __host__ __device__
// code the kernels inline): lets the CPU test-suite pin their integer arithmetic against the reference.
extern "C" int pb2h_sampler_draws(int sampler, int spp, int sampler_seed, int width, int height, int px, int py,
                                  int sample_index, const char* pattern, int n_samples, float* out_draws) {
  if (sampler != PB2_SAMPLER_SOBOL && sampler != PB2_SAMPLER_UNIFORM && sampler != PB2_SAMPLER_HALTON) {
    set_error("pb2h_sampler_draws: Sobol / Uniform / Halton samplers only (BlueSampler reads device tables)");
    return PB2_E_UNSUPPORTED;
  }
  if (spp <= 0 || !pattern || !out_draws) { set_error("invalid argument"); return PB2_E_INVALID; }
  const int log2_spp = floor_log2(spp);
  int pow2 = 1;
  while (pow2 < std::max(width, height)) pow2 <<= 1;
  const int nb4 = floor_log2(pow2) + (log2_spp + 1) / 2;
  for (int k = 0; k < n_samples; k++) {
    PixelSampler s{px, py, sample_index + k, 0};
    for (const char* c = pattern; *c; c++) {
      if (sampler == PB2_SAMPLER_SOBOL) {
        if (*c == '1') { *out_draws++ = sobol_get1d(s, log2_spp, nb4); }
        else { V2 u = sobol_get2d(s, log2_spp, nb4); *out_draws++ = u.x; *out_draws++ = u.y; }
      } else if (sampler == PB2_SAMPLER_HALTON) {
        const HaltonHost& H = halton_host();
        const HaltonTables t{H.primes.data(), H.sums.data(), H.perms.data()};
        if (*c == '1') { *out_draws++ = halton_get1d(t, s, H.inv0, H.inv1); }
        else { V2 u = halton_get2d(t, s, H.inv0, H.inv1); *out_draws++ = u.x; *out_draws++ = u.y; }
      } else {
        const int n = *c == '1' ? 1 : 2;
        for (int i = 0; i < n; i++) *out_draws++ = uniform_dim(s, sampler_seed, s.dim++);
      }
    }
  }
  return PB2_OK;
}

extern "C" int pb2_test_primary_rays(pb2_context* ctx, int sample_index, float* rays_host) {
  if (!ctx || ctx->device < 0 || !ctx->scene_ready) { set_error("no device scene"); return PB2_E_STATE; }
  PB2_CUDA_CHECK(cudaSetDevice(ctx->device));
  const size_t n = size_t(ctx->scene.camera.width) * ctx->scene.camera.height;
  PB2_CUDA_CHECK(ctx->rays.reserve(n * 32));
  k_test_primary<<<kNumSMs * 4, 256, 0, ctx->stream>>>(ctx->scene, sample_index, ctx->rays.as<float4>());
  ctx->counters.kernel_launches++;
  PB2_CUDA_CHECK(cudaGetLastError());
  PB2_CUDA_CHECK(cudaMemcpyAsync(rays_host, ctx->rays.ptr, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PB2_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  return PB2_OK;
}
