// wide_persistent.cuh -- the production traversal kernels' inner loop.
//
// Ray lengths in a path tracer are heavy tailed (ncu on the one-ray-per-thread version: 2.4 of 32
// lanes active per instruction in the shadow kernel), so the walk of traverse.cuh (2) is run as a
// per-lane state machine inside PERSISTENT warps: a lane that finishes its ray takes the next ray
// of the queue (one atomicAdd per warp, ranks from a ballot) instead of idling until the slowest
// lane of its warp is done. One iteration = at most one wide-node test plus the primitive tests it
// produced. Rays whose result the wide walk cannot certify (tie, gate failure, stack overflow --
// see traverse.cuh) are appended to a fallback list and re-traced in reference order by a second,
// tiny kernel, which keeps the slow path out of the hot loop's registers.
//
// IO concept (POD, passed by value):
//   __device__ Ray  load(int i) const;
//   __device__ void store_closest(const DScene&, int i, const Ray& ray_after, bool hit, Hit h) const;
//   __device__ void store_any(int i, bool occluded) const;
#pragma once
#include "traverse.cuh"

namespace pb2 {

#ifndef PB2_REFILL
#define PB2_REFILL 8
#endif
constexpr int kRefillThreshold = PB2_REFILL;  // refill a warp when at least this many lanes are idle
#ifndef PB2_TIGHT_LANES
#define PB2_TIGHT_LANES 16
#endif
#ifndef PB2_BATCH_FINISH
#define PB2_BATCH_FINISH 1  // 1: finished lanes certify and store their result together, when the warp refills
#endif

struct WideQueue {
  int* next;      // next unclaimed queue index (zero before the launch)
  int* fb_count;  // number of entries in fb_list (zero before the launch)
  int* fb_list;   // queue indices that need the reference-order walk
};

// Scheduling inside a warp (measured on B200, see DESIGN.md "what did not work"). One iteration =
// at most one wide-node test per lane, then primitive tests. After a node test only ~3 of 32 lanes
// of an incoherent warp own primitives, so those lanes test ONE primitive and go back to walking
// with the rest (a lane's remaining primitives wait in `tgroup`); when at least half the warp owns
// primitives (coherent rays, small scenes) every lane drains its whole group in a tight loop.
// Making lanes WAIT until a batch of primitive owners has formed was tried and lost 7% (fewer
// walking lanes per node test); `prefetch.global.L1` of the next primitive record gained 1.6% on
// the 10 M-triangle scene and cost 2x on scenes that live in L1, so neither is here.
// The order in which a ray's primitives are tested does not matter for the result (ties are
// detected and sent to the fallback, see traverse.cuh).
template <bool ANY, int RARE = kShapesAll, class IO>  // RARE: a kShapes* mask (shapes.cuh)
__device__ __forceinline__ void wide_persistent(const DScene& sc, int n, WideQueue wq, const IO& io) {
  const unsigned lane = threadIdx.x & 31u;
#if PB2_SMEM_TOP > 0
  __shared__ uint4 s_top[PB2_SMEM_TOP * 5];
  {
    const uint4* g = reinterpret_cast<const uint4*>(sc.nodes8);
    const int n5 = (sc.n_nodes8 < PB2_SMEM_TOP ? sc.n_nodes8 : PB2_SMEM_TOP) * 5;  // indices >= n_nodes8 never occur
    for (int i = threadIdx.x; i < n5; i += blockDim.x) s_top[i] = __ldg(g + i);
    __syncthreads();
  }
#else
  const uint4* s_top = nullptr;
#endif
  uint2 stack[kWideStack];
  int sp = 0;
#if PB2_BATCH_FINISH
  bool pending = false;
#endif
  bool active = false, drained = false;
  int idx = 0;
  Ray ray;
  Octant q;
  WideRay w;
  uint2 ngroup = make_uint2(0u, 0u);  // invariant while walking: has a node bit (y > 0x00ffffff)
  uint2 tgroup = make_uint2(0u, 0u);  // primitives waiting to be tested (y = 24-bit mask)
  bool have = false, ambiguous = false;
  float tie_t = -1.0f, tmax_in = 0.0f, t2 = kFloatMax;
  int win_leaf = -1;
  Hit h;
  h.geom = h.prim = -1;
#ifdef PB2_COUNT
  unsigned long long c_nodes = 0, c_prims = 0, c_rays = 0;
#endif

  // certification of a finished ray and its store (or its push onto the fallback list)
  auto finish = [&]() {
    if (ANY) {
      if (ambiguous) wq.fb_list[atomicAdd(wq.fb_count, 1)] = idx;
      else io.store_any(idx, have);
    } else {
#ifdef PB2_COUNT
      if (ambiguous && sc.walk_stats) atomicAdd(sc.walk_stats + 7, 1ull);
#endif
      if (have && !ambiguous) {
        if (tie_t == ray.tmax) ambiguous = true;
        else if (!leaf_gate(sc.leaves, win_leaf, q, ray.tmin, ray.tmax)) {
          // The exact leaf box can come out a few ulps BEHIND the hit it contains (flat boxes of
          // Rects and axis-aligned triangles: two different float expressions for one distance).
          // The reference still reaches the leaf if its tmax at that moment is large enough, and
          // that tmax is tmax_in or the t of another primitive hit before: >= U (traverse.cuh (ii')).
          const float U = pmin(pmin(tmax_in, t2), half_slack_of(ray.tmax));
          if (!leaf_gate(sc.leaves, win_leaf, q, ray.tmin, U)) ambiguous = true;
        }
#ifdef PB2_COUNT
        if (ambiguous && sc.walk_stats) atomicAdd(sc.walk_stats + (tie_t == ray.tmax ? 8 : 9), 1ull);
#endif
      }
      if (ambiguous) wq.fb_list[atomicAdd(wq.fb_count, 1)] = idx;
      else io.store_closest(sc, idx, ray, have, h);
    }
  };

  while (true) {
    // ---- refill idle lanes ----
    const unsigned idle = __ballot_sync(0xffffffffu, !active);
#if PB2_BATCH_FINISH
    if (idle == 0xffffffffu || (idle && !drained && __popc(idle) >= kRefillThreshold)) {
      if (pending) { finish(); pending = false; }
    }
#endif
    if (idle && !drained && (__popc(idle) >= kRefillThreshold || idle == 0xffffffffu)) {
      int base = 0;
      const int want = __popc(idle);
      if (lane == unsigned(__ffs(idle) - 1)) base = atomicAdd(wq.next, want);
      base = __shfl_sync(0xffffffffu, base, __ffs(idle) - 1);
      if (base + want >= n) drained = true;
      if (!active) {
        const int my = base + __popc(idle & ((1u << lane) - 1u));
        if (my < n) {
          idx = my;
          ray = io.load(idx);
          q = make_octant(ray);
          w = make_wide_ray(sc, ray, q);
          tmax_in = ray.tmax;
          ngroup = make_uint2(0u, 0x80000000u);
          tgroup = make_uint2(0u, 0u);
          sp = 0;
          have = false; ambiguous = false; tie_t = -1.0f; win_leaf = -1; t2 = kFloatMax;
          h.geom = h.prim = -1;
          active = true;
          if ((RARE & kShapesBoxes) && sc.n_box_shapes) {  // see DBoxShape and traverse.cuh (iii); a Rect-only scene has none
            bool touched = false;
            for (int i = 0; i < sc.n_box_shapes && !touched; i++) {
              const int4 b = __ldg(reinterpret_cast<const int4*>(sc.box_shapes) + i);
              if (ANY) {
                touched = shape_hit<kShapesBoxes>(sc, b.x, b.y, ray);
              } else {
                Ray r2 = ray;
                touched = shape_intersect<kShapesBoxes>(sc, b.x, b.y, r2);
              }
              touched = touched && leaf_gate(sc.leaves, b.z, q, ray.tmin, ray.tmax);
            }
            if (touched) {
              if (ANY) io.store_any(idx, true);
              else wq.fb_list[atomicAdd(wq.fb_count, 1)] = idx;
#ifdef PB2_COUNT
              if (!ANY && sc.walk_stats) atomicAdd(sc.walk_stats + 6, 1ull);
#endif
              active = false;
            }
          }
        }
      }
    }
    if (__ballot_sync(0xffffffffu, active) == 0u) {
      if (drained) break;
      continue;
    }

    bool done = false;
    // one primitive record against the lane's ray 
    auto test_prim = [&](const PrimRec& r) {
        if (ANY) {
        bool ok;
        if ((RARE & kShapesMesh) && r.prim >= 0) ok = tri_hit(ray, r.v0, r.v1, r.v2);
        else if (r.prim >= -1 - PB2_SHAPE_OBB) ok = false;  // AABB/OBB: tested at refill
        else ok = shape_hit<RARE>(sc, -1 - r.prim, __float_as_int(r.v0.x), ray);
        if (ok && leaf_gate(sc.leaves, r.leaf, q, ray.tmin, ray.tmax)) { have = true; done = true; }
      } else {
        // Accept candidates up to slack_of(best): t == best exposes ties, and every candidate a
        // little behind the best is remembered in t2 (second-smallest t seen), which the
        // certificate at the end needs.
        Ray r2 = ray;
        r2.tmax = slack_of(ray.tmax);
        bool ok;
        if ((RARE & kShapesMesh) && r.prim >= 0) ok = tri_intersect(r2, r.v0, r.v1, r.v2);
        else if (r.prim >= -1 - PB2_SHAPE_OBB) ok = false;  // AABB/OBB: rays that touch one never get here
        else {
          ok = shape_intersect<RARE>(sc, -1 - r.prim, __float_as_int(r.v0.x), r2);
          // Plane::intersect lets a NaN t through and Cylinder::intersect a far root below tmin (geometry.cpp:40-45,
          // 502-504): what the reference does next depends on its visiting order, so it decides
          if ((RARE & kShapesRare) && ok && !(r2.tmax >= ray.tmin)) { ambiguous = true; done = true; ok = false; }
        }
        if (ok) {
          if (r2.tmax < ray.tmax) {
            if (have) t2 = pmin(t2, ray.tmax);
            ray.tmax = r2.tmax; h.geom = r.geom; h.prim = r.prim >= 0 ? r.prim : 0; win_leaf = r.leaf; have = true;
          } else if (r2.tmax == ray.tmax) {
            if (have) tie_t = r2.tmax;
            else { ambiguous = true; done = true; }
          } else if (have) {
            t2 = pmin(t2, r2.tmax);
          }
        }
      }
    };
    // ---- node step: every lane that has no primitives left to test ----
    const bool walking = active && tgroup.y == 0u;
    if (walking) {
      const uint32_t hits = ngroup.y, imask = hits & 0xffu;
      const int bit = 31 - __clz(hits);
      ngroup.y &= ~(1u << bit);
      if (ngroup.y > 0x00ffffffu) {
        if (sp < kWideStack) stack[sp++] = ngroup;
        else ambiguous = true;
      }
      const uint32_t slot = uint32_t(bit - 24) ^ (w.octinv4 & 7u);
      const uint32_t rel = __popc(imask & ~(0xffffffffu << slot));
      uint32_t child_base, prim_base, cmask;
#ifdef PB2_COUNT
      c_nodes++;
#endif
      const uint32_t hm = test_node8<((RARE & kShapesUnbounded) != 0)>(sc.nodes8, ngroup.x + rel, w, ray.tmin, ANY ? ray.tmax : slack_of(ray.tmax),
                                     ANY ? ray.tmax : tmax_in, child_base, prim_base, cmask, s_top);
      ngroup = make_uint2(child_base, (hm & 0xff000000u) | cmask);
      tgroup = make_uint2(prim_base, hm & 0x00ffffffu);
      if (ambiguous) {
        done = true;
      } else if (tgroup.y == 0u && ngroup.y <= 0x00ffffffu) {
        if (sp == 0) done = true;
        else ngroup = stack[--sp];
      }
    }
    // ---- primitive step ----
    const bool owns = active && !done && tgroup.y != 0u;
    const bool tight = __popc(__ballot_sync(0xffffffffu, owns)) >= PB2_TIGHT_LANES;
    if (owns) {
      do {
        const int i = __ffs(tgroup.y) - 1;
        tgroup.y &= tgroup.y - 1;
        const PrimRec r = load_prim(sc.wide_prims, tgroup.x + i);
#ifdef PB2_COUNT
        c_prims++;
#endif
        test_prim(r);
      } while (tight && tgroup.y != 0u && !done);
      if (!done && tgroup.y == 0u && ngroup.y <= 0x00ffffffu) {  // back to walking: restore the invariant
        if (sp == 0) done = true;
        else ngroup = stack[--sp];
      }
    }
    if (done) {
#if PB2_BATCH_FINISH
      pending = true;
#else
      finish();
#endif
      active = false;
      tgroup.y = 0u;
#ifdef PB2_COUNT
      c_rays++;
#endif
    }
  }
#ifdef PB2_COUNT
  if (sc.walk_stats) {
    atomicAdd(sc.walk_stats + (ANY ? 2 : 0), c_nodes);
    atomicAdd(sc.walk_stats + (ANY ? 3 : 1), c_prims);
    atomicAdd(sc.walk_stats + (ANY ? 5 : 4), c_rays);
  }
#endif
}

// Reference-order re-trace of the rays the wide walk could not certify.
template <bool ANY, int RARE = kShapesAll, class IO>  // RARE: a kShapes* mask
__device__ __forceinline__ void wide_fallback(const DScene& sc, WideQueue wq, const IO& io) {
  const int n = *wq.fb_count;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    const int idx = wq.fb_list[k];
    Ray ray = io.load(idx);
    if (ANY) {
      io.store_any(idx, any_ref_order<RARE>(sc, ray));
    } else {
      Hit h;
      const bool hit = closest_ref_order<RARE>(sc, ray, h);
      io.store_closest(sc, idx, ray, hit, h);
    }
  }
}

// One ray per thread in reference order (accel = 1, and closest hits of order-dependent scenes).
template <bool ANY, int RARE = kShapesAll, class IO>  // RARE: a kShapes* mask
__device__ __forceinline__ void ref_order_all(const DScene& sc, int n, const IO& io) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    Ray ray = io.load(i);
    if (ANY) {
      io.store_any(i, any_ref_order<RARE>(sc, ray));
    } else {
      Hit h;
      const bool hit = closest_ref_order<RARE>(sc, ray, h);
      io.store_closest(sc, i, ray, hit, h);
    }
  }
}

}  // namespace pb2
