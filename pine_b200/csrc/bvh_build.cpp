// bvh_build.cpp -- see bvh_build.h. Restatement of BVHImpl::build_sah_binned
// (reference src/pine/impl/accel/bvh.cpp:43-147), same arithmetic, same tie-breaks.
//
// Parallelism: the recursion is a pure function of its primitive range, so big ranges hand their
// left half to another thread. Each half is built into its own fragment (nodes numbered from 0 in
// the reference's post-order) and the fragments are concatenated left, right, parent -- which is
// exactly the numbering the sequential recursion produces, so the tree is identical node for node
// whatever the thread count.
#include "bvh_build.h"

#if defined(__SSE2__)
#include <emmintrin.h>
#define PB2_BUILD_SSE 1
#endif

#include <atomic>
#include <future>
#include <thread>
#include <utility>

namespace pb2 {
namespace {

constexpr int kBuckets = 16;

struct Builder {
  Tree2& tree;

  int make_leaf(const BuildPrim* first, const BuildPrim* last) {
    Node2 node;
    node.leaf_offset = int(tree.leaf_prims.size());
    node.leaf_count = int(last - first);
    for (const BuildPrim* p = first; p != last; ++p) tree.leaf_prims.push_back(p->index);
    for (const BuildPrim* p = first; p != last; ++p) node.box[0].extend(p->box);
    node.box[1] = node.box[0];
    tree.nodes.push_back(node);
    return int(tree.nodes.size()) - 1;
  }

  static int bucket_of(const Box3& cbounds, const BuildPrim& prim, int axis) {
    return int(kBuckets * cbounds.relative_position(prim.box.centroid(axis), axis));
  }

  // Decides the split of [first,last) (bvh.cpp:50-131). Returns nullptr for "make a leaf",
  // else the partition point, with the two child boxes filled in.
  // The reference makes one pass over the primitives per axis and recomputes the bucket of every
  // primitive in the partition pass; here ONE pass fills the buckets of all three axes and remembers
  // each primitive's bucket (same float expressions per primitive and axis, and min/max and integer
  // counts do not depend on the order they are accumulated in, so every number is the reference's).
  std::vector<uint8_t> bucket_scratch;  // [3][count] of the range being split
  // `cb` (optional): the centroid bounds of [first,last) if the caller already has them -- the parent's
  // partition pass computes its children's while it computes their boxes, which saves the separate pass over
  // the range (min / max accumulate exactly, in any order). `child_cb` receives the children's.
  BuildPrim* choose_split(BuildPrim* first, BuildPrim* last, const Box3& bounds, Box3 (&child_box)[2],
                          const Box3* cb, Box3 (&child_cb)[2]) {
    const int count = int(last - first);
    if (count == 1) return nullptr;

#ifdef PB2_BUILD_SSE
    // Same float operations as the scalar statement below, three axes per instruction (IEEE add / sub / mul /
    // div are lane-wise identical; MINPS / MAXPS are `a < b ? a : b` / `a > b ? a : b`, i.e. pmin / pmax).
    // A box is read as two unaligned 4-float loads: lower.xyz + upper.x and upper.xyz + index; lane 3 is junk
    // and never leaves the registers.
    const __m128 half = _mm_set1_ps(0.5f);
    __m128 clo = _mm_set1_ps(kFloatMax), chi = _mm_set1_ps(-kFloatMax);
    Box3 cbounds;
    if (cb) {
      cbounds = *cb;
      clo = _mm_setr_ps(cb->lower.x, cb->lower.y, cb->lower.z, 0.0f);
      chi = _mm_setr_ps(cb->upper.x, cb->upper.y, cb->upper.z, 0.0f);
    } else {
      for (int i = 0; i < count; i++) {
        const __m128 c = _mm_mul_ps(_mm_add_ps(_mm_loadu_ps(&first[i].box.lower.x), _mm_loadu_ps(&first[i].box.upper.x)), half);
        clo = _mm_min_ps(clo, c);
        chi = _mm_max_ps(chi, c);
      }
      float t[4];
      _mm_storeu_ps(t, clo); cbounds.lower = V3(t[0], t[1], t[2]);
      _mm_storeu_ps(t, chi); cbounds.upper = V3(t[0], t[1], t[2]);
    }
    const float parent_area = bounds.surface_area();
    const bool live[3] = {!cbounds.degenerated(0), !cbounds.degenerated(1), !cbounds.degenerated(2)};
    int n_in[3][kBuckets] = {};
    unsigned occupied[3] = {0, 0, 0};  // bit b: bucket b of the axis is non-empty
    __m128 in_lo[3][kBuckets], in_hi[3][kBuckets];  // valid where n_in != 0 (min / max with the empty box's
                                                     // +-FLT_MAX returns the other operand, so "first touch
                                                     // assigns" equals "extend from empty")
    if (bucket_scratch.size() < size_t(count) * 3) bucket_scratch.resize(size_t(count) * 3);
    uint8_t* bk = bucket_scratch.data();
    // relative_position: dd > 0 ? o / dd : o -- dd > 0 on every live axis; dead axes get dd = 1 and are ignored
    const __m128 dd_raw = _mm_sub_ps(chi, clo);
    const __m128 dd = _mm_or_ps(_mm_and_ps(_mm_cmpgt_ps(dd_raw, _mm_setzero_ps()), dd_raw),
                                _mm_andnot_ps(_mm_cmpgt_ps(dd_raw, _mm_setzero_ps()), _mm_set1_ps(1.0f)));
    const __m128 nb = _mm_set1_ps(float(kBuckets));
    for (int i = 0; i < count; i++) {
      const __m128 lo = _mm_loadu_ps(&first[i].box.lower.x), hi = _mm_loadu_ps(&first[i].box.upper.x);
      const __m128 c = _mm_mul_ps(_mm_add_ps(lo, hi), half);
      const __m128i raw4 = _mm_cvttps_epi32(_mm_mul_ps(nb, _mm_div_ps(_mm_sub_ps(c, clo), dd)));
      alignas(16) int raw[4];
      _mm_store_si128(reinterpret_cast<__m128i*>(raw), raw4);
      for (int axis = 0; axis < 3; axis++) {
        if (!live[axis]) continue;
        bk[size_t(axis) * count + i] = uint8_t(raw[axis]);  // 0..16
        const int b = pmin(raw[axis], kBuckets - 1);
        if (n_in[axis][b]++ == 0) {
          occupied[axis] |= 1u << b;
          in_lo[axis][b] = lo;
          in_hi[axis][b] = hi;
        } else {
          in_lo[axis][b] = _mm_min_ps(in_lo[axis][b], lo);
          in_hi[axis][b] = _mm_max_ps(in_hi[axis][b], hi);
        }
      }
    }
    struct RawBox { float lower[3], upper[3]; };
    RawBox box_raw[3][kBuckets];  // uninitialised on purpose: only non-empty buckets are read below
    Box3(&box_in)[3][kBuckets] = reinterpret_cast<Box3(&)[3][kBuckets]>(box_raw);
    for (int a = 0; a < 3; a++)
      for (unsigned m = occupied[a]; m; m &= m - 1) {
        const int b = __builtin_ctz(m);
        float t[4];
        _mm_storeu_ps(t, in_lo[a][b]); box_in[a][b].lower = V3(t[0], t[1], t[2]);
        _mm_storeu_ps(t, in_hi[a][b]); box_in[a][b].upper = V3(t[0], t[1], t[2]);
      }
#else
    Box3 cbounds;
    if (cb) cbounds = *cb;
    else for (int i = 0; i < count; i++) cbounds.extend(first[i].box.centroid());
    const float parent_area = bounds.surface_area();

    const bool live[3] = {!cbounds.degenerated(0), !cbounds.degenerated(1), !cbounds.degenerated(2)};
    int n_in[3][kBuckets] = {};
    unsigned occupied[3] = {0, 0, 0};
    Box3 box_in[3][kBuckets];
    if (bucket_scratch.size() < size_t(count) * 3) bucket_scratch.resize(size_t(count) * 3);
    uint8_t* bk = bucket_scratch.data();
    for (int i = 0; i < count; i++) {
      const Box3& pb = first[i].box;
      for (int axis = 0; axis < 3; axis++) {
        if (!live[axis]) continue;
        const int raw = int(kBuckets * cbounds.relative_position(pb.centroid(axis), axis));
        bk[size_t(axis) * count + i] = uint8_t(raw);  // 0..16
        const int b = pmin(raw, kBuckets - 1);
        n_in[axis][b]++;
        occupied[axis] |= 1u << b;
        box_in[axis][b].extend(pb);
      }
    }

#endif

    float best_cost = kFloatMax;
    int best_axis = -1, best_bucket = -1;
    for (int axis = 0; axis < 3; axis++) {
      if (!live[axis]) continue;
      // cost[i] of the split after bucket i = 1 + (nL SA(L) + nR SA(R)) / SA(parent) (bvh.cpp:84-107). The
      // reference evaluates all 15; an empty side gives 0 * inf = NaN there, which never compares below the
      // running minimum, and between two consecutive non-empty buckets a < b both sides -- hence the cost --
      // are the same for every i in [a, b), of which the strict `<` keeps the first. So only i = a non-empty
      // bucket with another non-empty bucket after it can win: evaluate those, with the reference's operations.
      float right[kBuckets];  // right[i]: n * area of buckets >= i, valid where bucket i is non-empty
      {
        Box3 sweep;
        int n = 0;
        for (unsigned m = occupied[axis] & ~1u; m;) {  // descending over the non-empty buckets >= 1
          const int i = 31 - __builtin_clz(m);
          m &= ~(1u << i);
          sweep.extend(box_in[axis][i]);
          n += n_in[axis][i];
          right[i] = n * sweep.surface_area();
        }
      }
      float axis_cost = kFloatMax;
      int axis_bucket = -1;
      {
        Box3 sweep;
        int n = 0;
        for (unsigned m = occupied[axis]; m & (m - 1); m &= m - 1) {  // ascending, all but the last non-empty bucket
          const int i = __builtin_ctz(m);
          const int nxt = __builtin_ctz(m & (m - 1));
          sweep.extend(box_in[axis][i]);
          n += n_in[axis][i];
          float c = 0.0f;
          c += n * sweep.surface_area();
          c += right[nxt];
          c = 1.0f + c / parent_area;
          if (c < axis_cost) {
            axis_cost = c;
            axis_bucket = i;
          }
        }
      }
      if (axis_cost < best_cost) {
        best_cost = axis_cost;
        best_axis = axis;
        best_bucket = axis_bucket;
      }
    }

    if (best_cost > float(count)) return nullptr;

    // psl::partition (src/psl/algorithm.h:394-402): forward scan, swap matches to the front. The
    // remembered buckets travel with their primitives (only positions < it change before it is read).
    uint8_t* bb = bk + size_t(best_axis) * count;
    int mid_i = 0;
    for (int i = 0; i < count; i++) {
      int b = bb[i];
      if (b == kBuckets) b = kBuckets - 1;
      if (b <= best_bucket) {
        std::swap(first[mid_i], first[i]);
        std::swap(bb[mid_i], bb[i]);
        ++mid_i;
      }
    }
    BuildPrim* mid = first + mid_i;

#ifdef PB2_BUILD_SSE
    for (int c = 0; c < 2; c++) {
      __m128 lo = _mm_set1_ps(kFloatMax), hi = _mm_set1_ps(-kFloatMax);
      __m128 cl = lo, ch = hi;
      for (BuildPrim* p = c ? mid : first; p != (c ? last : mid); ++p) {
        const __m128 pl = _mm_loadu_ps(&p->box.lower.x), ph = _mm_loadu_ps(&p->box.upper.x);
        lo = _mm_min_ps(lo, pl);
        hi = _mm_max_ps(hi, ph);
        const __m128 ctr = _mm_mul_ps(_mm_add_ps(pl, ph), half);
        cl = _mm_min_ps(cl, ctr);
        ch = _mm_max_ps(ch, ctr);
      }
      float t[4];
      _mm_storeu_ps(t, lo); child_box[c].lower = V3(t[0], t[1], t[2]);
      _mm_storeu_ps(t, hi); child_box[c].upper = V3(t[0], t[1], t[2]);
      _mm_storeu_ps(t, cl); child_cb[c].lower = V3(t[0], t[1], t[2]);
      _mm_storeu_ps(t, ch); child_cb[c].upper = V3(t[0], t[1], t[2]);
    }
#else
    child_cb[0] = child_cb[1] = Box3();
    for (BuildPrim* p = first; p != mid; ++p) { child_box[0].extend(p->box); child_cb[0].extend(p->box.centroid()); }
    for (BuildPrim* p = mid; p != last; ++p) { child_box[1].extend(p->box); child_cb[1].extend(p->box.centroid()); }
#endif
    return mid;
  }

  int split(BuildPrim* first, BuildPrim* last, const Box3& bounds, const Box3* cb = nullptr) {
    Node2 node;
    Box3 child_cb[2];
    BuildPrim* mid = choose_split(first, last, bounds, node.box, cb, child_cb);
    if (!mid) return make_leaf(first, last);
    node.child[0] = split(first, mid, node.box[0], &child_cb[0]);
    node.child[1] = split(mid, last, node.box[1], &child_cb[1]);
    tree.nodes.push_back(node);
    return int(tree.nodes.size()) - 1;
  }
};

constexpr int kParallelMin = 1 << 15;  // ranges below this are built by the calling thread
std::atomic<int> g_spare_threads{0};   // threads the builder may still start

void append_fragment(Tree2& dst, Tree2&& src) {
  const int node_base = int(dst.nodes.size()), prim_base = int(dst.leaf_prims.size());
  dst.leaf_prims.insert(dst.leaf_prims.end(), src.leaf_prims.begin(), src.leaf_prims.end());
  dst.nodes.reserve(dst.nodes.size() + src.nodes.size() + 1);
  for (Node2 nd : src.nodes) {
    if (nd.leaf_count) nd.leaf_offset += prim_base;
    else { nd.child[0] += node_base; nd.child[1] += node_base; }
    dst.nodes.push_back(nd);
  }
}

// Builds [first,last) into `out` (which must be empty); the subtree root is out.nodes.back().
void split_parallel(BuildPrim* first, BuildPrim* last, const Box3& bounds, Tree2& out) {
  Builder b{out};
  if (last - first < kParallelMin) { b.split(first, last, bounds); return; }
  Node2 node;
  Box3 child_cb[2];
  BuildPrim* mid = b.choose_split(first, last, bounds, node.box, nullptr, child_cb);
  if (!mid) { b.make_leaf(first, last); return; }
  Tree2 right;
  if (g_spare_threads.fetch_sub(1) > 0) {
    auto fut = std::async(std::launch::async, [&] { split_parallel(first, mid, node.box[0], out); });
    split_parallel(mid, last, node.box[1], right);
    fut.get();
    g_spare_threads.fetch_add(1);
  } else {
    g_spare_threads.fetch_add(1);
    split_parallel(first, mid, node.box[0], out);
    split_parallel(mid, last, node.box[1], right);
  }
  node.child[0] = int(out.nodes.size()) - 1;
  append_fragment(out, std::move(right));
  node.child[1] = int(out.nodes.size()) - 1;
  out.nodes.push_back(node);
}

}  // namespace

void set_build_threads(int n) { g_spare_threads.store(n > 1 ? n - 1 : 0); }

void build_sah_binary(std::vector<BuildPrim> prims, Tree2& out) {
  out = Tree2();
  if (prims.empty()) return;
  Box3 bounds;
  for (auto& p : prims) bounds.extend(p.box);
  if (prims.size() < size_t(kParallelMin)) {
    out.nodes.reserve(prims.size());
    out.leaf_prims.reserve(prims.size());
  }
  split_parallel(prims.data(), prims.data() + prims.size(), bounds, out);
  out.root = int(out.nodes.size()) - 1;
}

}  // namespace pb2
