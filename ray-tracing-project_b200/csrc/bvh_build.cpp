// bvh_build.cpp - binned-SAH BVH2 builder (host, multithreaded).  See bvh_build.h.
#include "bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

namespace rtp {
namespace {

constexpr int kBins = 16;
constexpr float kCostNode = 1.0f;
float kCostTri = 2.0f;  // the exact reference triangle test is ~2-3x a slab pair (RTP_SAH_TRI_COST: tuning runs)
constexpr int kMedianDepth = 32;  // beyond this depth fall back to median splits (bounds the stack)

inline void box_reset(Aabb &b) {
  for (int k = 0; k < 3; ++k) {
    b.lo[k] = std::numeric_limits<float>::infinity();
    b.hi[k] = -std::numeric_limits<float>::infinity();
  }
}
inline void box_grow(Aabb &b, const Aabb &o) {
  for (int k = 0; k < 3; ++k) {
    b.lo[k] = std::min(b.lo[k], o.lo[k]);
    b.hi[k] = std::max(b.hi[k], o.hi[k]);
  }
}
inline float box_area(const Aabb &b) {
  float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f)) return 0.f;
  return 2.f * (dx * dy + dy * dz + dz * dx);
}

struct Builder {
  const Aabb *pb;
  std::vector<float> cen;  // 3 per prim
  int32_t *idx;
  BvhNode *nodes;  // sparse: a subtree over n prims owns n-1 consecutive slots
  int max_leaf;
  int par_min;  // 0: single-threaded
  std::atomic<int> depth_seen{0};

  static int32_t leaf_code(int first, int count) { return ~((first << 3) | (count - 1)); }

  void set_child_box(BvhNode &n, int which, const Aabb &b) {
    if (which == 0) {
      n.n0[0] = b.lo[0]; n.n0[1] = b.hi[0]; n.n0[2] = b.lo[1]; n.n0[3] = b.hi[1];
      n.n2[0] = b.lo[2]; n.n2[1] = b.hi[2];
    } else {
      n.n1[0] = b.lo[0]; n.n1[1] = b.hi[0]; n.n1[2] = b.lo[1]; n.n1[3] = b.hi[1];
      n.n2[2] = b.lo[2]; n.n2[3] = b.hi[2];
    }
  }

  Aabb range_bounds(int l, int r) const {
    Aabb b;
    box_reset(b);
    for (int i = l; i < r; ++i) box_grow(b, pb[idx[i]]);
    return b;
  }

  // Finds a split of [l, r); returns mid and the two child boxes.  force: never return "leaf".
  bool find_split(int l, int r, int depth, bool force, int &mid, Aabb &lb, Aabb &rb) {
    const int n = r - l;
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    Aabb whole;
    box_reset(whole);
    for (int i = l; i < r; ++i) {
      const float *c = &cen[3 * (size_t)idx[i]];
      for (int k = 0; k < 3; ++k) { clo[k] = std::min(clo[k], c[k]); chi[k] = std::max(chi[k], c[k]); }
      box_grow(whole, pb[idx[i]]);
    }
    float best_cost = INFINITY;
    int best_axis = -1, best_bin = -1;
    if (depth < kMedianDepth) {
      const float inv_area = 1.0f / std::max(box_area(whole), 1e-30f);
      // one sweep bins every primitive on the three axes at once
      int cnt[3][kBins] = {{0}};
      Aabb bb[3][kBins];
      float scale[3];
      bool live[3];
      for (int axis = 0; axis < 3; ++axis) {
        float ext = chi[axis] - clo[axis];
        live[axis] = ext > 0.f;
        scale[axis] = live[axis] ? kBins * (1.0f - 1e-6f) / ext : 0.f;
        for (auto &b : bb[axis]) box_reset(b);
      }
      for (int i = l; i < r; ++i) {
        const int p = idx[i];
        const Aabb &pbox = pb[p];
        const float *c = &cen[3 * (size_t)p];
        for (int axis = 0; axis < 3; ++axis) {
          if (!live[axis]) continue;
          int b = (int)((c[axis] - clo[axis]) * scale[axis]);
          b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
          cnt[axis][b]++;
          box_grow(bb[axis][b], pbox);
        }
      }
      for (int axis = 0; axis < 3; ++axis) {
        if (!live[axis]) continue;
        float rarea[kBins];
        int rcnt[kBins];
        Aabb acc;
        box_reset(acc);
        int c = 0;
        for (int b = kBins - 1; b > 0; --b) {
          box_grow(acc, bb[axis][b]);
          c += cnt[axis][b];
          rarea[b] = box_area(acc);
          rcnt[b] = c;
        }
        box_reset(acc);
        c = 0;
        for (int b = 0; b < kBins - 1; ++b) {
          box_grow(acc, bb[axis][b]);
          c += cnt[axis][b];
          if (c == 0 || rcnt[b + 1] == 0) continue;
          float cost = kCostNode + kCostTri * (c * box_area(acc) + rcnt[b + 1] * rarea[b + 1]) * inv_area;
          if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
        }
      }
    }
    if (best_axis >= 0) {
      if (!force && n <= max_leaf && kCostTri * n <= best_cost) return false;
      float ext = chi[best_axis] - clo[best_axis];
      float scale = kBins * (1.0f - 1e-6f) / ext;
      float lo = clo[best_axis];
      int32_t *m = std::partition(idx + l, idx + r, [&](int32_t p) {
        int b = (int)((cen[3 * (size_t)p + best_axis] - lo) * scale);
        b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
        return b <= best_bin;
      });
      mid = (int)(m - idx);
    }
    if (best_axis < 0 || mid == l || mid == r) {
      if (!force && n <= max_leaf) return false;
      // object-median on the widest centroid axis (or plain index split for coincident centroids)
      int axis = 0;
      for (int k = 1; k < 3; ++k)
        if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
      mid = l + n / 2;
      std::nth_element(idx + l, idx + mid, idx + r, [&](int32_t a, int32_t b) {
        return cen[3 * (size_t)a + axis] < cen[3 * (size_t)b + axis];
      });
    }
    lb = range_bounds(l, mid);
    rb = range_bounds(mid, r);
    return true;
  }

  // Builds the subtree over [l, r) with its node at `slot`.  Only called when a split exists.
  void build_node(int l, int r, int mid, const Aabb &lb, const Aabb &rb, int slot, int depth) {
    int seen = depth_seen.load(std::memory_order_relaxed);
    while (depth > seen && !depth_seen.compare_exchange_weak(seen, depth)) {}
    BvhNode &node = nodes[slot];
    std::memset(&node, 0, sizeof node);
    set_child_box(node, 0, lb);
    set_child_box(node, 1, rb);
    auto child = [this, depth](int cl, int cr, int cslot, int32_t *out) {
      int cm;
      Aabb a, b;
      if (cr - cl > 0 && find_split(cl, cr, depth + 1, cr - cl > max_leaf, cm, a, b)) {
        *out = cslot;
        build_node(cl, cr, cm, a, b, cslot, depth + 1);
      } else {
        *out = leaf_code(cl, cr - cl);
      }
    };
    const int nl = mid - l;
    int32_t *c0 = &node.child0, *c1 = &node.child1;
    // one task per subtree above par_min primitives: SAH splits are uneven, so there are several
    // times more tasks than threads and the parent works on the other half meanwhile
    if (par_min > 0 && nl > par_min && r - mid > par_min) {
      std::thread th(child, l, mid, slot + 1, c0);
      child(mid, r, slot + nl, c1);
      th.join();
    } else {
      child(l, mid, slot + 1, c0);
      child(mid, r, slot + nl, c1);
    }
  }
};

}  // namespace

void build_bvh(const std::vector<Aabb> &prim_bounds, int max_leaf, int n_threads, Bvh &out) {
  const int n = (int)prim_bounds.size();
  out.nodes.clear();
  out.order.resize(n);
  for (int i = 0; i < n; ++i) out.order[i] = i;
  if (n == 0) return;
  if (const char *e = std::getenv("RTP_SAH_TRI_COST")) kCostTri = (float)std::atof(e);
  if (max_leaf > 8) max_leaf = 8;
  if (max_leaf < 1) max_leaf = 1;
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  if (n_threads <= 0) n_threads = 1;

  const bool timing = std::getenv("RTP_BUILD_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[build_bvh]     %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };
  Builder b;
  b.pb = prim_bounds.data();
  b.cen.resize(3 * (size_t)n);
  for (int i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) b.cen[3 * (size_t)i + k] = 0.5f * (prim_bounds[i].lo[k] + prim_bounds[i].hi[k]);
  b.idx = out.order.data();
  b.max_leaf = max_leaf;
  b.par_min = n_threads > 1 ? std::max(8192, n / (8 * n_threads)) : 0;

  lap("centroids");
  std::vector<BvhNode> sparse((size_t)std::max(n - 1, 1));
  b.nodes = sparse.data();
  if (n == 1) {
    // root must be an inner node: both children are the same single-triangle leaf (the second
    // visit finds an equal t with an equal rank and is ignored)
    BvhNode &r = sparse[0];
    std::memset(&r, 0, sizeof r);
    b.set_child_box(r, 0, prim_bounds[0]);
    b.set_child_box(r, 1, prim_bounds[0]);
    r.child0 = r.child1 = Builder::leaf_code(0, 1);
  } else {
    int mid;
    Aabb lb, rb;
    b.find_split(0, n, 0, /*force=*/true, mid, lb, rb);
    b.build_node(0, n, mid, lb, rb, 0, 0);
  }

  lap("split recursion");
  // compact the sparse array in DFS pre-order
  out.nodes.reserve((size_t)n / 2 + 1);
  struct Item { int sparse_slot; int parent; int which; int depth; };
  std::vector<Item> stack;
  stack.push_back({0, -1, 0, 0});
  out.max_depth = 0;
  out.n_leaves = 0;
  while (!stack.empty()) {
    Item it = stack.back();
    stack.pop_back();
    int me = (int)out.nodes.size();
    out.nodes.push_back(sparse[it.sparse_slot]);
    if (it.parent >= 0) (it.which ? out.nodes[it.parent].child1 : out.nodes[it.parent].child0) = me;
    out.max_depth = std::max(out.max_depth, it.depth + 1);
    const BvhNode nd = sparse[it.sparse_slot];
    // push child1 first so child0 is laid out right after its parent
    if (nd.child1 >= 0) stack.push_back({nd.child1, me, 1, it.depth + 1}); else out.n_leaves++;
    if (nd.child0 >= 0) stack.push_back({nd.child0, me, 0, it.depth + 1}); else out.n_leaves++;
  }
  lap("pre-order compaction");
}

}  // namespace rtp
