// bvh_build.h - host-side binned-SAH builder for the linear BVH2 the traversal kernels walk.
//
// This replaces the reference's acceleration structure for *culling only*: the reference keeps a
// flat list of overlapping AABBs and tests every one for every ray (src/flyscene.cpp:381-393,
// src/BoundingBox.cpp:109-149).  Which faces a ray may see is still decided per candidate by the
// reference's own box test (rtp_device.h: ref_box_pass), so the hierarchy built here only has to
// be conservative.
#pragma once
#include <cstdint>
#include <vector>

namespace rtp {

struct Aabb {
  float lo[3], hi[3];
};

struct BvhNode {  // 64 bytes, the device layout (see rtp_device.h)
  float n0[4], n1[4], n2[4];
  int32_t child0, child1, pad0, pad1;
};

struct Bvh {
  std::vector<BvhNode> nodes;   // nodes[0] is the root (always an inner node)
  std::vector<int32_t> order;   // leaf order: slot -> input primitive index
  int max_depth = 0;
  int n_leaves = 0;
};

// prim_bounds: one (already padded) box per primitive.  max_leaf <= 8.
// n_threads <= 0: std::thread::hardware_concurrency().
void build_bvh(const std::vector<Aabb> &prim_bounds, int max_leaf, int n_threads, Bvh &out);

}  // namespace rtp
