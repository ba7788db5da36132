// scene_host.h - host-side flattening of the reference's mesh into the device layout
// (rtp_device.h) plus the BVH over it.  Pure C++ (no CUDA) so tests/hostsim can reuse it.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "../../include/rtp.h"
#include "bvh_build.h"
#include "rtp_device.h"

namespace rtp {

// std::vector whose resize() leaves trivially-constructible elements uninitialised: the per-triangle
// arrays are hundreds of MB and every element is written by the (threaded) loops that follow, so a
// single-threaded zero fill would only serialise the first touch of those pages.
template <class T>
struct DefaultInit : std::allocator<T> {
  template <class U> struct rebind { using other = DefaultInit<U>; };
  template <class U> void construct(U *p) noexcept(noexcept(::new ((void *)p) U)) { ::new ((void *)p) U; }
  template <class U, class... A> void construct(U *p, A &&...a) { ::new ((void *)p) U(std::forward<A>(a)...); }
};
template <class T> using BigVec = std::vector<T, DefaultInit<T>>;

struct HostScene {
  BigVec<float4> nodes;           // 4 per node
  BigVec<float4> tris;            // 4 per triangle, leaf order
  BigVec<float4> vnorm;           // 3 per triangle, leaf order
  std::vector<float4> boxes;      // 2 per reference box
  BigVec<int32_t> face_of;        // slot -> face id
  std::vector<float4> materials;  // 3 per entry, entry 0 = Flyscene defaults (material_id -1)
  float minv[12];
  // reference BoundingBox list (kept for rtp_scene_get_boxes)
  std::vector<float> box_bounds;  // B x 6
  std::vector<int32_t> box_off, box_faces;
  int n_tris = 0, n_nodes = 0, n_boxes = 0, bvh_depth = 0;
  int n_flagged = 0;     // triangles that need the zero-normal check
  int n_degenerate = 0;  // triangles whose face normal is not finite/unit (edge-case list)
  float pad = 0.f;       // AABB padding used
  float maxabs = 1.f;    // largest finite |coordinate| of the world-space vertices (the pad is relative to it)
  float scene_lo[3] = {0, 0, 0}, scene_hi[3] = {0, 0, 0};  // bounds of the finite (unpadded) triangles
};

// Returns "" on success, else an error message.  with_bvh = false: no hierarchy is built; tris / vnorm /
// face_of stay in reference rank order and nodes stays empty (the device builder takes over).
std::string flatten_scene(const rtp_mesh &m, HostScene &out, int n_threads = 0, bool with_bvh = true);

// Camera::getCenter, getViewMatrix().inverse(), perspective scale (camera.hpp:115-118,155-173,263-266)
void derive_camera(const rtp_camera &c, DeviceCamera &out);

// Flyscene::lights then dirLights, in the order calculateColor visits them (flyscene.cpp:606-612)
void pack_lights(const rtp_lights &l, std::vector<float4> &out);

// Screen tiles owned by `rank` in increasing tile id; local_of_tile[tile id] = index within the owner
void tile_tables(int width, int height, int rank, int world, std::vector<int32_t> &tiles_of_rank,
                 std::vector<int32_t> &local_of_tile);

}  // namespace rtp
