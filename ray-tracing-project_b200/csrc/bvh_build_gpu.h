// bvh_build_gpu.h - LBVH construction on the device (bvh_build_gpu.cu).
#pragma once
#include <cuda_runtime.h>

namespace rtp {

// d_tris_in / d_vnorm_in / d_face_in: the flattened triangles in reference rank order (4 / 3 float4 and
// one face id per triangle).  Outputs, all caller-allocated device memory: the same arrays in leaf order
// (d_tris, d_vnorm, d_face_of: n entries) and d_nodes (room for n - 1 nodes of 4 float4).  *n_nodes =
// nodes written (node 0 = root), *height = height of the radix tree (bounds the traversal stack).
// Synchronises `st` before returning.
cudaError_t build_lbvh_device(const float4 *d_tris_in, const float4 *d_vnorm_in, const int *d_face_in, int n, float pad,
                              const float scene_lo[3], const float scene_hi[3], int max_leaf, float4 *d_tris, float4 *d_vnorm,
                              int *d_face_of, float4 *d_nodes, int *n_nodes, int *height, cudaStream_t st);

}  // namespace rtp
