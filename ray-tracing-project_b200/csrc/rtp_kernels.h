// rtp_kernels.h - launch interface of the wavefront ray-tracing kernels (rtp_kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "rtp_device.h"

namespace rtp {

// counters[] layout (ints)
constexpr int kCtrPaths = 0;    // [depth]   number of live paths entering depth
constexpr int kCtrFetch = 16;   // [2*depth + {0: closest, 1: shade}] persistent-warp batch cursors
constexpr int kCtrAux = 48;     // ray-batch API cursor
constexpr int kCtrInts = 64;
// stats[] layout (unsigned long long)
constexpr int kStatPrimary = 0, kStatShadow = 1, kStatReflection = 2, kStatHits = 3, kStatCount = 4;

struct FrameParams {
  DeviceScene S;
  DeviceCamera cam;
  int width, height, tiles_x;
  int n_px;             // local pixels = owned tiles * 1024 (tile-major order)
  const int *tile_ids;  // owned tile -> global tile id
  const float *user_rays;  // non-null: depth-0 rays come from here (n x 6) instead of the camera
  int n_user_rays;
  int max_depth;
  float bg[3];
  float4 *fb;           // out, tile-major: r, g, b, bit-cast primary face id
  int *primface;        // [n_px]
  float4 *ksdeep;       // [n_px] ks of the deepest hit, .w = hit-chain length bits
  float4 *direct;       // [max_depth][n_px] clamped direct colour per depth
  float4 *ray_o[2];     // ping-pong path buffers: (origin, bit-cast local pixel)
  float4 *ray_d[2];     //                         (direction, -)
  float *hit_t;         // [n_px] closest-hit t of the current depth
  int *hit_slot;        // [n_px] closest-hit triangle slot (-1 miss, -2 pixel outside the image)
  int *counters;
  unsigned long long *stats;
};

struct LaunchCfg {
  int blocks_closest = 0, blocks_shade = 0, threads = 128;
};
void query_launch_cfg(int device, LaunchCfg &cfg);

void launch_closest(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st);
void launch_shade(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st);
void launch_resolve(const FrameParams &p, cudaStream_t st);
// tile-major buffers of all ranks -> row-major planes (any of the outputs may be null).
// only_rank >= 0: `all` holds just that rank's tiles; pixels of other ranks are left untouched.
void launch_untile(const float4 *all, size_t rank_stride_f4, const int *local_of_tile, int width, int height,
                   int world, int only_rank, float *rgb, unsigned char *rgb8, int *hit, cudaStream_t st);

// ray-batch API (rtp_trace_rays): rays = n x (o, d) floats on the device
void launch_rays_closest(const DeviceScene &S, const float *rays, long long n, int *face, float *t, int *cursor,
                         const LaunchCfg &cfg, cudaStream_t st);
void launch_rays_anyhit(const DeviceScene &S, const float *rays, long long n, unsigned char *occ, int *cursor,
                        const LaunchCfg &cfg, cudaStream_t st);

}  // namespace rtp
