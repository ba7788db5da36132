// rtp_kernels.h - launch interface of the wavefront ray-tracing kernels (rtp_kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "rtp_device.h"

namespace rtp {

// counters[] layout (ints), zeroed at the start of every frame
constexpr int kCtrPaths = 0;          // [depth] rays entering depth (depth 0: n_px, kept on the host)
constexpr int kCtrHits = 16;          // [depth] hits found at depth
constexpr int kCtrFetchClosest = 32;  // [depth] persistent-warp ray cursor of k_closest
constexpr int kCtrFetchAny = 48;      // [depth] persistent-warp ray cursor of k_anyhit
constexpr int kCtrAux = 62;           // ray-batch API cursor
constexpr int kCtrPark = 64;          // parked-ray list, 2 ints per trace launch of a frame ([depth] closest,
                                      // [8 + depth] any-hit, [16] ray-batch API): rays parked, k_coop cursor
constexpr int kCtrInts = kCtrPark + 2 * 17 + 2;
constexpr int kParkSlotBytes = 288;   // one parked ray (rtp_kernels.cu: kParkInts)

struct FrameParams {
  DeviceScene S;
  DeviceCamera cam;
  int width, height, tiles_x;
  int n_px;                // local paths at depth 0 = owned tiles * 1024 (tile-major) or user rays
  const int *tile_ids;     // owned tile -> global tile id
  const float *user_rays;  // non-null: depth-0 rays come from here (n x 6) instead of the camera
  int n_user_rays;
  int max_depth;
  float bg[3];
  float4 *fb;        // out, tile-major: r, g, b, bit-cast primary face id (may be null when img_* is set)
  float *img_rgb;           // out, optional: the finished image as floats, row-major W x H x 3
  unsigned char *img_rgb8;  // out, optional: the finished image, row-major W x H x 3, quantised like the PPM writer;
  int *img_hit;             //   and the primary face ids, W x H.  Either may live on ANOTHER GPU (peer memory
                            //   mapped through CUDA IPC): every rank stores its own tiles straight into GPU 0's image
  int *primface;     // [n_px] primary face id (-1 background, -2 pixel outside the image)
  float4 *ksdeep;    // [n_px] ks of the deepest hit, .w = hit-chain length bits
  float4 *direct;    // [max_depth][n_px] clamped direct colour per depth
  float4 *ray_o[2];  // ping-pong ray lists: (origin, bit-cast local pixel)
  float4 *ray_d[2];  //                      (direction, 1 = live / 0 = no ray)
  int *hit_slot;     // [n_px] per path of the current depth: index of its hit record (-1 miss, -2 no ray)
  float4 *pbuf;      // [n_px] per hit record: (P, bit-cast triangle slot)
  float4 *dbuf;      //        (ray direction, bit-cast local pixel)
  int *hitq;         //        path index at this depth (compacted levels; a camera level's record h is pixel h)
  unsigned char *occl;  // [n_px * n_lights] per hit and light: 1 = shadowed
  int *counters;
};

struct LaunchCfg {
  int blocks_closest = 0, blocks_any = 0, blocks_coop = 0, blocks_paths = 0, blocks_camera = 0, blocks_rays_any = 0, threads = 128;
  int *park_slots = nullptr;  // parked-ray list storage (park_cap * kParkSlotBytes), owned by the scene
  int park_cap = 0;
  int groups = 1;  // tile groups of the frame this launch belongs to (a group's n_px x groups = the rank's pixels)
  bool scene_exceeds_l2 = false;  // nodes + triangles larger than half the L2: traversal chains run on cold lines
};
void query_launch_cfg(int device, LaunchCfg &cfg);

// depth-0 rays of a caller-supplied batch into the ray list (camera frames generate theirs inside launch_closest)
void launch_raygen(const FrameParams &p, cudaStream_t st);
// calculateMinimumFace for every ray of `depth` (flyscene.cpp:373-396), at depth 0 of a camera frame with the ray
// generation of traceRayThread (flyscene.cpp:299-314); per hit P, interpolated normal and the compacted hit record
// (flyscene.cpp:464,572-600).  Returns the kernels launched (k_trace, plus k_coop when stragglers are parked)
int launch_closest(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st);
// shadow() for every (hit, light) pair (flyscene.cpp:510-526)
int launch_anyhit(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st);
// launch_anyhit(0) with the camera level's shading folded in (replaces launch_accum(0) and, at max_depth 1, launch_resolve);
// anyhit_shade_ok: the frame qualifies (camera rays, 1-32 lights, a launch that does not park shadow rays)
bool anyhit_shade_ok(const FrameParams &p, const LaunchCfg &cfg);
int launch_anyhit_shade(const FrameParams &p, const LaunchCfg &cfg, cudaStream_t st);
// The camera level of a frame in one kernel (k_camera: primary rays, their shadow rays, shading, and - max_depth 1 - the
// image; max_depth > 1: the level-0 records and level 1's ray list), plus k_coop_camera for parked pixels: replaces
// launch_closest(0) + launch_anyhit(0) + launch_accum(0) and, at max_depth 1, launch_resolve.  camera_fusable: the frame
// qualifies (camera rays, at most 32 lights, fewer than 2^24 pixels per launch).  Returns the kernels launched.
bool camera_fusable(const FrameParams &p);
// phase 0: both kernels, 1: k_camera only, 2: k_coop_camera only (launch timelines stamp between the two)
int launch_camera(const FrameParams &p, const LaunchCfg &cfg, cudaStream_t st, int phase = 0);
// every bounce level from first_depth on, path by path, in one persistent kernel (k_paths): replaces the per-level
// launch_closest / launch_anyhit / launch_accum sequence below the camera level
void launch_paths(const FrameParams &p, int first_depth, const LaunchCfg &cfg, cudaStream_t st);
// calcSingleColor / calculateColor + reflection-ray spawn (flyscene.cpp:542-566,603-614,355-363); not needed when
// max_depth is 1 (launch_resolve shades the single level itself)
void launch_accum(const FrameParams &p, int depth, cudaStream_t st);
// fold-back of the per-depth colours (flyscene.cpp:366-370) -> tile-major framebuffer and / or the finished image planes
void launch_resolve(const FrameParams &p, cudaStream_t st);
// tile-major buffers of all ranks -> row-major planes (any of the outputs may be null).
// only_rank >= 0: `all` holds just that rank's tiles; pixels of other ranks are left untouched.
void launch_untile(const float4 *all, size_t rank_stride_f4, const int *local_of_tile, int width, int height,
                   int world, int only_rank, float *rgb, unsigned char *rgb8, int *hit, cudaStream_t st);

// widens every child box of the hierarchy by `extra` (far cameras / ray origins: see k_inflate_nodes)
void launch_inflate_nodes(float4 *nodes, int n_nodes, float extra, cudaStream_t st);

// arrive-and-wait of one rank on a counter in (peer) device memory; see k_peer_barrier
void launch_peer_barrier(unsigned *counter, unsigned target, unsigned long long timeout_ns, int *timeouts, cudaStream_t st);

// ray-batch API (rtp_trace_rays): rays = n x (o, d) floats on the device
int launch_rays_closest(const DeviceScene &S, const float *rays, long long n, int *face, float *t, int *cursor,
                         const LaunchCfg &cfg, cudaStream_t st);
int launch_rays_anyhit(const DeviceScene &S, const float *rays, long long n, unsigned char *occ, int *cursor,
                        const LaunchCfg &cfg, cudaStream_t st);

// The same two launches with their scratch (ray cursor, parked-ray counters and slots) in a file-scope __device__ array
// of rtp_kernels.cu instead of the scene's cudaMalloc'ed blocks: nothing is allocated, every call on a device shares the
// array.  `device` must be current.  any: 0 = closest hit (face, t), 1 = any hit (occ).
constexpr int kNoAllocParkCap = 4096;  // parked-ray slots in the array (a ray that finds the list full carries on)
cudaError_t launch_rays_noalloc(int device, const DeviceScene &S, const float *rays, long long n, int any, int *face,
                                float *t, unsigned char *occ, LaunchCfg cfg, cudaStream_t st, int *launches);

}  // namespace rtp
