/*
 * rtp.h - C ABI of the B200-native ray-tracing hot path (libflyscene_b200.so).
 *
 * This is the drop-in boundary for plindhorst/Ray-Tracing-Project's one data-parallel path:
 * Flyscene::raytraceScene / traceRayThread / traceRay and everything under them
 * (reference src/flyscene.cpp:250-614, src/BoundingBox.cpp:7-161).  The reference has no
 * FFI of its own - its boundary is the Flyscene / BoundingBox C++ class surface - so each
 * entry point below names the reference member it replaces.  The C++ mirror classes in
 * ray-tracing-project_b200/host/ (Flyscene, BoundingBox, Flycamera) are thin callers of this
 * ABI; INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * Conventions: plain pointers and sizes, caller owns all host buffers, the library owns all
 * device memory behind the opaque rtp_scene handle.  Every function returns 0 on success and
 * a negative rtp_status on failure; rtp_last_error() gives the message (thread-local).  No
 * exceptions cross the ABI.  There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with RTP_ERR_CUDA.
 */
#ifndef RTP_H
#define RTP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTP_ABI_VERSION 7

typedef enum rtp_status {
  RTP_OK = 0,
  RTP_ERR_INVALID = -1, /* bad argument                                   */
  RTP_ERR_CUDA = -2,    /* CUDA runtime error or no device                */
  RTP_ERR_NOMEM = -3,
  RTP_ERR_STATE = -4    /* call order (e.g. render before set_camera)     */
} rtp_status;

typedef struct rtp_scene rtp_scene; /* opaque */

/* The mesh exactly as the reference's Tucano::Mesh holds it after loadObjFile +
 * normalizeModelMatrix (flyscene.cpp:18-22; tucano/mesh.hpp:238-248,387,395,426).
 * Object space; the library flattens it once into world-space SoA float4 + a linear BVH. */
typedef struct rtp_mesh {
  int32_t n_verts, n_faces, n_mats, n_boxes;
  const float *verts;       /* V x 4  Mesh::getVertex(i), w = 1                          */
  const float *normals;     /* V x 3  Mesh::getNormal(i) (not necessarily unit)          */
  const int32_t *faces;     /* F x 3  Face::vertex_ids                                   */
  const int32_t *face_mat;  /* F      Face::material_id (-1 = none)                      */
  const float *face_normal; /* F x 3  Face::normal, or NULL: built as mesh.hpp:475-477   */
  const float *materials;   /* M x 12 per Mtl: ka[3] kd[3] ks[3] Ns Ni d (mtl.hpp:17-37) */
  float shape[12];          /* Model::getShapeModelMatrix(), 3x4 row-major (model.hpp:102) */
  /* The BoundingBox list in creation order (BoundingBox.h:13,16,17,26).  It fixes which faces
   * a ray may see (flyscene.cpp:381-393) and the first-visited-wins tie order.  n_boxes = 0
   * with NULL pointers: the library builds it with the reference's splitter
   * (flyscene.cpp:399-418) using min_faces / max_boxes below. */
  const float *boxes;         /* B x 6: low[3] high[3], object space */
  const int32_t *box_off;     /* B + 1 */
  const int32_t *box_faces;   /* F face ids in box order */
  int32_t min_faces;          /* Flyscene::MIN_FACES (300) when 0 */
  int32_t max_boxes;          /* Flyscene::MAX_BOXES (INT_MAX) when 0 */
  int32_t build;              /* RTP_BUILD_*: where the traversal hierarchy over the faces is built */
  int32_t reserved;
} rtp_mesh;
#define RTP_BUILD_HOST_SAH 0 /* binned SAH on the host threads: best trees (default)                          */
#define RTP_BUILD_GPU_LBVH 1 /* Morton sort + radix tree on the device: ~50x faster to build, slower to trace; */
                             /* the image is the same either way (the hierarchy only culls)                    */

/* Tucano::Camera state read by screenToWorld / getCenter (camera.hpp:115-118,155-173). */
typedef struct rtp_camera {
  float view[12];    /* Camera::view_matrix, 3x4 row-major                */
  float viewport[4]; /* Camera::viewport (x, y, w, h)                     */
  float fovy;        /* degrees, as given to setPerspectiveMatrix         */
  float aspect;      /* aspect_ratio as given to setPerspectiveMatrix     */
} rtp_camera;

/* Flyscene::lights / Flyscene::dirLights (flyscene.hpp:132,135): 6 floats each,
 * position-or-vector[3] then rgb[3].  Directional vectors are used verbatim, unnormalised
 * (flyscene.cpp:611). */
typedef struct rtp_lights {
  int32_t n_point, n_dir;
  const float *point;
  const float *dir;
} rtp_lights;

typedef struct rtp_render_opts {
  int32_t width, height; /* image size; the camera's viewport drives ray generation       */
  int32_t max_depth;     /* Flyscene::max_depth (2): reflection depth D <=> max_depth D+1  */
  float background[3];   /* Flyscene::BACKGROUND_COLOR                                    */
  int32_t rank, world;   /* tile sharding: this process renders tiles with owner == rank   */
  int32_t flags;         /* RTP_RENDER_* */
} rtp_render_opts;
#define RTP_RENDER_DEFAULT 0
#define RTP_RENDER_PROFILE 1 /* bracket each kernel with CUDA events -> ms_closest / ms_anyhit / ms_shade */

typedef struct rtp_stats {
  int64_t rays_primary, rays_shadow, rays_reflection, hits;
  float ms_trace;       /* device time of the traced region (CUDA events)                  */
  float ms_total;       /* device time incl. output conversion, excl. H2D/D2H              */
  int32_t kernel_launches;
  int32_t n_bvh_nodes, n_triangles;
  /* with RTP_RENDER_PROFILE: summed device time per kernel kind */
  float ms_closest; /* closest-hit traversal (calculateMinimumFace)                            */
  float ms_anyhit;  /* shadow-ray traversal (shadow)                                           */
  float ms_shade;   /* streaming kernels: Phong + reflection spawn of the camera level, resolve + image   */
  float ms_paths;   /* bounce levels >= 1, path by path in one persistent kernel (traceRay's recursion) */
  float ms_camera;  /* the camera level in one kernel (primary rays + their shadow rays + shading [+ image]); when it   */
                    /* runs, ms_closest / ms_anyhit / ms_shade hold the levels below it and the fold only               */
} rtp_stats;

/* ---- lifetime ---------------------------------------------------------------------------- */
/* Replaces Flyscene::initialize's scene side + generateBoundingBoxes (flyscene.cpp:18-22,62,
 * 399-428).  device = CUDA ordinal (-1: current). */
int rtp_scene_create(const rtp_mesh *mesh, int device, rtp_scene **out);
int rtp_scene_destroy(rtp_scene *s); /* ~Flyscene / BoundingBox::deconstruct */

/* The same scene on n_devices GPUs of one box, driven from ONE process: replaces the fan-out of the frame over
 * Flyscene::THREADS row-band threads and their join inside raytraceScene (flyscene.cpp:266-289).  The mesh is
 * flattened and its hierarchy built once, then replicated device to device; rtp_set_lights / rtp_set_camera reach
 * every replica; rtp_render cuts the frame into interleaved 32x32 screen tiles (owner = (tx + ty) % n_devices), each
 * device renders its tiles on its own stream, fed by its own host thread, and its resolve kernel stores the finished
 * pixels straight into devices[0]'s image over NVLink (peer access); fork and join are CUDA events, then one copy to
 * the host.  The image is bitwise the one a single device renders.  devices[] may name a device more than once (the
 * ranks then share that GPU: how the sharded path is exercised on a one-GPU box).  rtp_render_async, the *_device
 * entry points and rtp_trace_rays use devices[0] only. */
#define RTP_MAX_DEVICES 64
int rtp_scene_create_multi(const rtp_mesh *mesh, const int *devices, int n_devices, rtp_scene **out);
int rtp_scene_device_count(const rtp_scene *s); /* 1 for rtp_scene_create */
int rtp_device_count(void);                     /* CUDA devices visible to the process; negative rtp_status on error */

/* Flyscene::lights, dirLights (public members, flyscene.hpp:132,135). */
int rtp_set_lights(rtp_scene *s, const rtp_lights *l);
/* getCamera()->setViewport / setPerspectiveMatrix / view matrix (flyscene.cpp:14-15). */
int rtp_set_camera(rtp_scene *s, const rtp_camera *c);

/* Tuning knob with no counterpart in the reference (its 20 row-band threads, flyscene.cpp:266-285, are the
 * nearest relative): a frame's owned tiles are cut into `groups` bands that run as independent sub-frames
 * on their own CUDA streams, so that one band's kernels fill the idle tail of another's.  0 = automatic
 * (default), 1 = one launch sequence for the whole frame, at most 8.  The image does not depend on it. */
int rtp_set_tile_groups(rtp_scene *s, int groups);
int rtp_get_tile_groups(const rtp_scene *s); /* bands the most recently enqueued frame was cut into */

/* ---- the render entry point --------------------------------------------------------------- */
/* Replaces Flyscene::raytraceScene + traceRayThread (flyscene.cpp:250-314) for the full
 * W x H frame.  Host outputs, each optional (NULL to skip):
 *   rgb   W*H*3 float  pixel_data[j][i] (row-major)
 *   rgb8  W*H*3 uint8  the values writePPMImage prints: min(255,(int)(255*c)) (ppmIO.hpp:150)
 *   hit   W*H   int32  primary-hit face id, -1 = background
 * With world > 1 (one virtual rank of a sharded frame, scenes of rtp_scene_create only: tests and external
 * multi-process drivers) only this rank's tiles are rendered and copied; pixels of other ranks are left untouched in
 * the host buffers.  A scene of rtp_scene_create_multi shards the frame itself: world must be 1.
 * Limit: pixels per call x max(lights, 1) < 2^31 (shadow rays of a level are indexed in 32 bits) - RTP_ERR_INVALID
 * otherwise.  The camera may be arbitrarily far from the scene: the hierarchy's padding follows ulp(|eye|). */
int rtp_render(rtp_scene *s, const rtp_render_opts *o, float *rgb, uint8_t *rgb8, int32_t *hit,
               rtp_stats *st);

/* Multi-frame renders (a camera path): rtp_render_async returns as soon as the frame is enqueued, rtp_render_wait
 * blocks until the OLDEST outstanding frame has landed in its host buffers and returns its statistics.  Two
 * frames may be outstanding: the device image planes are double-buffered and the band copies run on their own
 * streams, so frame k+1 is traced while frame k is still on its way to the host.  rgb8 / hit as in rtp_render
 * (at least one; page-locked memory, e.g. cudaHostAlloc, for a real overlap); camera and lights are the ones set
 * when the call is made.  One GPU per scene (world must be 1).  No counterpart in the reference, whose
 * raytraceScene renders one frame per key press (main.cpp:68-69). */
int rtp_render_async(rtp_scene *s, const rtp_render_opts *o, uint8_t *rgb8, int32_t *hit);
int rtp_render_wait(rtp_scene *s, rtp_stats *st /* may be NULL */);

/* Device-resident variant used by multi-GPU drivers: renders this rank's interleaved screen
 * tiles into a compact tile-major buffer of float4 (r, g, b, bit-cast primary face id) in
 * device memory.  d_tiles: rtp_tile_buffer_bytes() bytes of device memory owned by the caller
 * (e.g. a torch tensor), stream: a cudaStream_t (NULL = default stream).  Asynchronous. */
size_t rtp_tile_buffer_bytes(int width, int height, int rank, int world);
/* Host-side description of the sharding (no device needed): the 32x32 screen tiles of the frame are
 * numbered row-major (tile id = ty * tiles_x + tx, tiles_x = ceil(width / 32)); tile (tx, ty) belongs
 * to rank (tx + ty) % world.  Fills tiles_of_rank (ids owned by `rank`, ascending - their order in
 * the rank's tile-major buffer) and local_of_tile (for EVERY tile id, its index within its owner's
 * buffer); either may be NULL.  Inside a tile, pixel (x, y) sits at float4 index
 * ((y >> 2) * 4 + (x >> 3)) * 32 + (y & 3) * 8 + (x & 7)  (8x4-pixel warp footprints).
 * Returns the number of tiles `rank` owns, or a negative rtp_status. */
int rtp_tile_table(int width, int height, int rank, int world, int32_t *tiles_of_rank, int32_t *local_of_tile);
int rtp_render_tiles_device(rtp_scene *s, const rtp_render_opts *o, void *d_tiles, void *stream,
                            rtp_stats *st);
/* After the per-rank tile buffers have been gathered (NCCL) into d_all = [world][max tile
 * bytes], scatter them to row-major device planes on the gathering rank.  Any output NULL to
 * skip.  d_rgb: W*H*3 float, d_rgb8: W*H*3 uint8, d_hit: W*H int32. */
int rtp_untile_device(rtp_scene *s, int width, int height, int world, const void *d_all,
                      size_t rank_stride_bytes, void *d_rgb, void *d_rgb8, void *d_hit, void *stream);

/* Fused variant of the two calls above: this rank's tiles are resolved straight into the finished
 * row-major image - d_rgb8 (W*H*3 uint8, the values writePPMImage prints, ppmIO.hpp:150) and/or d_hit
 * (W*H int32 primary face ids), either may be NULL - with 96-byte row segments per warp.  The
 * pointers may address ANOTHER GPU's memory (rtp_shared_open below): every rank then stores its tiles
 * into GPU 0's image over NVLink from inside the resolve kernel and no gather / un-tile pass is left;
 * the caller only has to order GPU 0's readers after all ranks' streams (a barrier).  Replaces the
 * join + image assembly of raytraceScene (flyscene.cpp:261-264,287-292).  Asynchronous on `stream`. */
int rtp_render_image_device(rtp_scene *s, const rtp_render_opts *o, void *d_rgb8, void *d_hit, void *stream,
                            rtp_stats *st);
/* Device buffers shareable between the processes of one box (CUDA IPC): the owner allocates and
 * exports a 64-byte handle, peers map it (peer access over NVLink is enabled on first use). */
#define RTP_SHARED_HANDLE_BYTES 64
int rtp_shared_alloc(int device, size_t bytes, void **d_ptr, void *handle /* RTP_SHARED_HANDLE_BYTES */);
int rtp_shared_open(int device, const void *handle, void **d_ptr);
int rtp_shared_free(int device, void *d_ptr, int opened /* 1: came from rtp_shared_open */);

/* Frame barrier of the peer-store route, made of the library's own kernel instead of a collective: `d_counter` is a
 * zero-initialised uint32 in memory every rank can reach (e.g. the tail of GPU 0's rtp_shared_alloc block).  Enqueued
 * on `stream` behind this rank's frame, one thread publishes the rank's image stores, adds 1 to the counter with a
 * system-scope atomic over NVLink and waits until it has reached `target` (= world x frame number, frames counted
 * from 1; the counter is never reset).  Work enqueued behind it on GPU 0 sees every rank's pixels.  A wait that is not
 * satisfied within timeout_ms gives up (no hang) and is counted: rtp_peer_barrier_timeouts (synchronises the device)
 * returns how many did, or a negative rtp_status.  Replaces the join of raytraceScene's threads (flyscene.cpp:287-289).
 * A timeout is STICKY: the counter stays short of the missing arrivals, so the frame it happened in is incomplete on
 * GPU 0 and every later wait would spin for the full timeout.  The next rtp_peer_barrier / rtp_render_image_device on
 * that scene therefore fails with RTP_ERR_STATE (the count lives in page-locked mapped memory: the check does not
 * synchronise the device).  To go on, every rank re-creates / re-opens the shared block (fresh counter) and calls
 * rtp_peer_barrier_reset. */
int rtp_peer_barrier(rtp_scene *s, void *d_counter, uint32_t target, int timeout_ms, void *stream);
int rtp_peer_barrier_timeouts(rtp_scene *s);
int rtp_peer_barrier_reset(rtp_scene *s);

/* ---- single-ray API ---------------------------------------------------------------------- */
typedef enum rtp_ray_mode {
  RTP_RAYS_SHADE = 0,   /* Flyscene::traceRay(origin, dir, 0, false)  -> rgb n x 3 (flyscene.cpp:317) */
  RTP_RAYS_CLOSEST = 1, /* calculateMinimumFace -> face n (int32), t n (flyscene.cpp:373)  */
  RTP_RAYS_ANYHIT = 2   /* shadow(P, L) -> occluded n (uint8)                (flyscene.cpp:510)  */
} rtp_ray_mode;
/* rays: n x 6 floats (origin, direction) in HOST memory; outputs in host memory. */
int rtp_trace_rays(rtp_scene *s, const float *rays, int64_t n, int mode, int max_depth,
                   const float background[3], float *rgb, int32_t *face, float *t, uint8_t *occluded,
                   rtp_stats *st);
/* Same with rays and outputs already resident on the device (bench / AO stress). */
int rtp_trace_rays_device(rtp_scene *s, const void *d_rays, int64_t n, int mode, void *d_face,
                          void *d_t, void *d_occluded, void *stream, rtp_stats *st);

/* rtp_trace_rays_device for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same rays, modes (CLOSEST or ANYHIT) and outputs as rtp_trace_rays_device, no statistics.  The call allocates
 * nothing, on the device or through any library: the scratch the trace kernels need (ray cursor, parked-ray counters
 * and RTP_NOALLOC_PARK_SLOTS parked-ray slots, 1 180 160 bytes) is a static __device__ array of the library that every
 * GPU loading it holds until exit.  Limits: n <= RTP_NOALLOC_MAX_RAYS per call (RTP_ERR_INVALID beyond, split the
 * batch).  Every call on a device shares that one array - also calls on different scenes: calls that overlap on the GPU
 * (different streams, no ordering between them) race and give undefined hits; calls on one stream, or ordered by
 * events, are fine.  The array keeps the last call's contents; the call clears what must start at zero itself
 * (cudaMemsetAsync on `stream`).  It does not wait for or order itself against the scene's frames (it reads the scene
 * arrays only), and - as rtp_trace_rays_device - it does not widen the hierarchy for far-away ray origins. */
#define RTP_NOALLOC_MAX_RAYS (1 << 22)
#define RTP_NOALLOC_PARK_SLOTS 4096
int rtp_trace_rays_device_noalloc(rtp_scene *s, const void *d_rays, int64_t n, int mode, void *d_face, void *d_t,
                                  void *d_occluded, void *stream);

/* ---- introspection ------------------------------------------------------------------------ */
/* BoundingBox::boxes view of the scene (BoundingBox.h:13): sizes, then copies. */
int rtp_scene_box_count(const rtp_scene *s);
int rtp_scene_get_boxes(const rtp_scene *s, float *boxes /* B x 6 */, int32_t *box_off /* B+1 */,
                        int32_t *box_faces /* F */);
int rtp_scene_info(const rtp_scene *s, rtp_stats *st);
/* Waits for the most recently enqueued frame (rtp_render_tiles_device / rtp_render) and returns
 * its ray counters and, if it ran with RTP_RENDER_PROFILE, the per-kernel device times. */
int rtp_frame_stats(rtp_scene *s, rtp_stats *st);
const char *rtp_last_error(void);
int rtp_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RTP_H */
