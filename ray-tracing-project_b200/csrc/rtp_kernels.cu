// rtp_kernels.cu - hand-written sm_100a kernels for the ray-tracing hot path.
//
// Wavefront pipeline, one bounce level = two persistent-warp kernels:
//   k_closest(d): depth 0 generates the primary ray of every owned pixel on the fly
//                 (traceRayThread, flyscene.cpp:299-314); deeper levels read the compacted
//                 reflection-ray list.  BVH2 closest-hit traversal = calculateMinimumFace
//                 (flyscene.cpp:373-396).
//   k_shade(d)  : per hit: P, interpolated normal, then for every light one any-hit
//                 traversal (= shadow, flyscene.cpp:510-526) and the Phong term
//                 (calcSingleColor / calculateColor, flyscene.cpp:542-566,603-614); spawns the
//                 reflection ray of the next level (flyscene.cpp:361-363) through a
//                 warp-aggregated append.
//   k_resolve   : folds the per-depth direct colours back with the reference's ks-of-deepest-hit
//                 rule (flyscene.cpp:355-370) and writes r,g,b,face into the tile-major
//                 framebuffer.
// FP32 CUDA cores only; no tensor cores (not a dense contraction).
#include "rtp_kernels.h"

namespace rtp {

namespace {

constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ int warp_fetch(int *cursor, int lane) {
  int base = 0;
  if (lane == 0) base = atomicAdd(cursor, 32);
  return __shfl_sync(kFull, base, 0);
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(kFull, v, o);
  return v;
}

// Primary ray of local path q: pixel (i, j) of an owned screen tile through the camera
// (traceRayThread, flyscene.cpp:301-308), or the q-th caller-supplied ray (traceRay API).
__device__ __forceinline__ bool primary_ray(const FrameParams &p, int q, v3 &o, v3 &d) {
  if (p.user_rays) {
    if (q >= p.n_user_rays) return false;
    const float *r = p.user_rays + 6 * (size_t)q;
    o = mk3(r[0], r[1], r[2]);
    d = mk3(r[3], r[4], r[5]);
    return true;
  }
  int tile = __ldg(p.tile_ids + (q >> 10));
  int x, y;
  tile_pixel(q & (kTilePixels - 1), x, y);
  const int i = (tile % p.tiles_x) * kTileDim + x;
  const int j = (tile / p.tiles_x) * kTileDim + y;
  o = mk3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
  if (i >= p.width || j >= p.height) return false;
  d = primary_dir(p.cam, i, j);
  return true;
}

__global__ void __launch_bounds__(128) k_closest(const FrameParams p, const int depth) {
  const int lane = threadIdx.x & 31;
  const int count = depth == 0 ? p.n_px : p.counters[kCtrPaths + depth];
  int *cursor = p.counters + kCtrFetch + 2 * depth;
  const float4 *ro = p.ray_o[depth & 1], *rd = p.ray_d[depth & 1];
  unsigned long long n_primary = 0;
  for (;;) {
    const int base = warp_fetch(cursor, lane);
    if (base >= count) break;
    const int q = base + lane;
    if (q >= count) continue;
    v3 o, d;
    bool valid = true;
    if (depth == 0) {
      valid = primary_ray(p, q, o, d);
      p.ksdeep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      o = xyz(ro[q]);
      d = xyz(rd[q]);
    }
    float t = INFINITY;
    int slot = -2;
    if (valid) {
      traverse<false>(p.S, o, d, o, t, slot);
      ++n_primary;
    }
    p.hit_t[q] = t;
    p.hit_slot[q] = slot;
    if (depth == 0) p.primface[q] = slot >= 0 ? __ldg(p.S.face_of + slot) : (valid ? -1 : -2);
  }
  if (depth == 0) {
    n_primary = warp_sum(n_primary);
    if (lane == 0 && n_primary) atomicAdd(p.stats + kStatPrimary, n_primary);
  }
}

__global__ void __launch_bounds__(128) k_shade(const FrameParams p, const int depth) {
  const int lane = threadIdx.x & 31;
  const int count = depth == 0 ? p.n_px : p.counters[kCtrPaths + depth];
  int *cursor = p.counters + kCtrFetch + 2 * depth + 1;
  const float4 *ro = p.ray_o[depth & 1], *rd = p.ray_d[depth & 1];
  float4 *no = p.ray_o[(depth + 1) & 1], *nd = p.ray_d[(depth + 1) & 1];
  const bool spawn_level = depth + 1 < p.max_depth;
  const int n_lights = p.S.n_lights;
  unsigned long long n_hits = 0;
  for (;;) {
    const int base = warp_fetch(cursor, lane);
    if (base >= count) break;
    const int q = base + lane;
    const int slot = q < count ? p.hit_slot[q] : -1;
    const bool hit = slot >= 0;
    v3 o2 = mk3(0.f, 0.f, 0.f), r = mk3(0.f, 0.f, 0.f);
    int pixel = 0;
    if (hit) {
      ++n_hits;
      v3 o, d;
      if (depth == 0) {
        pixel = q;
        primary_ray(p, q, o, d);
      } else {
        float4 a = ro[q];
        o = xyz(a);
        pixel = f2i(a.w);
        d = xyz(rd[q]);
      }
      const float t = p.hit_t[q];
      const v3 P = vadd(o, vmuls(d, t));  // flyscene.cpp:464
      const float4 *T = p.S.tris + 4 * (size_t)slot;
      const float4 *Nn = p.S.vnorm + 3 * (size_t)slot;
      const float4 T2 = __ldg(T + 2);
      const v3 normal = interpolate_normal(xyz(__ldg(T)), xyz(__ldg(T + 1)), xyz(T2), xyz(__ldg(T + 3)),
                                           xyz(__ldg(Nn)), xyz(__ldg(Nn + 1)), xyz(__ldg(Nn + 2)), P);
      const unsigned mat = (unsigned)f2i(T2.w) & kTriMatMask;
      const float4 m0 = __ldg(p.S.materials + 3 * mat), m1 = __ldg(p.S.materials + 3 * mat + 1),
                   m2 = __ldg(p.S.materials + 3 * mat + 2);
      v3 sum = mk3(0.f, 0.f, 0.f);
      for (int li = 0; li < n_lights; ++li) {  // calculateColor, flyscene.cpp:606-612
        const float4 l0 = __ldg(p.S.lights + 2 * li), l1 = __ldg(p.S.lights + 2 * li + 1);
        v3 L = xyz(l0);
        if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));  // :607
        const v3 inter = vadd(P, vmuls(L, (float)0.003));     // :512
        float ts;
        int ss;
        if (traverse<true>(p.S, inter, L, P, ts, ss)) continue;  // shadowed: contributes nothing (:543-545)
        sum = vadd(sum, phong(xyz(m0), xyz(m1), xyz(m2), m0.w, normal, L, xyz(l1), o, P));
      }
      p.direct[(size_t)depth * p.n_px + pixel] = make_float4(clamp01(sum.x), clamp01(sum.y), clamp01(sum.z), 0.f);
      p.ksdeep[pixel] = make_float4(m2.x, m2.y, m2.z, i2f(depth + 1));
      if (spawn_level) {
        r = reflect_dir(d, normal);                 // :361
        o2 = vadd(P, vmuls(r, (float)0.001));       // :362
      }
    }
    const bool spawn = hit && spawn_level;
    const unsigned m = __ballot_sync(kFull, spawn);
    if (m) {
      const int leader = __ffs(m) - 1;
      int dst = 0;
      if (lane == leader) dst = atomicAdd(p.counters + kCtrPaths + depth + 1, __popc(m));
      dst = __shfl_sync(kFull, dst, leader) + __popc(m & ((1u << lane) - 1u));
      if (spawn) {
        no[dst] = make_float4(o2.x, o2.y, o2.z, i2f(pixel));
        nd[dst] = make_float4(r.x, r.y, r.z, 0.f);
      }
    }
  }
  n_hits = warp_sum(n_hits);
  if (lane == 0 && n_hits) {
    atomicAdd(p.stats + kStatHits, n_hits);
    atomicAdd(p.stats + kStatShadow, n_hits * (unsigned long long)n_lights);
    if (spawn_level) atomicAdd(p.stats + kStatReflection, n_hits);
  }
}

__global__ void __launch_bounds__(256) k_resolve(const FrameParams p) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.n_px) return;
  const float4 kd = p.ksdeep[q];
  const int k = f2i(kd.w);
  const int face = p.primface[q];
  float4 out;
  if (face == -2) {
    out = make_float4(0.f, 0.f, 0.f, i2f(-1));
  } else if (k == 0) {
    out = make_float4(p.bg[0], p.bg[1], p.bg[2], i2f(-1));  // flyscene.cpp:327-330
  } else {
    // deepest level: clamp(direct + 0 * ks) (the recursion below it returned (0,0,0))
    float4 dc = p.direct[(size_t)(k - 1) * p.n_px + q];
    v3 c = mk3(clamp01(fadd(dc.x, fmul(0.f, kd.x))), clamp01(fadd(dc.y, fmul(0.f, kd.y))),
               clamp01(fadd(dc.z, fmul(0.f, kd.z))));
    for (int d = k - 2; d >= 0; --d) {  // flyscene.cpp:366-370 with the member ks of the deepest hit
      dc = p.direct[(size_t)d * p.n_px + q];
      c = mk3(clamp01(fadd(dc.x, fmul(c.x, kd.x))), clamp01(fadd(dc.y, fmul(c.y, kd.y))),
              clamp01(fadd(dc.z, fmul(c.z, kd.z))));
    }
    out = make_float4(c.x, c.y, c.z, i2f(face));
  }
  p.fb[q] = out;
}

__global__ void __launch_bounds__(256) k_untile(const float4 *all, size_t rank_stride_f4, const int *local_of_tile,
                                                int width, int height, int tiles_x, int world, int only_rank,
                                                float *rgb, unsigned char *rgb8, int *hit) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)width * height) return;
  const int i = (int)(idx % width), j = (int)(idx / width);
  const int tx = i / kTileDim, ty = j / kTileDim, x = i % kTileDim, y = j % kTileDim;
  const int owner = tile_owner(tx, ty, world);
  if (only_rank >= 0 && owner != only_rank) return;
  const int local = __ldg(local_of_tile + ty * tiles_x + tx);
  const int pix = (((y >> 2) * 4 + (x >> 3)) << 5) + ((y & 3) << 3) + (x & 7);
  const float4 v = all[(size_t)(only_rank >= 0 ? 0 : owner) * rank_stride_f4 + (size_t)local * kTilePixels + pix];
  if (rgb) { rgb[3 * idx] = v.x; rgb[3 * idx + 1] = v.y; rgb[3 * idx + 2] = v.z; }
  if (rgb8) {  // ppmIO.hpp:150: min(255, (int)(255*c)); NaN -> 0 (undefined in the reference)
    float c[3] = {v.x, v.y, v.z};
    for (int k = 0; k < 3; ++k) {
      float s = fmul(255.f, c[k]);
      int qv = (s != s) ? 0 : __float2int_rz(s);
      rgb8[3 * idx + k] = (unsigned char)min(255, max(qv, 0));
    }
  }
  if (hit) hit[idx] = f2i(v.w);
}

__global__ void __launch_bounds__(128) k_rays_closest(const DeviceScene S, const float *rays, long long n, int *face,
                                                      float *t, int *cursor) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    const long long base = warp_fetch(cursor, lane);
    if (base >= n) break;
    const long long q = base + lane;
    if (q >= n) continue;
    const float *r = rays + 6 * q;
    v3 o = mk3(r[0], r[1], r[2]), d = mk3(r[3], r[4], r[5]);
    float tt;
    int slot;
    traverse<false>(S, o, d, o, tt, slot);
    face[q] = slot >= 0 ? __ldg(S.face_of + slot) : -1;
    if (t) t[q] = tt;
  }
}

__global__ void __launch_bounds__(128) k_rays_anyhit(const DeviceScene S, const float *rays, long long n,
                                                     unsigned char *occ, int *cursor) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    const long long base = warp_fetch(cursor, lane);
    if (base >= n) break;
    const long long q = base + lane;
    if (q >= n) continue;
    const float *r = rays + 6 * q;
    v3 P = mk3(r[0], r[1], r[2]), L = mk3(r[3], r[4], r[5]);
    const v3 inter = vadd(P, vmuls(L, (float)0.003));  // shadow(P, L), flyscene.cpp:512
    float tt;
    int slot;
    occ[q] = traverse<true>(S, inter, L, P, tt, slot) ? 1 : 0;
  }
}

}  // namespace

void query_launch_cfg(int device, LaunchCfg &cfg) {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  int b0 = 1, b1 = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b0, k_closest, cfg.threads, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, k_shade, cfg.threads, 0);
  cfg.blocks_closest = sms * (b0 > 0 ? b0 : 1);
  cfg.blocks_shade = sms * (b1 > 0 ? b1 : 1);
}

void launch_closest(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  k_closest<<<cfg.blocks_closest, cfg.threads, 0, st>>>(p, depth);
}
void launch_shade(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  k_shade<<<cfg.blocks_shade, cfg.threads, 0, st>>>(p, depth);
}
void launch_resolve(const FrameParams &p, cudaStream_t st) {
  k_resolve<<<(p.n_px + 255) / 256, 256, 0, st>>>(p);
}
void launch_untile(const float4 *all, size_t rank_stride_f4, const int *local_of_tile, int width, int height,
                   int world, int only_rank, float *rgb, unsigned char *rgb8, int *hit, cudaStream_t st) {
  const long long n = (long long)width * height;
  const int tiles_x = (width + kTileDim - 1) / kTileDim;
  k_untile<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(all, rank_stride_f4, local_of_tile, width, height, tiles_x,
                                                        world, only_rank, rgb, rgb8, hit);
}
void launch_rays_closest(const DeviceScene &S, const float *rays, long long n, int *face, float *t, int *cursor,
                         const LaunchCfg &cfg, cudaStream_t st) {
  k_rays_closest<<<cfg.blocks_closest, cfg.threads, 0, st>>>(S, rays, n, face, t, cursor);
}
void launch_rays_anyhit(const DeviceScene &S, const float *rays, long long n, unsigned char *occ, int *cursor,
                        const LaunchCfg &cfg, cudaStream_t st) {
  k_rays_anyhit<<<cfg.blocks_closest, cfg.threads, 0, st>>>(S, rays, n, occ, cursor);
}

}  // namespace rtp
