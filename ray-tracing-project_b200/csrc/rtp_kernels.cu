// rtp_kernels.cu - hand-written sm_100a kernels for the ray-tracing hot path.
//
// Wavefront pipeline.  Per bounce level d:
//   k_trace<FrameClosest>  calculateMinimumFace for every ray of the level (flyscene.cpp:373-396).  At level 0 the
//                          ray is generated from the pixel id when a lane takes it (traceRayThread + screenToWorld,
//                          flyscene.cpp:299-314: a primary ray is a pure function of the pixel); when a ray retires, the
//                          lane that still holds the triangle computes P = o + t*d and the interpolated normal
//                          (flyscene.cpp:464,572-600) and appends the hit record to the compacted hit list.
//   k_trace<FrameAny>      shadow() for every (hit, light) pair (flyscene.cpp:510-526)
//   k_accum                calcSingleColor / calculateColor and the reflection-ray spawn
//                          (flyscene.cpp:542-566,603-614,355-363)
// and k_resolve_image after the last level (fold-back with the reference's ks-of-deepest-hit rule,
// flyscene.cpp:366-370, quantise, store into the image).  With max_depth 1 the last two are ONE kernel
// (k_shade_image): a frame is closest -> shadow -> shade+image, three dependent launches (+ k_coop).
// (Round 1 ran k_raygen / k_prep / k_accum / k_resolve as four separate streaming kernels: 20 % of the C3 frame.)
//
// The two trace kernels are persistent: one CTA slot per SM resident set, every warp pulls rays
// from a global cursor and REFILLS idle lanes (ballot + warp-aggregated atomic) whenever fewer
// than kRefillBelow lanes still hold a ray, so divergent ray lengths do not leave lanes empty.
// Per-thread short stack in local memory (L1-resident), 128-bit read-only loads of 64-byte nodes
// and triangles.  FP32 CUDA cores only; no tensor cores (not a dense contraction).
#include "rtp_kernels.h"

#include <algorithm>
#include <cstdlib>
#include <mutex>

namespace rtp {

namespace {

constexpr unsigned kFull = 0xffffffffu;

// resident CTAs per SM of the persistent trace kernels (128 threads each): 8 = 64 registers
#ifndef RTP_TRACE_MINB
#define RTP_TRACE_MINB 8
#endif
#ifndef RTP_RAYS_ANY_MINB
#define RTP_RAYS_ANY_MINB 9
#endif
#ifndef RTP_TRACE_MINB_ANY  // resident CTAs per SM of the any-hit instantiations (experiment builds: 9 = 56 registers)
#define RTP_TRACE_MINB_ANY RTP_TRACE_MINB
#endif

// ---- jobs: where a trace kernel gets its rays from and where results go ----------------------
// CAM = true: level 0 of a camera frame - ray q is pixel q of this rank's tile-major pixel list, generated here.
// CAM = false: rays come from the level's ray list (reflection levels, caller-supplied rays).
template <bool CAM>
struct FrameClosest {
  static constexpr bool kAny = false;
  static constexpr bool kSub = false;
  static constexpr bool kFinish = false;  // no per-unit epilogue in k_trace
  static constexpr int kMinBlocks = RTP_TRACE_MINB;
  DeviceScene S;
  // CAM
  DeviceCamera cam;
  const int *tile_ids;
  int tiles_x, width, height;
  // !CAM
  const float4 *ro, *rd;
  // results: per ray the index of its hit record (-1 miss, -2 no ray); per hit record (P, triangle slot),
  // (direction, pixel) and the path index.  CAM: record h of pixel q is q itself (the level's any-hit and shading
  // kernels walk the pixels in order and skip the misses: neighbouring records are neighbouring pixels, no atomic
  // in the retire path); otherwise the records are compacted (reflection levels lose most of their paths).
  int *hit_slot;
  float4 *pbuf, *dbuf;
  int *hitq;
  int *hits_ctr;
  int *primface;   // level 0 only: primary face id per pixel (-1 background, -2 pixel outside the image)
  float4 *ksdeep;  // level 0 only: zeroed for pixels without a hit (k_accum / k_shade_image write the others)
  int *cursor;
  const int *count_ptr;  // device counter (depth > 0) or null
  int count_fixed;       // = n_px: the capacity of every per-path and per-hit-record buffer above
  __device__ __forceinline__ int count() const { return count_ptr ? *count_ptr : count_fixed; }
  __device__ __forceinline__ bool init(int q, Trav &t, int *stack) const {
    RTP_CHECK((unsigned)q < (unsigned)count_fixed, 10, q, count_fixed);
    if (CAM) {
      const int tile = __ldg(tile_ids + (q >> 10));
      int x, y;
      tile_pixel(q & (kTilePixels - 1), x, y);
      const int i = (tile % tiles_x) * kTileDim + x;
      const int j = (tile / tiles_x) * kTileDim + y;
      if (i >= width || j >= height) {  // tile padding: pixel outside the image
        hit_slot[q] = -2;
        primface[q] = -2;
        ksdeep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        return false;
      }
      trav_init(t, mk3(cam.eye[0], cam.eye[1], cam.eye[2]), primary_dir(cam, i, j), stack);
      return true;
    }
    const float4 d = rd[q];
    if (d.w == 0.f) {  // no ray (padding of a caller-supplied batch; k_raygen set primface / ksdeep)
      hit_slot[q] = -2;
      return false;
    }
    trav_init(t, xyz(ro[q]), xyz(d), stack);
    return true;
  }
  __device__ __forceinline__ v3 box_origin(int, const Trav &t) const { return t.o; }
  // the deferred BoundingBox check of the winner (closest_needs_rerun): true = the ray runs again, careful
  __device__ __forceinline__ bool rerun(Trav &t, int *stack) const { return closest_needs_rerun(S, t, stack); }
  // Retires the rays of the lanes with fin set; called by all 32 lanes together.
  __device__ __forceinline__ void commit(int q, const Trav &t, bool fin) const {
    const int lane = threadIdx.x & 31;
    const bool hit = fin && t.sbest >= 0;
    const unsigned m = __ballot_sync(kFull, hit);
    int h = q;
    if (m) {
      const int leader = __ffs(m) - 1;
      if (CAM) {
        if (lane == leader) atomicAdd(hits_ctr, __popc(m));  // statistics only: nobody waits for the result
      } else {
        if (lane == leader) h = atomicAdd(hits_ctr, __popc(m));
        h = __shfl_sync(kFull, h, leader) + __popc(m & ((1u << lane) - 1u));
      }
    }
    if (hit) {
      RTP_CHECK((unsigned)h < (unsigned)count_fixed, 11, h, count_fixed);
      RTP_CHECK((unsigned)t.sbest < (unsigned)S.n_tris, 12, t.sbest, S.n_tris);
      const v3 P = vadd(t.o, vmuls(t.d, t.tbest));  // flyscene.cpp:464
      pbuf[h] = make_float4(P.x, P.y, P.z, i2f(t.sbest));
      dbuf[h] = make_float4(t.d.x, t.d.y, t.d.z, i2f(CAM ? q : f2i(ro[q].w)));
      if (!CAM) hitq[h] = q;
    }
    if (fin) {
      hit_slot[q] = hit ? h : -1;
      if (primface) {
        primface[q] = hit ? __ldg(S.face_of + t.sbest) : -1;
        if (!hit) ksdeep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
};

// DENSE: the level's hit records are indexed by pixel (camera rays): shadow ray q = pixel * n_lights + light, and pixels
// without a hit have no shadow rays.  In batch mode the warp then takes 32 PIXELS from the cursor and walks the lights
// itself (k_trace, kSubs): one atomic per 32 pixels whatever the light count, 32 rays towards the SAME light per batch
// (coherent), and a batch of pixels without any hit costs a ballot, not n_lights trips to the cursor.
template <bool DENSE>
struct FrameAny {
  static constexpr bool kAny = true;
  static constexpr bool kSub = DENSE;
  static constexpr bool kFinish = false;  // no per-unit epilogue in k_trace
  static constexpr int kMinBlocks = RTP_TRACE_MINB_ANY;
  DeviceScene S;
  const float4 *pbuf;
  unsigned char *occl;
  int *cursor;
  const int *hits_ptr;
  const int *hit_slot;  // DENSE: hit_slot[pixel] < 0 = no hit, no shadow ray
  int n_px;             // DENSE: pixels of the level
  __device__ __forceinline__ int units() const { return n_px; }
  __device__ __forceinline__ int subs() const { return S.n_lights; }
  __device__ __forceinline__ int count() const { return (DENSE ? n_px : *hits_ptr) * S.n_lights; }
  __device__ __forceinline__ v3 light_dir(int q, v3 P) const {
    const int li = q % S.n_lights;
    const float4 l0 = __ldg(S.lights + 2 * li);
    v3 L = xyz(l0);
    if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));  // flyscene.cpp:607
    return L;
  }
  __device__ __forceinline__ bool init(int q, Trav &t, int *stack) const {
    const int h = q / S.n_lights;
    RTP_CHECK((unsigned)h < (unsigned)n_px, 13, h, n_px);
    if (DENSE && hit_slot[h] < 0) return false;
    const v3 P = xyz(pbuf[h]);
    const v3 L = light_dir(q, P);
    trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);  // flyscene.cpp:512
    return true;
  }
  __device__ __forceinline__ v3 box_origin(int q, const Trav &) const { return xyz(pbuf[q / S.n_lights]); }
  __device__ __forceinline__ bool rerun(Trav &, int *) const { return false; }
  __device__ __forceinline__ void commit(int q, const Trav &t, bool fin) const { if (fin) occl[q] = t.sbest >= 0 ? 1 : 0; }
};

struct RaysClosest {  // rtp_trace_rays(CLOSEST): rays n x 6 floats
  static constexpr bool kAny = false;
  static constexpr bool kSub = false;
  static constexpr bool kFinish = false;  // no per-unit epilogue in k_trace
  static constexpr int kMinBlocks = RTP_TRACE_MINB;
  DeviceScene S;
  const float *rays;
  int *face;
  float *tout;
  int *cursor;
  int n;
  __device__ __forceinline__ int count() const { return n; }
  __device__ __forceinline__ bool init(int q, Trav &t, int *stack) const {
    RTP_CHECK((unsigned)q < (unsigned)n, 14, q, n);
    const float *r = rays + 6 * (size_t)q;
    trav_init(t, mk3(r[0], r[1], r[2]), mk3(r[3], r[4], r[5]), stack);
    return true;
  }
  __device__ __forceinline__ v3 box_origin(int, const Trav &t) const { return t.o; }
  __device__ __forceinline__ bool rerun(Trav &t, int *stack) const { return closest_needs_rerun(S, t, stack); }
  __device__ __forceinline__ void commit(int q, const Trav &t, bool fin) const {
    if (!fin) return;
    face[q] = t.sbest >= 0 ? __ldg(S.face_of + t.sbest) : -1;
    if (tout) tout[q] = t.tbest;
  }
};

struct RaysAny {  // rtp_trace_rays(ANYHIT): shadow(P, L) for rays (P, L)
  static constexpr bool kAny = true;
  static constexpr bool kSub = false;
  static constexpr bool kFinish = false;  // no per-unit epilogue in k_trace
  // scattered any-hit batches are the one launch that gains from a ninth CTA per SM (56 registers, 16 bytes of spill
  // outside the node loop): C5 20.40 -> 19.75 ms; the frame kernels lose 1 % (profiles/r02_design_items.md)
  static constexpr int kMinBlocks = RTP_RAYS_ANY_MINB;
  DeviceScene S;
  const float *rays;
  unsigned char *occ;
  int *cursor;
  int n;
  __device__ __forceinline__ int count() const { return n; }
  __device__ __forceinline__ bool init(int q, Trav &t, int *stack) const {
    RTP_CHECK((unsigned)q < (unsigned)n, 15, q, n);
    const float *r = rays + 6 * (size_t)q;
    const v3 P = mk3(r[0], r[1], r[2]), L = mk3(r[3], r[4], r[5]);
    trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);
    return true;
  }
  __device__ __forceinline__ v3 box_origin(int q, const Trav &) const {
    const float *r = rays + 6 * (size_t)q;
    return mk3(r[0], r[1], r[2]);
  }
  __device__ __forceinline__ bool rerun(Trav &, int *) const { return false; }
  __device__ __forceinline__ void commit(int q, const Trav &t, bool fin) const { if (fin) occ[q] = t.sbest >= 0 ? 1 : 0; }
};

// ---- stragglers: step budget, parking list, warp-cooperative traversal (k_coop) ------------------
// A ray that is still running when every other lane of its warp has finished is a chain of
// dependent node fetches (a cold node costs a DRAM round trip, ~0.6 us on B200) executed by one lane
// while 31 sit idle; a single 650-node ray used to set the duration of the whole closest-hit kernel,
// and the max-of-32 last batches that of the shadow kernel (profiles/r01_tail.md).
//
// k_trace therefore PARKS stragglers: when a warp has spent more than a budget of steps on its rays
// and only a few lanes still hold one, those rays - id, best hit so far, pending traversal stack -
// are appended to a global list and the lanes are free.  k_coop, launched right behind, finishes
// every parked ray with a whole warp (coop_trace): a shared-memory stack, every iteration the 32
// lanes pop up to 32 entries, test them in parallel (inner node: both child slabs; leaf: the exact
// triangle tests) and push the surviving children back, the near child of the top entry last (a
// 32-wide speculative depth-first order).  Candidates merge through a lexicographic (t, reference
// rank) warp minimum - exactly the order-independent rule of trav_step - so the result is the one
// the serial traversal gives.
// (Tried first and dropped: finishing parked rays inside k_trace.  A shared ring polled by idle
// warps made every reservation cost 10-70 us; an in-kernel call or inlined copy of coop_trace pushed
// the node loop of trav_step into local-memory spills at 64 registers, -25 % on the whole frame.)
constexpr int kCoopSoft = 160;                             // above this fill level only one entry is popped per iteration
constexpr int kCoopStack = kCoopSoft + kStackSize + 32;    // depth-first growth bound + slack
constexpr int kParkInts = 5 + kStackSize + 3;              // (ray id + 1) | careful << 31, tbest, rank, slot, n, entries
static_assert(kParkInts * sizeof(int) == kParkSlotBytes, "parked-ray slot size");

struct ParkList {
  int *slots;   // kParkInts ints per parked ray
  int *ctr;     // [0] rays parked by k_trace, [1] k_coop's cursor (zeroed per frame, one pair per launch)
  int cap;      // slots available; a ray that finds the list full simply carries on
  int budget;   // steps a warp spends on a batch before its last lanes count as stragglers (huge = never park)
  int lanes;    // ... and at most this many lanes still hold a ray
  int step_nodes;  // inner nodes per trav_step (kMaxNodesPerStep for coherent batches; RTP_STEP_NODES for scattered rays)
};

// Finishes the (warp-uniform) ray in t with all 32 lanes, starting from the n pending entries in ws;
// t.tbest/rbest/sbest hold the best hit found so far on entry and the final one on return.
template <bool ANY>
__device__ __forceinline__ void coop_trace(const DeviceScene &S, Trav &t, const v3 bo, int *ws, int n) {
  const int lane = threadIdx.x & 31;
  const unsigned gt = lane == 31 ? 0u : (kFull << (lane + 1));  // lanes above this one
  const float idx = t.idx, idy = t.idy, idz = t.idz, oodx = t.oodx, oody = t.oody, oodz = t.oodz;
  const v3 o = t.o, d = t.d;
  const bool careful = t.careful;
  float tb = t.tbest;
  unsigned rb = t.rbest;
  int sb = t.sbest;
  while (n > 0) {
    const int room = kCoopSoft - n;
    const int k = min(min(32, n), room > 1 ? room : 1);
    const int my = lane < k ? ws[n - 1 - lane] : kSentinel;  // lane 0 takes the top of the stack
    n -= k;
    __syncwarp();
    int push_far = kSentinel, push_near = kSentinel;
    bool improved = false;
    if (lane < k) {
      if (my >= 0) {
        const float4 *Np = S.nodes + 4 * (size_t)my;
        RTP_CHECK(my < S.n_nodes, 21, my, S.n_nodes);
        const float4 n0 = __ldg(Np), n1 = __ldg(Np + 1), n2 = __ldg(Np + 2), n3 = __ldg(Np + 3);
        const float c0lox = fmaf(n0.x, idx, -oodx), c0hix = fmaf(n0.y, idx, -oodx);
        const float c0loy = fmaf(n0.z, idy, -oody), c0hiy = fmaf(n0.w, idy, -oody);
        const float c0loz = fmaf(n2.x, idz, -oodz), c0hiz = fmaf(n2.y, idz, -oodz);
        const float c1lox = fmaf(n1.x, idx, -oodx), c1hix = fmaf(n1.y, idx, -oodx);
        const float c1loy = fmaf(n1.z, idy, -oody), c1hiy = fmaf(n1.w, idy, -oody);
        const float c1loz = fmaf(n2.z, idz, -oodz), c1hiz = fmaf(n2.w, idz, -oodz);
        const float c0min = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), 0.f));
        const float c0max = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), tb));
        const float c1min = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), 0.f));
        const float c1max = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), tb));
        const bool h0 = c0min <= c0max, h1 = c1min <= c1max;
        const int ch0 = f2i(n3.x), ch1 = f2i(n3.y);
        if (h0 && h1) {
          const bool swap = c1min < c0min;
          push_near = swap ? ch1 : ch0;
          push_far = swap ? ch0 : ch1;
        } else if (h0 || h1) {
          push_near = h0 ? ch0 : ch1;
        }
      } else {
        const int enc = ~my;
        const int first = enc >> 3, cnt = (enc & 7) + 1;
        RTP_CHECK(first + cnt <= S.n_tris, 22, first + cnt, S.n_tris);
        for (int j = 0; j < cnt; ++j) {
          float tt;
          float4 T0, T1;
          if (!tri_exact(S, first + j, o, d, tb, tt, T0, T1)) continue;
          const unsigned rank = (unsigned)f2i(T0.w);
          if (!ANY && !(tt < tb || (sb >= 0 && rank < rb))) continue;
          if ((ANY || careful) && !ref_box_pass(S, pinned(bo), pinned(d), f2i(T1.w))) continue;
          tb = ANY ? tb : tt;
          rb = rank;
          sb = first + j;
          improved = true;
          if (ANY) break;
        }
      }
    }
    const unsigned found = __ballot_sync(kFull, improved);
    if (found) {
      if constexpr (ANY) {
        sb = __shfl_sync(kFull, sb, __ffs(found) - 1);
        break;
      } else {
#pragma unroll
      for (int m = 16; m > 0; m >>= 1) {  // lexicographic (t, rank) minimum; every lane started from the warp's best
        const float t2 = __shfl_xor_sync(kFull, tb, m);
        const unsigned r2 = __shfl_xor_sync(kFull, rb, m);
        const int s2 = __shfl_xor_sync(kFull, sb, m);
        if (s2 >= 0 && (sb < 0 || t2 < tb || (t2 == tb && r2 < rb))) { tb = t2; rb = r2; sb = s2; }
      }
      }
    }
    const bool pf = push_far != kSentinel, pn = push_near != kSentinel;
    const unsigned bf = __ballot_sync(kFull, pf), bn = __ballot_sync(kFull, pn);
    const int pos = n + __popc(bf & gt) + __popc(bn & gt);  // entries of higher lanes go below
    RTP_CHECK(pos + 1 < kCoopStack, 20, pos + 1, kCoopStack);
    if (pf) ws[pos] = push_far;
    if (pn) ws[pos + (pf ? 1 : 0)] = push_near;
    n += __popc(bf) + __popc(bn);
    __syncwarp();
  }
  __syncwarp();
  t.tbest = tb; t.rbest = rb; t.sbest = sb;
  t.node = kSentinel; t.sp = 0;
}

// ---- the persistent trace kernel -----------------------------------------------------------------
#ifdef RTP_TRACE_TIMING  // debug builds only (scratch A/B runs): per-warp start / end / last-fetch times
__device__ unsigned long long g_warp_timing[2 * 8192 * 8];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#endif
template <class Job, int kRefillBelow>
__global__ void __launch_bounds__(128, Job::kMinBlocks) k_trace(const Job job, const ParkList pl) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  // kSubs (batch mode of a job with sub-batches): the cursor hands out 32 UNITS (pixels) and the warp walks each
  // unit's `n_sub` rays (lights) itself: ray q = unit * n_sub + sub
  constexpr bool kSubs = Job::kSub && kRefillBelow == 1;
  int n_sub = 1, sub = 1, ubase = 0;
  if constexpr (kSubs) { n_sub = job.subs(); sub = n_sub; }
  int count;
  if constexpr (kSubs) count = job.units(); else count = job.count();
#ifdef RTP_TRACE_TIMING
  const unsigned long long tm_start = gtime();
  unsigned long long tm_fetch = tm_start;
  unsigned tm_nb = 0, tm_nc = 0;
#endif
#ifdef RTP_SMEM_STACK
  __shared__ int s_stack[RTP_SMEM_STACK * 128];
  int *const stack = s_stack + threadIdx.x;
#else
  int stack[kStackSize];
#endif
  Trav t;
  int q = 0, age = 0;  // age: steps since this warp last took rays (warp-uniform)
  bool have = false, exhausted = false;
  // Job::kFinish (a job with sub-batches in batch mode): bit s of occ = sub-ray s of this lane's unit found a hit; when the
  // warp has walked all subs of its 32 units, job.finish(unit, occ) runs the units' epilogue (FrameAnyShade: shading)
  unsigned occ = 0;
  bool open = false;  // a batch of units has been taken and not finished yet
  for (;;) {
    // lanes whose traversal is still going (a finished one keeps its result in t until it is retired)
    unsigned act = __ballot_sync(kFull, have && t.node != kSentinel);
    if (__popc(act) < kRefillBelow) {
      // retire: every lane whose traversal has ended writes its result NOW, together - the deferred
      // BoundingBox check of a closest hit is ~60 instructions and used to run one lane at a time
      bool fin = have && t.node == kSentinel;
      if (fin && job.rerun(t, stack)) fin = false;  // careful re-run: node = root again
      job.commit(q, t, fin);                        // all lanes: the warp's hit records are appended together
      if constexpr (Job::kFinish) {
        if (fin && t.sbest >= 0) occ |= 1u << (sub - 1);
      }
      if (fin) have = false;
      act = __ballot_sync(kFull, have);
      if constexpr (kSubs) {
        if (act == 0 && (!exhausted || sub < n_sub)) {
          if (sub >= n_sub) {  // next 32 units
            if constexpr (Job::kFinish) {
              if (open) job.finish(ubase + lane, occ);
              occ = 0;
            }
            int base = 0;
            if (lane == 0) base = atomicAdd(job.cursor, 32);
            ubase = __shfl_sync(kFull, base, 0);
            exhausted = ubase + 32 >= count;
            open = ubase < count;
            sub = 0;
          }
          const int u = ubase + lane;
          if (u < count) {
            q = u * n_sub + sub;
            have = job.init(q, t, stack);
          }
          act = __ballot_sync(kFull, have);
          sub = (sub == 0 && act == 0) ? n_sub : sub + 1;  // no unit of the batch has rays at all: skip its other subs
          age = 0;
        }
      } else
      if (!exhausted && __popc(act) < kRefillBelow) {
        // refill every idle lane with the next rays of the global list
        const unsigned need = ~act;
        const int n_need = __popc(need);
        const int leader = __ffs(need) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(job.cursor, n_need);
        base = __shfl_sync(kFull, base, leader);
        exhausted = base + n_need >= count;
#ifdef RTP_TRACE_TIMING
        if (base < count) { tm_fetch = gtime(); ++tm_nb; }
#endif
        if (!have) {
          q = base + __popc(need & lt);
          if (q < count) have = job.init(q, t, stack);
        }
        act = __ballot_sync(kFull, have);
        age = 0;
      }
    }
    if (act == 0) {
      if (exhausted && sub >= n_sub) {
        if constexpr (Job::kFinish) {
          if (open) job.finish(ubase + lane, occ);
        }
        break;
      }
      continue;
    }
    if (++age > pl.budget && __popc(act) <= pl.lanes) {
      // only stragglers are left in this warp: park them for k_coop
      if (have && t.node != kSentinel) {
        const int at = atomicAdd(pl.ctr, 1);
        if (at < pl.cap) {
          int *e = pl.slots + (size_t)at * kParkInts;
          e[0] = (q + 1) | (t.careful ? (int)0x80000000 : 0);
          e[1] = f2i(t.tbest);
          e[2] = (int)t.rbest;
          e[3] = t.sbest;
          int n = 0;
          for (int i = 1; i <= t.sp; ++i) e[5 + n++] = RTP_STK(stack, i);
          if (t.node != kSentinel) e[5 + n++] = t.node;
          RTP_CHECK(5 + n <= kParkInts, 23, 5 + n, kParkInts);
          e[4] = n;
          have = false;
#ifdef RTP_TRACE_TIMING
          ++tm_nc;
#endif
        }
      }
      age = 0;
      continue;
    }
    if (have && t.node != kSentinel) trav_step<Job::kAny ? 1 : 0>(job.S, t, stack, [&] { return job.box_origin(q, t); }, pl.step_nodes);
  }
#ifdef RTP_TRACE_TIMING
  if (lane == 0) {
    const unsigned wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid < 8192) {
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      unsigned long long *o = g_warp_timing + ((Job::kAny ? 8192 : 0) + wid) * 8;
      o[0] = tm_start; o[1] = gtime(); o[2] = tm_fetch - tm_start;
      o[3] = smid | ((unsigned long long)tm_nb << 16) | ((unsigned long long)tm_nc << 32);
      o[4] = 0; o[5] = 0; o[6] = 0; o[7] = 0;
    }
  }
#endif
}

// Finishes the rays k_trace parked, one warp per ray.
template <class Job>
__global__ void __launch_bounds__(128, 4) k_coop(const Job job, const ParkList pl) {
  __shared__ int s_ws[4][kCoopStack];
  int *const ws = s_ws[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int n_parked = min(*pl.ctr, pl.cap);
  int stack[1];  // job.rerun -> closest_needs_rerun -> trav_init writes stack[0] only
  for (;;) {
    int k = 0;
    if (lane == 0) k = atomicAdd(pl.ctr + 1, 1);
    k = __shfl_sync(kFull, k, 0);
    if (k >= n_parked) break;
    const int *e = pl.slots + (size_t)k * kParkInts;
    const int ex = e[0];
    const int q = (ex & 0x7fffffff) - 1;
    Trav t;
    job.init(q, t, stack);  // the same ray in every lane
    t.tbest = i2f(e[1]); t.rbest = (unsigned)e[2]; t.sbest = e[3]; t.careful = ex < 0;
    int n = e[4];
    RTP_CHECK(n >= 0 && 5 + n <= kParkInts && (unsigned)q < 0x7f000000u, 24, n, kParkInts - 5);
    for (int i = lane; i < n; i += 32) ws[i] = e[5 + i];
    __syncwarp();
    for (;;) {
      coop_trace<Job::kAny>(job.S, t, job.box_origin(q, t), ws, n);
      if (!job.rerun(t, stack)) break;
      if (lane == 0) ws[0] = 0;  // closest-hit: the winner's box failed - start over from the root, careful
      n = 1;
      __syncwarp();
    }
    job.commit(q, t, lane == 0);  // every lane holds the same ray: one of them retires it
    __syncwarp();
  }
}

// ---- streaming kernels ----------------------------------------------------------------------------
// Depth-0 rays of a caller-supplied batch (traceRay API): the q-th ray into the level-0 ray list.  Camera frames do not
// come through here: their primary rays are generated inside k_trace (FrameClosest<true>::init).
__global__ void __launch_bounds__(256) k_raygen(const FrameParams p) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.n_px) return;
  const bool valid = q < p.n_user_rays;
  const float *r = p.user_rays + 6 * (size_t)(valid ? q : 0);
  p.ray_o[0][q] = make_float4(r[0], r[1], r[2], i2f(q));
  p.ray_d[0][q] = make_float4(r[3], r[4], r[5], valid ? 1.f : 0.f);
  p.ksdeep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  p.primface[q] = valid ? -1 : -2;
}

// calculateColor for hit record h (flyscene.cpp:603-614): interpolated normal (flyscene.cpp:572-600), Phong over the
// unshadowed lights, clamped.  ks_out: the material's specular colour (the reflection weight of SURVEY.md Q5).
// shadowed(li): the outcome of shadow() for this hit and light li.
template <class Shadowed>
__device__ __forceinline__ v3 shade_hit_with(const FrameParams &p, Shadowed shadowed, v3 o, v3 P, int slot, v3 &normal, float4 &ks_out) {
  normal = interpolate_normal_slot(p.S, slot, P);
  const unsigned mat = (unsigned)f2i(__ldg(p.S.tris + 4 * (size_t)slot + 2).w) & kTriMatMask;
  const float4 m0 = __ldg(p.S.materials + 3 * mat), m1 = __ldg(p.S.materials + 3 * mat + 1),
               m2 = __ldg(p.S.materials + 3 * mat + 2);
  const int n_lights = p.S.n_lights;
  v3 sum = mk3(0.f, 0.f, 0.f);
  for (int li = 0; li < n_lights; ++li) {  // calculateColor, flyscene.cpp:606-612
    if (shadowed(li)) continue;  // shadowed: contributes nothing (:543-545)
    const float4 l0 = __ldg(p.S.lights + 2 * li), l1 = __ldg(p.S.lights + 2 * li + 1);
    v3 L = xyz(l0);
    if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));  // :607
    sum = vadd(sum, phong(xyz(m0), xyz(m1), xyz(m2), m0.w, normal, L, xyz(l1), o, P));
  }
  ks_out = m2;
  return mk3(clamp01(sum.x), clamp01(sum.y), clamp01(sum.z));
}
__device__ __forceinline__ v3 shade_hit(const FrameParams &p, int h, v3 o, v3 P, int slot, v3 &normal, float4 &ks_out) {
  const int n_lights = p.S.n_lights;
  return shade_hit_with(p, [&](int li) { return p.occl[(size_t)h * n_lights + li] != 0; }, o, P, slot, normal, ks_out);
}

// origin of path q at `depth` (camera frames: the eye at depth 0; otherwise the level's ray list)
__device__ __forceinline__ v3 path_origin(const FrameParams &p, int depth, int q) {
  if (depth == 0 && !p.user_rays) return mk3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
  return xyz(p.ray_o[depth & 1][q]);
}

// the hit records of camera rays (level 0 of a frame) are dense over the pixels, all others are compacted
__device__ __forceinline__ bool dense_level(const FrameParams &p, int depth) { return depth == 0 && !p.user_rays; }

// Per hit: normal, Phong over the unshadowed lights, clamp, remember ks, spawn the reflection ray.
__global__ void __launch_bounds__(256) k_accum(const FrameParams p, const int depth) {
  const int lane = threadIdx.x & 31;
  const bool dense = dense_level(p, depth);
  const int count = dense ? p.n_px : p.counters[kCtrHits + depth];
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (blockIdx.x * blockDim.x >= count) return;  // block-uniform
  const bool live = h < count && (!dense || p.hit_slot[h] >= 0);
  const bool spawn = live && depth + 1 < p.max_depth;
  v3 o2 = mk3(0.f, 0.f, 0.f), r = o2;
  int pixel = 0;
  if (live) {
    const float4 pb = p.pbuf[h], db = p.dbuf[h];
    const v3 P = xyz(pb), d = xyz(db);
    pixel = f2i(db.w);
    const v3 o = path_origin(p, depth, dense ? h : p.hitq[h]);
    float4 ks;
    v3 normal;
    const v3 c = shade_hit(p, h, o, P, f2i(pb.w), normal, ks);
    RTP_CHECK((unsigned)pixel < (unsigned)p.n_px && (unsigned)depth < (unsigned)p.max_depth, 33, pixel, p.n_px);
    p.direct[(size_t)depth * p.n_px + pixel] = make_float4(c.x, c.y, c.z, 0.f);
    p.ksdeep[pixel] = make_float4(ks.x, ks.y, ks.z, i2f(depth + 1));
    if (spawn) {
      r = reflect_dir(d, normal);            // :361
      o2 = vadd(P, vmuls(r, (float)0.001));  // :362
    }
  }
  const unsigned m = __ballot_sync(kFull, spawn);
  if (m) {
    const int leader = __ffs(m) - 1;
    int dst = 0;
    if (lane == leader) dst = atomicAdd(p.counters + kCtrPaths + depth + 1, __popc(m));
    dst = __shfl_sync(kFull, dst, leader) + __popc(m & ((1u << lane) - 1u));
    if (spawn) {
      RTP_CHECK((unsigned)dst < (unsigned)p.n_px && (unsigned)pixel < (unsigned)p.n_px, 30, dst, p.n_px);
      p.ray_o[(depth + 1) & 1][dst] = make_float4(o2.x, o2.y, o2.z, i2f(pixel));
      p.ray_d[(depth + 1) & 1][dst] = make_float4(r.x, r.y, r.z, 1.f);
    }
  }
}

// ---- bounce levels >= 1 in ONE persistent kernel ---------------------------------------------------------------------
// The wavefront above pays per level: three dependent launches, each as long as its longest ray however few rays the
// level holds (C4, 10 M triangles: levels 1-5 hold a third of the rays and took two thirds of the frame; on a 1/8 share of
// the frame every level cost 0.24 ms of pure latency).  Below level 0 the rays are incoherent anyway - there is no
// screen-space order left to protect - so k_paths drops the level structure: a lane takes a PATH (a reflection ray that
// k_accum spawned at level 0) and carries it to its end - closest hit (calculateMinimumFace), hit point and
// interpolated normal, one shadow ray per light (shadow), Phong (calcSingleColor), the reflection ray
// (flyscene.cpp:317-371), level after level.  A warp's lanes mix closest-hit and shadow rays (trav_step<2>: the kind is a
// per-ray flag); the state a path carries between traversals lives in shared memory so that the node loop keeps its
// registers; lanes that have finished a traversal wait until a few of them can go through the (divergent, heavy)
// shading code together; idle lanes refill from the level-1 ray list.  Same functions, same arithmetic, same results
// as the wavefront kernels (RTP_DEEP=wavefront selects those; the parity suite runs under both).
enum { PS_OX, PS_OY, PS_OZ, PS_DX, PS_DY, PS_DZ, PS_PX, PS_PY, PS_PZ, PS_NX, PS_NY, PS_NZ, PS_SX, PS_SY, PS_SZ, PS_FLOATS };
enum { PS_PIXEL, PS_DEPTH, PS_LIGHT, PS_SLOT, PS_INTS };
int g_path_wait = 6;  // lanes that must be waiting for the shading block ... (RTP_PATH_WAIT)
int g_path_run = 12;  // ... unless fewer than this many lanes still traverse (RTP_PATH_RUN)

template <int kRefillBelow>
__global__ void __launch_bounds__(128, RTP_TRACE_MINB) k_paths(const FrameParams p, const int first_depth, const int kPathWait,
                                                               const int kPathRun) {
  __shared__ float s_f[PS_FLOATS][128];
  __shared__ int s_i[PS_INTS][128];
  __shared__ int s_cnt[2 * kMaxDepth];  // [d] hits at depth d, [kMaxDepth + d] rays spawned into depth d
  const int tid = threadIdx.x, lane = tid & 31;
  const unsigned lt = (1u << lane) - 1u;
  if (tid < 2 * kMaxDepth) s_cnt[tid] = 0;
  __syncthreads();
#define PF(k) s_f[k][tid]
#define PI(k) s_i[k][tid]
  const DeviceScene &S = p.S;
  const int n_lights = S.n_lights;
  const int count = p.counters[kCtrPaths + first_depth];
  int *const cursor = p.counters + kCtrFetchClosest + first_depth;
  const float4 *ro = p.ray_o[first_depth & 1], *rd = p.ray_d[first_depth & 1];
  int stack[kStackSize];
  Trav t;
  t.node = kSentinel;
  t.any = false;
  bool have = false, exhausted = false;
  for (;;) {
    unsigned running = __ballot_sync(kFull, have && t.node != kSentinel);
    const unsigned waiting = __ballot_sync(kFull, have && t.node == kSentinel);
    if (waiting && (__popc(waiting) >= kPathWait || __popc(running) < kPathRun)) {
      if (have && t.node == kSentinel) {
        bool next_light = false;
        int li = 0;
        v3 P, sum;
        if (!t.any) {  // a closest-hit traversal has ended
          if (closest_needs_rerun(S, t, stack)) {
            // careful re-run from the root: keeps traversing
          } else if (t.sbest < 0) {
            have = false;  // the path leaves the scene: nothing below contributes (flyscene.cpp:327-332)
          } else {
            P = vadd(t.o, vmuls(t.d, t.tbest));  // flyscene.cpp:464
            const v3 N = interpolate_normal_slot(S, t.sbest, P);
            PF(PS_PX) = P.x; PF(PS_PY) = P.y; PF(PS_PZ) = P.z;
            PF(PS_NX) = N.x; PF(PS_NY) = N.y; PF(PS_NZ) = N.z;
            PI(PS_SLOT) = t.sbest;
            sum = mk3(0.f, 0.f, 0.f);
            atomicAdd(&s_cnt[PI(PS_DEPTH)], 1);
            next_light = true;
          }
        } else {  // the shadow ray towards light PS_LIGHT has ended
          li = PI(PS_LIGHT);
          P = mk3(PF(PS_PX), PF(PS_PY), PF(PS_PZ));
          sum = mk3(PF(PS_SX), PF(PS_SY), PF(PS_SZ));
          if (t.sbest < 0) {  // not shadowed: calcSingleColor (flyscene.cpp:546-565)
            const unsigned mat = (unsigned)f2i(__ldg(S.tris + 4 * (size_t)PI(PS_SLOT) + 2).w) & kTriMatMask;
            const float4 m0 = __ldg(S.materials + 3 * mat), m1 = __ldg(S.materials + 3 * mat + 1), m2 = __ldg(S.materials + 3 * mat + 2);
            const float4 l1 = __ldg(S.lights + 2 * li + 1);
            sum = vadd(sum, phong(xyz(m0), xyz(m1), xyz(m2), m0.w, mk3(PF(PS_NX), PF(PS_NY), PF(PS_NZ)), t.d, xyz(l1),
                                  mk3(PF(PS_OX), PF(PS_OY), PF(PS_OZ)), P));
          }
          ++li;
          next_light = true;
        }
        if (next_light) {
          if (li < n_lights) {  // shadow(P, L) for the next light (flyscene.cpp:510-526,606-612)
            const float4 l0 = __ldg(S.lights + 2 * li);
            v3 L = xyz(l0);
            if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));  // flyscene.cpp:607
            PI(PS_LIGHT) = li;
            PF(PS_SX) = sum.x; PF(PS_SY) = sum.y; PF(PS_SZ) = sum.z;
            trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);  // flyscene.cpp:512
            t.any = true;
          } else {  // the level is done: its clamped direct colour, ks of the (so far) deepest hit, the reflection ray
            const int depth = PI(PS_DEPTH), pixel = PI(PS_PIXEL);
            const unsigned mat = (unsigned)f2i(__ldg(S.tris + 4 * (size_t)PI(PS_SLOT) + 2).w) & kTriMatMask;
            const float4 m2 = __ldg(S.materials + 3 * mat + 2);
            RTP_CHECK((unsigned)pixel < (unsigned)p.n_px && (unsigned)depth < (unsigned)p.max_depth, 31, pixel, p.n_px);
            p.direct[(size_t)depth * p.n_px + pixel] = make_float4(clamp01(sum.x), clamp01(sum.y), clamp01(sum.z), 0.f);
            p.ksdeep[pixel] = make_float4(m2.x, m2.y, m2.z, i2f(depth + 1));
            if (depth + 1 < p.max_depth) {
              const v3 r = reflect_dir(mk3(PF(PS_DX), PF(PS_DY), PF(PS_DZ)), mk3(PF(PS_NX), PF(PS_NY), PF(PS_NZ)));  // :361
              const v3 o2 = vadd(P, vmuls(r, (float)0.001));                                                          // :362
              PF(PS_OX) = o2.x; PF(PS_OY) = o2.y; PF(PS_OZ) = o2.z;
              PF(PS_DX) = r.x; PF(PS_DY) = r.y; PF(PS_DZ) = r.z;
              PI(PS_DEPTH) = depth + 1;
              atomicAdd(&s_cnt[kMaxDepth + depth + 1], 1);
              trav_init(t, o2, r, stack);
            } else {
              have = false;
            }
          }
        }
      }
      running = __ballot_sync(kFull, have && t.node != kSentinel);
    }
    const unsigned act = __ballot_sync(kFull, have);
    if (!exhausted && __popc(act) < kRefillBelow) {
      const unsigned need = ~act;
      const int n_need = __popc(need);
      const int leader = __ffs(need) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(cursor, n_need);
      base = __shfl_sync(kFull, base, leader);
      exhausted = base + n_need >= count;
      if (!have) {
        const int q = base + __popc(need & lt);
        if (q < count) {
          const float4 o4 = ro[q], d4 = rd[q];
          PF(PS_OX) = o4.x; PF(PS_OY) = o4.y; PF(PS_OZ) = o4.z;
          PF(PS_DX) = d4.x; PF(PS_DY) = d4.y; PF(PS_DZ) = d4.z;
          PI(PS_PIXEL) = f2i(o4.w);
          PI(PS_DEPTH) = first_depth;
          trav_init(t, xyz(o4), xyz(d4), stack);
          have = true;
        }
      }
      running = __ballot_sync(kFull, have && t.node != kSentinel);
    }
    if (running == 0) {
      if (__ballot_sync(kFull, have) == 0 && exhausted) break;
      continue;
    }
    if (have && t.node != kSentinel)
      // origin of the reference's BoundingBox test: the un-offset hit point for a shadow ray, the ray's own origin
      // for a (careful) closest-hit ray
      trav_step<2>(S, t, stack, [&] { return t.any ? mk3(PF(PS_PX), PF(PS_PY), PF(PS_PZ)) : t.o; });
  }
#undef PF
#undef PI
  __syncthreads();
  if (tid < 2 * kMaxDepth && s_cnt[tid]) atomicAdd(p.counters + (tid < kMaxDepth ? kCtrHits + tid : kCtrPaths + tid - kMaxDepth), s_cnt[tid]);
}

// the fold of traceRay's return values (flyscene.cpp:366-370 with the member ks of the deepest hit)
__device__ __forceinline__ float4 resolve_pixel(const FrameParams &p, int q) {
  const float4 kd = p.ksdeep[q];
  const int k = f2i(kd.w);
  const int face = p.primface[q];
  if (face == -2) return make_float4(0.f, 0.f, 0.f, i2f(-2));  // pixel outside the image
  if (k == 0) return make_float4(p.bg[0], p.bg[1], p.bg[2], i2f(-1));  // flyscene.cpp:327-330
  // deepest level: clamp(direct + 0 * ks) (the recursion below it returned (0,0,0))
  float4 dc = p.direct[(size_t)(k - 1) * p.n_px + q];
  v3 c = mk3(clamp01(fadd(dc.x, fmul(0.f, kd.x))), clamp01(fadd(dc.y, fmul(0.f, kd.y))),
             clamp01(fadd(dc.z, fmul(0.f, kd.z))));
  for (int d = k - 2; d >= 0; --d) {
    dc = p.direct[(size_t)d * p.n_px + q];
    c = mk3(clamp01(fadd(dc.x, fmul(c.x, kd.x))), clamp01(fadd(dc.y, fmul(c.y, kd.y))),
            clamp01(fadd(dc.z, fmul(c.z, kd.z))));
  }
  return make_float4(c.x, c.y, c.z, i2f(face));
}

// ppmIO.hpp:150: min(255, (int)(255*c)); NaN -> 0 (undefined in the reference)
__device__ __forceinline__ unsigned char quantise(float c) {
  const float s = fmul(255.f, c);
  const int qv = (s != s) ? 0 : __float2int_rz(s);
  return (unsigned char)min(255, max(qv, 0));
}

// max_depth 1: there is one level, so the shading of its hits (k_accum) and the fold (clamp(direct + 0 * ks), the
// deepest-level case of resolve_pixel) are done here, per pixel, without the round trip through direct[] / ksdeep[].
__device__ __forceinline__ float4 shade_pixel_depth1(const FrameParams &p, int q) {
  const int face = p.primface[q];
  if (face == -2) return make_float4(0.f, 0.f, 0.f, i2f(-2));  // pixel outside the image / no ray
  const int h = p.hit_slot[q];
  if (h < 0) return make_float4(p.bg[0], p.bg[1], p.bg[2], i2f(-1));  // flyscene.cpp:327-330
  const float4 pb = p.pbuf[h];
  float4 ks;
  v3 normal;
  const v3 dc = shade_hit(p, h, path_origin(p, 0, q), xyz(pb), f2i(pb.w), normal, ks);
  return make_float4(clamp01(fadd(dc.x, fmul(0.f, ks.x))), clamp01(fadd(dc.y, fmul(0.f, ks.y))),
                     clamp01(fadd(dc.z, fmul(0.f, ks.z))), i2f(face));
}

template <bool SHADE1>
__global__ void __launch_bounds__(256) k_resolve(const FrameParams p) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.n_px) return;
  float4 out = SHADE1 ? shade_pixel_depth1(p, q) : resolve_pixel(p, q);
  if (f2i(out.w) == -2) out.w = i2f(-1);
  p.fb[q] = out;
}

// Resolve fused with the image assembly: thread r of a tile handles pixel (r & 31, r >> 5) of that
// tile, so a warp writes 32 consecutive pixels of one image row (96 contiguous bytes) - the image may
// be peer memory on GPU 0, reached over NVLink, and then this kernel IS the framebuffer gather.
// SHADE1 (max_depth 1): k_shade_image - the level's shading is done here too (shade_pixel_depth1).
template <bool SHADE1>
__global__ void __launch_bounds__(256) k_resolve_image(const FrameParams p) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p.n_px) return;  // n_px is a multiple of 1024: whole warps leave together
  const int lane = threadIdx.x & 31;
  const int local = r >> 10, x = r & 31, y = (r >> 5) & 31;
  const int tile = __ldg(p.tile_ids + local);
  const int i = (tile % p.tiles_x) * kTileDim + x, j = (tile / p.tiles_x) * kTileDim + y;
  const bool valid = i < p.width && j < p.height;
  const int q = (local << 10) + ((((y >> 2) * 4 + (x >> 3)) << 5) + ((y & 3) << 3) + (x & 7));
  float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) out = SHADE1 ? shade_pixel_depth1(p, q) : resolve_pixel(p, q);
  const size_t idx = (size_t)j * p.width + i;
  RTP_CHECK(!valid || (idx < (size_t)p.width * p.height && q < p.n_px), 32, idx, p.width * p.height);
  if (valid && p.fb) p.fb[q] = out;
  if (valid && p.img_rgb) {  // float plane, row-major: a warp stores 384 consecutive bytes
    p.img_rgb[3 * idx] = out.x;
    p.img_rgb[3 * idx + 1] = out.y;
    p.img_rgb[3 * idx + 2] = out.z;
  }
  if (p.img_rgb8) {
    // the warp's 32 pixels are 96 consecutive bytes of one image row: lanes 0-23 store them as 24 aligned words
    // (one full-width transaction towards HBM, a peer GPU over NVLink or page-locked host memory over PCIe)
    // instead of 96 single bytes; rows whose start is not word-aligned and the ragged right edge store bytes
    const unsigned packed = (unsigned)quantise(out.x) | ((unsigned)quantise(out.y) << 8) | ((unsigned)quantise(out.z) << 16);
    unsigned char *seg = p.img_rgb8 + 3 * (idx - (size_t)lane);  // byte address of the warp's first pixel
    const bool whole = __all_sync(kFull, valid) && (((size_t)seg & 3) == 0);
    const int b0 = 4 * (lane < 24 ? lane : 0);  // first byte of this lane's word
    const int p0 = b0 / 3, c0 = b0 - 3 * p0;
    const unsigned lo = __shfl_sync(kFull, packed, p0), hi = __shfl_sync(kFull, packed, (p0 + 1) & 31);
    if (whole) {
      if (lane < 24) {
        const unsigned long long both = (unsigned long long)lo | ((unsigned long long)hi << 24);
        reinterpret_cast<unsigned *>(seg)[lane] = (unsigned)(both >> (8 * c0));
      }
    } else if (valid) {
      p.img_rgb8[3 * idx] = (unsigned char)(packed & 0xff);
      p.img_rgb8[3 * idx + 1] = (unsigned char)((packed >> 8) & 0xff);
      p.img_rgb8[3 * idx + 2] = (unsigned char)(packed >> 16);
    }
  }
  if (valid && p.img_hit) p.img_hit[idx] = f2i(out.w);
}

// ---- the camera level of a frame in ONE kernel -----------------------------------------------------------------------
// k_trace<FrameClosest> -> (k_coop) -> k_trace<FrameAny> -> (k_coop) -> k_accum / k_resolve_image are dependent launches,
// each with a latency-bound tail (its last batches of dependent node fetches leave most SMs idle), and the pixels' hit
// records travel through HBM between them (hit_slot, pbuf, dbuf, occl: ~100 B per pixel).  k_camera runs the whole
// camera level of traceRay (flyscene.cpp:299-371 at depth 0) per warp batch instead: a warp takes an 8x4-pixel footprint,
// generates the 32 primary rays, finds their closest hits (trav_step<0>), then sends the 32 shadow rays towards light 0,
// light 1, ... (trav_step<1>), shades its pixels and - max_depth 1 - stores them straight into the image (which may be
// GPU 0's, over NVLink); with max_depth > 1 it leaves the level-0 records the fold needs (direct, ksdeep, primface) and
// appends the reflection rays to level 1's ray list.  The phases are batch-synchronous inside the warp, as in k_trace's
// batch mode - neighbouring pixels stay together, the rays of a phase are coherent - but different warps are in
// different phases, so the only tail left is the frame's.  The hit record of a pixel (P, triangle slot, occlusion bits)
// never leaves the SM: shared memory between the phases.  Same functions and arithmetic as the per-phase kernels
// (RTP_FUSED=0 selects those; the parity suite runs under both).
// Stragglers are parked as in k_trace; k_coop_camera finishes a parked PIXEL with a whole warp: the rest of its primary
// ray or of its shadow ray towards light li (coop_trace from the saved stack), the remaining lights (coop_trace from the
// root), shading, output.  Parked record (kParkInts ints): [0] (pixel + 1) | li << 24 | kCamAnyFlag | careful << 31,
// primary ray: [1..3] best t / rank / slot so far; shadow ray: [1..3] the hit point P, [70] its triangle slot, [71] the
// occlusion bits of lights < li; [4] n, [5..] the pending stack.
constexpr int kCamAnyFlag = 1 << 29;
constexpr int kCamMaxLights = 32;         // occlusion bits of a pixel live in one register
constexpr int kCamMaxPixels = 1 << 24;    // pixel id field of a parked record
struct CamPark {
  int *slots, *ctr;
  int cap;
  int budget, lanes;          // primary rays: steps a warp spends on a batch before the rays still running are parked
  int budget_any, lanes_any;  // shadow rays
};

// pixel q of this rank's tile-major pixel list -> image coordinates; false: tile padding outside the image
__device__ __forceinline__ bool cam_pixel(const FrameParams &p, int q, int &i, int &j) {
  const int tile = __ldg(p.tile_ids + (q >> 10));
  int x, y;
  tile_pixel(q & (kTilePixels - 1), x, y);
  i = (tile % p.tiles_x) * kTileDim + x;
  j = (tile / p.tiles_x) * kTileDim + y;
  return i < p.width && j < p.height;
}

__device__ __forceinline__ v3 cam_light_dir(const DeviceScene &S, int li, v3 P) {
  const float4 l0 = __ldg(S.lights + 2 * li);
  v3 L = xyz(l0);
  if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));  // flyscene.cpp:607
  return L;
}

__device__ __forceinline__ bool cam_park(const CamPark &pk, int head, int a1, int a2, int a3, const Trav &t, const int *stack,
                                         int x0, int x1) {
  const int at = atomicAdd(pk.ctr, 1);
  if (at >= pk.cap) return false;  // list full: the ray simply carries on
  int *e = pk.slots + (size_t)at * kParkInts;
  e[0] = head; e[1] = a1; e[2] = a2; e[3] = a3;
  int n = 0;
  for (int i = 1; i <= t.sp; ++i) e[5 + n++] = RTP_STK(stack, i);
  if (t.node != kSentinel) e[5 + n++] = t.node;
  RTP_CHECK(5 + n <= kParkInts - 2, 25, 5 + n, kParkInts - 2);
  e[4] = n;
  e[kParkInts - 2] = x0; e[kParkInts - 1] = x1;
  return true;
}

// What is left of a camera pixel once its rays are traced: calculateColor over the unshadowed lights (occ: bit li =
// shadow(P, light li)), then - DEPTH1 - the fold of the only level (clamp(direct + 0 * ks), as shade_pixel_depth1) and the
// outputs of k_resolve / k_resolve_image; otherwise the level-0 records of the fold and the reflection ray (o2, r) of
// flyscene.cpp:355-363: returns true when that ray has to be appended to level 1's list (the caller compacts).
template <bool DEPTH1>
__device__ __forceinline__ bool camera_finish(const FrameParams &p, int q, int i, int j, bool valid, bool hit, v3 P, int slot,
                                              unsigned occ, v3 &o2, v3 &r) {
  const int face = !valid ? -2 : hit ? __ldg(p.S.face_of + slot) : -1;
  float4 ks = make_float4(0.f, 0.f, 0.f, 0.f);
  v3 c = mk3(0.f, 0.f, 0.f), normal = c;
  if (hit)
    c = shade_hit_with(p, [&](int li) { return ((occ >> li) & 1u) != 0u; }, mk3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]), P, slot,
                       normal, ks);
  if (DEPTH1) {
    float4 out = make_float4(0.f, 0.f, 0.f, i2f(-1));                    // pixel outside the image
    if (valid && !hit) out = make_float4(p.bg[0], p.bg[1], p.bg[2], i2f(-1));  // flyscene.cpp:327-330
    if (hit) out = make_float4(clamp01(fadd(c.x, fmul(0.f, ks.x))), clamp01(fadd(c.y, fmul(0.f, ks.y))),
                               clamp01(fadd(c.z, fmul(0.f, ks.z))), i2f(face));
    if (p.fb) p.fb[q] = out;
    if (valid) {
      const size_t idx = (size_t)j * p.width + i;
      RTP_CHECK(idx < (size_t)p.width * p.height, 34, idx, p.width * p.height);
      if (p.img_rgb) { p.img_rgb[3 * idx] = out.x; p.img_rgb[3 * idx + 1] = out.y; p.img_rgb[3 * idx + 2] = out.z; }
      if (p.img_rgb8) {  // the 8 lanes of a footprint row store 24 consecutive bytes
        p.img_rgb8[3 * idx] = quantise(out.x); p.img_rgb8[3 * idx + 1] = quantise(out.y); p.img_rgb8[3 * idx + 2] = quantise(out.z);
      }
      if (p.img_hit) p.img_hit[idx] = f2i(out.w);
    }
    return false;
  }
  p.primface[q] = face;
  if (!hit) {
    p.ksdeep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    return false;
  }
  p.direct[q] = make_float4(c.x, c.y, c.z, 0.f);  // level 0
  p.ksdeep[q] = make_float4(ks.x, ks.y, ks.z, i2f(1));
  r = reflect_dir(primary_dir(p.cam, i, j), normal);  // :361
  o2 = vadd(P, vmuls(r, (float)0.001));               // :362
  return true;
}

// EXPERIMENT (RTP_ANY_SHADE=1, off by default: parity-green, no gain - the tile groups already hide the shading kernel).
// The camera level's shadow launch with the level's shading folded in (big frames, where the three phases stay separate
// launches): FrameAny<DENSE> whose warps, once they have walked all lights of their 32 pixels, shade those pixels from
// the occlusion bits in their registers - k_accum's work at level 0 (max_depth > 1) or the whole of k_resolve_image
// (max_depth 1: shading, fold, quantisation, image store).  No occl[] round trip, one launch and one tail less per tile
// group.  Only for launches that do not park shadow rays (a parked ray's outcome arrives after the warp has moved on)
// and at most 32 lights; otherwise FrameAny + k_accum / k_resolve_image as before.
template <bool DEPTH1>
struct FrameAnyShade : FrameAny<true> {
  static constexpr bool kFinish = true;
  FrameParams p;
  __device__ __forceinline__ void commit(int, const Trav &, bool) const {}  // the outcome stays in the lane (k_trace: occ)
  __device__ __forceinline__ void finish(int q, unsigned occ) const {
    const int lane = threadIdx.x & 31;
    int i, j;
    const bool valid = cam_pixel(p, q, i, j);
    const bool hit = this->hit_slot[q] >= 0;
    const float4 pb = hit ? this->pbuf[q] : make_float4(0.f, 0.f, 0.f, i2f(-1));
    v3 o2 = mk3(0.f, 0.f, 0.f), r = o2;
    const bool spawn = camera_finish<DEPTH1>(p, q, i, j, valid, hit, xyz(pb), f2i(pb.w), occ, o2, r);
    if (!DEPTH1) {
      const unsigned m = __ballot_sync(kFull, spawn);
      if (m) {
        const int leader = __ffs(m) - 1;
        int dst = 0;
        if (lane == leader) dst = atomicAdd(p.counters + kCtrPaths + 1, __popc(m));
        dst = __shfl_sync(kFull, dst, leader) + __popc(m & ((1u << lane) - 1u));
        if (spawn) {
          RTP_CHECK((unsigned)dst < (unsigned)p.n_px, 30, dst, p.n_px);
          p.ray_o[1][dst] = make_float4(o2.x, o2.y, o2.z, i2f(q));
          p.ray_d[1][dst] = make_float4(r.x, r.y, r.z, 1.f);
        }
      }
    }
  }
};

template <bool DEPTH1>
__global__ void __launch_bounds__(128, RTP_TRACE_MINB) k_camera(const FrameParams p, const CamPark pk) {
  __shared__ float s_P[3][128];
  __shared__ int s_slot[128];
  const int tid = threadIdx.x, lane = tid & 31;
  const DeviceScene &S = p.S;
  const int n_lights = S.n_lights;
  int *const cursor = p.counters + kCtrFetchClosest;
  int stack[kStackSize];
  for (;;) {
    int base = 0;
    if (lane == 0) base = atomicAdd(cursor, 32);
    base = __shfl_sync(kFull, base, 0);
    if (base >= p.n_px) break;  // n_px is a multiple of 1024: whole batches
    const int q = base + lane;
    Trav t;
    t.node = kSentinel;
    t.sbest = -1;
    bool have;
    {
      int i, j;
      have = cam_pixel(p, q, i, j);
      if (have) trav_init(t, mk3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]), primary_dir(p.cam, i, j), stack);
    }
    // ---- phase A: calculateMinimumFace for the 32 primary rays (flyscene.cpp:373-396)
    bool parked = false;
    for (int age = 0;;) {
      const unsigned act = __ballot_sync(kFull, have && t.node != kSentinel);
      if (act == 0) {  // all traversals have ended: the deferred BoundingBox check of the winners, together
        const bool again = have && closest_needs_rerun(S, t, stack);  // careful re-run from the root
        if (!__any_sync(kFull, again)) break;
        continue;
      }
      if (++age > pk.budget && __popc(act) <= pk.lanes) {
        if (have && t.node != kSentinel &&
            cam_park(pk, (q + 1) | (t.careful ? (int)0x80000000 : 0), f2i(t.tbest), (int)t.rbest, t.sbest, t, stack, 0, 0)) {
          have = false;
          parked = true;
        }
        age = 0;
        continue;
      }
      if (have && t.node != kSentinel) trav_step<0>(S, t, stack, [&] { return t.o; });
    }
    const bool hit = have && t.sbest >= 0;
    {
      const unsigned m = __ballot_sync(kFull, hit);
      if (m && lane == __ffs(m) - 1) atomicAdd(p.counters + kCtrHits, __popc(m));
    }
    if (hit) {
      RTP_CHECK((unsigned)t.sbest < (unsigned)S.n_tris, 12, t.sbest, S.n_tris);
      const v3 P = vadd(t.o, vmuls(t.d, t.tbest));  // flyscene.cpp:464
      s_P[0][tid] = P.x; s_P[1][tid] = P.y; s_P[2][tid] = P.z;
      s_slot[tid] = t.sbest;
    }
    // ---- phase B: shadow(P, light) for the batch, light after light (flyscene.cpp:510-526,606-612)
    unsigned occ = 0;
    for (int li = 0; li < n_lights; ++li) {
      bool run = hit && !parked;
      if (!__any_sync(kFull, run)) break;
      t.node = kSentinel;
      if (run) {
        const v3 P = mk3(s_P[0][tid], s_P[1][tid], s_P[2][tid]);
        const v3 L = cam_light_dir(S, li, P);
        trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);  // flyscene.cpp:512
      }
      for (int age = 0;;) {
        const unsigned act = __ballot_sync(kFull, run && t.node != kSentinel);
        if (act == 0) break;
        if (++age > pk.budget_any && __popc(act) <= pk.lanes_any) {
          if (run && t.node != kSentinel &&
              cam_park(pk, (q + 1) | (li << 24) | kCamAnyFlag, f2i(s_P[0][tid]), f2i(s_P[1][tid]), f2i(s_P[2][tid]), t, stack,
                       s_slot[tid], (int)occ)) {
            run = false;
            parked = true;
          }
          age = 0;
          continue;
        }
        if (run && t.node != kSentinel)
          trav_step<1>(S, t, stack, [&] { return mk3(s_P[0][tid], s_P[1][tid], s_P[2][tid]); });
      }
      if (run && t.sbest >= 0) occ |= 1u << li;
    }
    // ---- phase C: shading and output (parked pixels: k_coop_camera)
    v3 o2 = mk3(0.f, 0.f, 0.f), r = o2;
    bool spawn = false;
    if (!parked) {
      int i, j;
      const bool valid = cam_pixel(p, q, i, j);
      const v3 P = hit ? mk3(s_P[0][tid], s_P[1][tid], s_P[2][tid]) : o2;
      spawn = camera_finish<DEPTH1>(p, q, i, j, valid, hit, P, hit ? s_slot[tid] : -1, occ, o2, r);
    }
    if (!DEPTH1) {
      const unsigned m = __ballot_sync(kFull, spawn);
      if (m) {
        const int leader = __ffs(m) - 1;
        int dst = 0;
        if (lane == leader) dst = atomicAdd(p.counters + kCtrPaths + 1, __popc(m));
        dst = __shfl_sync(kFull, dst, leader) + __popc(m & ((1u << lane) - 1u));
        if (spawn) {
          RTP_CHECK((unsigned)dst < (unsigned)p.n_px, 30, dst, p.n_px);
          p.ray_o[1][dst] = make_float4(o2.x, o2.y, o2.z, i2f(q));
          p.ray_d[1][dst] = make_float4(r.x, r.y, r.z, 1.f);
        }
      }
    }
  }
}

// Finishes the pixels k_camera parked, one warp per pixel.
template <bool DEPTH1>
__global__ void __launch_bounds__(128, 4) k_coop_camera(const FrameParams p, const CamPark pk) {
  __shared__ int s_ws[4][kCoopStack];
  int *const ws = s_ws[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const DeviceScene &S = p.S;
  const int n_lights = S.n_lights;
  const int n_parked = min(*pk.ctr, pk.cap);
  int stack[1];  // trav_init / closest_needs_rerun write stack[0] only
  for (;;) {
    int k = 0;
    if (lane == 0) k = atomicAdd(pk.ctr + 1, 1);
    k = __shfl_sync(kFull, k, 0);
    if (k >= n_parked) break;
    const int *e = pk.slots + (size_t)k * kParkInts;
    const int head = e[0];
    const int q = (head & (kCamMaxPixels - 1)) - 1;
    RTP_CHECK((unsigned)q < (unsigned)p.n_px, 26, q, p.n_px);
    int i, j;
    cam_pixel(p, q, i, j);  // a parked pixel is inside the image
    int n = e[4];
    RTP_CHECK(n >= 0 && 5 + n <= kParkInts - 2, 24, n, kParkInts - 7);
    for (int x = lane; x < n; x += 32) ws[x] = e[5 + x];
    __syncwarp();
    Trav t;
    v3 P = mk3(0.f, 0.f, 0.f);
    int slot = -1, li = 0;
    unsigned occ = 0;
    bool hit;
    if (!(head & kCamAnyFlag)) {  // the pixel's primary ray: the same ray in every lane
      trav_init(t, mk3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]), primary_dir(p.cam, i, j), stack);
      t.tbest = i2f(e[1]); t.rbest = (unsigned)e[2]; t.sbest = e[3]; t.careful = head < 0;
      for (;;) {
        coop_trace<false>(S, t, t.o, ws, n);
        if (!closest_needs_rerun(S, t, stack)) break;
        if (lane == 0) ws[0] = 0;  // the winner's box failed: start over from the root, careful
        n = 1;
        __syncwarp();
      }
      hit = t.sbest >= 0;
      if (hit) {
        P = vadd(t.o, vmuls(t.d, t.tbest));  // flyscene.cpp:464
        slot = t.sbest;
        if (lane == 0) atomicAdd(p.counters + kCtrHits, 1);
      }
    } else {  // its shadow ray towards light li: hit point, slot and the bits of the lights before it come with the record
      hit = true;
      P = mk3(i2f(e[1]), i2f(e[2]), i2f(e[3]));
      slot = e[kParkInts - 2];
      occ = (unsigned)e[kParkInts - 1];
      li = (head >> 24) & (kCamMaxLights - 1);
      const v3 L = cam_light_dir(S, li, P);
      trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);
      coop_trace<true>(S, t, P, ws, n);
      if (t.sbest >= 0) occ |= 1u << li;
      ++li;
    }
    if (hit)
      for (; li < n_lights; ++li) {  // the lights that are left, each from the root
        const v3 L = cam_light_dir(S, li, P);
        trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);
        __syncwarp();
        if (lane == 0) ws[0] = 0;
        __syncwarp();
        coop_trace<true>(S, t, P, ws, 1);
        if (t.sbest >= 0) occ |= 1u << li;
      }
    if (lane == 0) {
      v3 o2, r;
      if (camera_finish<DEPTH1>(p, q, i, j, true, hit, P, slot, occ, o2, r)) {
        const int dst = atomicAdd(p.counters + kCtrPaths + 1, 1);
        RTP_CHECK((unsigned)dst < (unsigned)p.n_px, 30, dst, p.n_px);
        p.ray_o[1][dst] = make_float4(o2.x, o2.y, o2.z, i2f(q));
        p.ray_d[1][dst] = make_float4(r.x, r.y, r.z, 1.f);
      }
    }
    __syncwarp();
  }
}

// k_paths with HELPERS: parallelism inside a path.  EXPERIMENT, not the default: parity-green (bit-equal to k_paths and to
// the per-level kernels) and slower on everything measured (C2 levels >= 1: 0.393 -> 0.432 ms, a C4 1/8 share 1.57 -> 1.66,
// the whole C4 frame through it 5.15 -> 6.45): the claim / report traffic of every scheduling round and the scattered
// helper rays cost more than the idle lanes were worth, profiles/r02_camera.md section 6.  k_paths walks a hit's shadow rays one after the other in the lane that
// owns the path - a chain of (1 + lights) traversals per level, of cold node fetches on a big scene - and a small share of
// a frame (a rank of 8, a small image) is as long as its longest chain while most lanes of the warp sit idle.  Here a lane
// without a path (the ray list is exhausted, or it is above the refill threshold) takes one of the shadow rays that an
// owner of its warp has not started yet: the owner publishes the hit (P in shared memory) and hands out light indices
// through a per-owner counter; whoever traces light li reports the outcome as bit li of the owner's occlusion word; when
// all lights have reported, the owner adds up calcSingleColor over the unshadowed lights IN LIGHT ORDER (the sum the
// sequential walk forms, bit for bit) and goes on to the reflection ray.  With every lane busy it degenerates to the
// sequential walk.  At most 32 lights (one occlusion word); frames with more take the per-level kernels.
template <int kRefillBelow>
__global__ void __launch_bounds__(128, RTP_TRACE_MINB) k_paths_help(const FrameParams p, const int first_depth, const int kPathWait,
                                                                    const int kPathRun) {
  __shared__ float s_f[PS_SX][128];  // O, D, P, N of the lane's path (no running sum here)
  __shared__ int s_i[PS_INTS][128];
  __shared__ int s_next[128], s_done[128];  // per owner: next light to hand out, lights that have reported
  __shared__ unsigned s_occ[128];           //            bit li: light li is shadowed
  __shared__ int s_cnt[2 * kMaxDepth];
  const int tid = threadIdx.x, lane = tid & 31;
  const unsigned lt = (1u << lane) - 1u;
  if (tid < 2 * kMaxDepth) s_cnt[tid] = 0;
  s_next[tid] = 0x40000000;  // no offer
  __syncthreads();
#define PF(k) s_f[k][tid]
#define PI(k) s_i[k][tid]
  const DeviceScene &S = p.S;
  const int n_lights = S.n_lights;
  const int count = p.counters[kCtrPaths + first_depth];
  int *const cursor = p.counters + kCtrFetchClosest + first_depth;
  const float4 *ro = p.ray_o[first_depth & 1], *rd = p.ray_d[first_depth & 1];
  int stack[kStackSize];
  Trav t;
  t.node = kSentinel;
  t.any = false;
  bool have = false, exhausted = false, helping = false;
  int h_owner = 0, h_light = 0;  // helping: thread (in the CTA) that owns the path, light this lane traces for it
  for (;;) {
    if (helping && t.node == kSentinel) {  // a helper reports as soon as its ray has ended
      if (t.sbest >= 0) atomicOr(&s_occ[h_owner], 1u << h_light);
      __threadfence_block();
      atomicAdd(&s_done[h_owner], 1);
      helping = false;
    }
    __syncwarp();
    unsigned running = __ballot_sync(kFull, (have || helping) && t.node != kSentinel);
    const unsigned waiting = __ballot_sync(kFull, have && t.node == kSentinel);
    if (waiting && (__popc(waiting) >= kPathWait || __popc(running) < kPathRun)) {
      if (have && t.node == kSentinel) {
        bool start_light = false, level_ready = false;
        int li = 0;
        if (!t.any) {  // a closest-hit traversal has ended
          if (closest_needs_rerun(S, t, stack)) {
            // careful re-run from the root: keeps traversing
          } else if (t.sbest < 0) {
            have = false;  // the path leaves the scene (flyscene.cpp:327-332)
          } else {
            const v3 P = vadd(t.o, vmuls(t.d, t.tbest));  // flyscene.cpp:464
            const v3 N = interpolate_normal_slot(S, t.sbest, P);
            PF(PS_PX) = P.x; PF(PS_PY) = P.y; PF(PS_PZ) = P.z;
            PF(PS_NX) = N.x; PF(PS_NY) = N.y; PF(PS_NZ) = N.z;
            PI(PS_SLOT) = t.sbest;
            atomicAdd(&s_cnt[PI(PS_DEPTH)], 1);
            s_occ[tid] = 0u;
            s_done[tid] = 0;
            t.any = true;  // shadow phase (t.node stays at the sentinel until a light is started)
            if (n_lights > 0) {
              PI(PS_LIGHT) = 0;  // the owner takes light 0 itself, the others are on offer
              __threadfence_block();
              s_next[tid] = 1;
              start_light = true;
            } else {
              PI(PS_LIGHT) = n_lights;
              level_ready = true;
            }
          }
        } else {  // shadow phase: the owner's own ray has ended, or it only waits for its helpers
          const int mine = PI(PS_LIGHT);
          if (mine < n_lights) {
            if (t.sbest >= 0) atomicOr(&s_occ[tid], 1u << mine);
            atomicAdd(&s_done[tid], 1);
            li = atomicAdd(&s_next[tid], 1);
            if (li < n_lights) start_light = true;
            else PI(PS_LIGHT) = n_lights;  // every light is with somebody
          }
          if (!start_light) level_ready = atomicAdd(&s_done[tid], 0) >= n_lights;
        }
        if (start_light) {  // shadow(P, L) for light li (flyscene.cpp:510-526,606-612)
          const v3 P = mk3(PF(PS_PX), PF(PS_PY), PF(PS_PZ));
          const v3 L = cam_light_dir(S, li, P);
          PI(PS_LIGHT) = li;
          trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);  // flyscene.cpp:512
          t.any = true;
        } else if (level_ready) {  // calculateColor in light order, the level's records, the reflection ray
          s_next[tid] = 0x40000000;
          const unsigned occ = atomicOr(&s_occ[tid], 0u);
          const v3 P = mk3(PF(PS_PX), PF(PS_PY), PF(PS_PZ));
          const v3 N = mk3(PF(PS_NX), PF(PS_NY), PF(PS_NZ));
          const unsigned mat = (unsigned)f2i(__ldg(S.tris + 4 * (size_t)PI(PS_SLOT) + 2).w) & kTriMatMask;
          const float4 m0 = __ldg(S.materials + 3 * mat), m1 = __ldg(S.materials + 3 * mat + 1), m2 = __ldg(S.materials + 3 * mat + 2);
          v3 sum = mk3(0.f, 0.f, 0.f);
          for (int l = 0; l < n_lights; ++l) {
            if ((occ >> l) & 1u) continue;  // shadowed (flyscene.cpp:543-545)
            const float4 l1 = __ldg(S.lights + 2 * l + 1);
            sum = vadd(sum, phong(xyz(m0), xyz(m1), xyz(m2), m0.w, N, cam_light_dir(S, l, P), xyz(l1),
                                  mk3(PF(PS_OX), PF(PS_OY), PF(PS_OZ)), P));
          }
          const int depth = PI(PS_DEPTH), pixel = PI(PS_PIXEL);
          RTP_CHECK((unsigned)pixel < (unsigned)p.n_px && (unsigned)depth < (unsigned)p.max_depth, 31, pixel, p.n_px);
          p.direct[(size_t)depth * p.n_px + pixel] = make_float4(clamp01(sum.x), clamp01(sum.y), clamp01(sum.z), 0.f);
          p.ksdeep[pixel] = make_float4(m2.x, m2.y, m2.z, i2f(depth + 1));
          if (depth + 1 < p.max_depth) {
            const v3 r = reflect_dir(mk3(PF(PS_DX), PF(PS_DY), PF(PS_DZ)), N);  // :361
            const v3 o2 = vadd(P, vmuls(r, (float)0.001));                       // :362
            PF(PS_OX) = o2.x; PF(PS_OY) = o2.y; PF(PS_OZ) = o2.z;
            PF(PS_DX) = r.x; PF(PS_DY) = r.y; PF(PS_DZ) = r.z;
            PI(PS_DEPTH) = depth + 1;
            atomicAdd(&s_cnt[kMaxDepth + depth + 1], 1);
            trav_init(t, o2, r, stack);  // t.any = false: closest hit of the next level
          } else {
            have = false;
          }
        }
      }
    }
    const unsigned act = __ballot_sync(kFull, have || helping);
    if (!exhausted && __popc(act) < kRefillBelow) {
      const unsigned need = ~act;
      const int n_need = __popc(need);
      const int leader = __ffs(need) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(cursor, n_need);
      base = __shfl_sync(kFull, base, leader);
      exhausted = base + n_need >= count;
      if (!have && !helping) {
        const int q = base + __popc(need & lt);
        if (q < count) {
          const float4 o4 = ro[q], d4 = rd[q];
          PF(PS_OX) = o4.x; PF(PS_OY) = o4.y; PF(PS_OZ) = o4.z;
          PF(PS_DX) = d4.x; PF(PS_DY) = d4.y; PF(PS_DZ) = d4.z;
          PI(PS_PIXEL) = f2i(o4.w);
          PI(PS_DEPTH) = first_depth;
          trav_init(t, xyz(o4), xyz(d4), stack);
          have = true;
        }
      }
    }
    __syncwarp();
    {  // lanes that are still without work take a shadow ray off an owner of this warp
      const bool idle = !have && !helping;
      const bool offers = have && t.any && atomicAdd(&s_next[tid], 0) < n_lights;
      const unsigned idle_m = __ballot_sync(kFull, idle), off_m = __ballot_sync(kFull, offers);
      if (idle && off_m) {
        const int ol = __fns(off_m, 0, (__popc(idle_m & lt) % __popc(off_m)) + 1);  // spread the idle lanes over the owners
        const int otid = (tid & ~31) | ol;
        const int li = atomicAdd(&s_next[otid], 1);
        if (li < n_lights) {
          helping = true;
          h_owner = otid;
          h_light = li;
          const v3 P = mk3(s_f[PS_PX][otid], s_f[PS_PY][otid], s_f[PS_PZ][otid]);
          const v3 L = cam_light_dir(S, li, P);
          trav_init(t, vadd(P, vmuls(L, (float)0.003)), L, stack);
          t.any = true;
        }
      }
    }
    running = __ballot_sync(kFull, (have || helping) && t.node != kSentinel);
    if (running == 0) {
      if (__ballot_sync(kFull, have || helping) == 0 && exhausted) break;
      continue;
    }
    if ((have || helping) && t.node != kSentinel)
      // origin of the reference's BoundingBox test: the un-offset hit point for a shadow ray (the owner's, for a helper),
      // the ray's own origin for a (careful) closest-hit ray
      trav_step<2>(S, t, stack, [&] {
        const int o = helping ? h_owner : tid;
        return t.any ? mk3(s_f[PS_PX][o], s_f[PS_PY][o], s_f[PS_PZ][o]) : t.o;
      });
  }
#undef PF
#undef PI
  __syncthreads();
  if (tid < 2 * kMaxDepth && s_cnt[tid]) atomicAdd(p.counters + (tid < kMaxDepth ? kCtrHits + tid : kCtrPaths + tid - kMaxDepth), s_cnt[tid]);
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
  if (rgb8) {
    rgb8[3 * idx] = quantise(v.x);
    rgb8[3 * idx + 1] = quantise(v.y);
    rgb8[3 * idx + 2] = quantise(v.z);
  }
  if (hit) hit[idx] = f2i(v.w);
}

// Frame barrier over peer memory.  `counter` lives in GPU 0's shared image block (CUDA IPC); every rank's frame ends
// with one thread that publishes the rank's image stores (they were made by earlier kernels of this stream, so they
// are complete; the fence orders them before the arrival for every observer in the system), adds 1 to the counter
// with a system-scope atomic over NVLink, and then polls it until all `world` ranks of this frame have arrived.
// The counter only ever grows (target = world * frame number), so there is no reset to race with.  A rank that never
// arrives does not hang the others: after timeout_ns the wait gives up and counts the event in *timeouts.
__global__ void k_peer_barrier(unsigned *counter, unsigned target, unsigned long long timeout_ns, int *timeouts) {
  __threadfence_system();
  atomicAdd_system(counter, 1u);
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    if ((int)(v - target) >= 0) break;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) {
      atomicAdd(timeouts, 1);
      break;
    }
    __nanosleep(100);
  }
  __threadfence_system();
}

// Grows the padding of every child box of the hierarchy by `extra` (in place, once, when a camera or a ray batch
// starts so far from the scene that the rounding of P = o + t*d exceeds the padding the tree was built with:
// ulp(|o|) grows with the distance, the build-time pad does not).  Unbounded boxes (non-finite faces) stay as they are.
__global__ void __launch_bounds__(256) k_inflate_nodes(float4 *nodes, int n_nodes, float extra) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  auto lo = [&](float v) { return fabsf(v) < 3.0e38f ? v - extra : v; };
  auto hi = [&](float v) { return fabsf(v) < 3.0e38f ? v + extra : v; };
  float4 *N = nodes + 4 * (size_t)i;
  float4 a = N[0], b = N[1], c = N[2];
  N[0] = make_float4(lo(a.x), hi(a.y), lo(a.z), hi(a.w));
  N[1] = make_float4(lo(b.x), hi(b.y), lo(b.z), hi(b.w));
  N[2] = make_float4(lo(c.x), hi(c.y), lo(c.z), hi(c.w));
}

int g_refill = -2;  // RTP_REFILL overrides the per-launch policy below (tuning runs)
// Straggler policy (measured on B200, profiles/r01_tail.md): parking pays for closest-hit launches on
// scenes that do not fit the L2 (cold nodes make the serial chain slow: C3 closest-hit 0.47 -> 0.40 ms);
// on L2-resident scenes and for shadow rays k_coop's launch costs more than the shorter tail saves.
// v5 sweep (profiles/r01_parking.md): every ray still running after 6 steps is parked, however many lanes are left
// (C3 frame 0.770 -> 0.735 ms, one rank's 1/8 share 0.328 -> 0.255 ms); shadow rays only in small launches, where
// the launch is as long as its longest ray (1/8 share 0.255 -> 0.23 ms) - in a full-frame launch their k_coop pass
// costs more than the tail it removes (0.735 -> 0.79 ms).
int g_budget = 6;       // RTP_BUDGET: steps a warp spends on a batch before the rays still running are parked (0: never)
bool g_budget_from_env = false;
int g_park_lanes = 32;  // RTP_PARK_LANES: ... if at most this many lanes are still running
int g_park_any = 0;     // RTP_PARK_ANY=1: park shadow rays in every launch
int g_any_small = 400000;  // RTP_ANY_SMALL: ... or in launches over at most this many paths
int g_budget_any = 6;   // RTP_BUDGET_ANY: step budget of shadow-ray launches (a step of an unoccluded ray is 32 nodes)
int g_park_lanes_any = 32;
int g_park_small = 0;   // RTP_PARK_SMALL=1: park on L2-resident scenes too
int g_step_nodes = kMaxNodesPerStep;  // RTP_STEP_NODES: inner nodes per step of scattered-ray launches (lane refill mode)

// Refill policy: a warp refills its idle lanes from the global ray list when fewer than `refill`
// lanes still hold a ray.  Coherent waves (camera rays, their point-light shadow rays) run
// fastest in batch mode (refill only when the whole warp is idle: neighbouring pixels stay
// together, measured +6% on C3); scattered waves (reflection levels, caller-supplied rays)
// keep at least 24 lanes busy.
void read_policy_env() {
  if (g_refill != -2) return;
  const char *e = std::getenv("RTP_REFILL");
  g_refill = e ? std::atoi(e) : -1;
  if ((e = std::getenv("RTP_BUDGET"))) { g_budget = g_budget_any = std::atoi(e); g_budget_from_env = true; }
  if ((e = std::getenv("RTP_PARK_LANES"))) g_park_lanes = g_park_lanes_any = std::atoi(e);
  if ((e = std::getenv("RTP_PARK_ANY"))) g_park_any = std::atoi(e);
  if ((e = std::getenv("RTP_ANY_SMALL"))) g_any_small = std::atoi(e);
  if ((e = std::getenv("RTP_BUDGET_ANY"))) g_budget_any = std::atoi(e);
  if ((e = std::getenv("RTP_PARK_LANES_ANY"))) g_park_lanes_any = std::atoi(e);
  if ((e = std::getenv("RTP_PARK_SMALL"))) g_park_small = std::atoi(e);
  if ((e = std::getenv("RTP_STEP_NODES"))) g_step_nodes = std::min(std::max(std::atoi(e), 1), kMaxNodesPerStep);
}

template <class Job>
int launch_trace(const Job &job, int blocks, const LaunchCfg &cfg, int *counters, int launch_slot, bool coherent,
                 long long n_paths, cudaStream_t st) {
  const int threads = cfg.threads;
  read_policy_env();
  static int refill_any = -2, refill_closest = -2;
  if (refill_any == -2) {
    const char *e = std::getenv("RTP_REFILL_ANY");
    refill_any = e ? std::atoi(e) : -1;
    e = std::getenv("RTP_REFILL_CLOSEST");
    refill_closest = e ? std::atoi(e) : -1;
  }
  int refill = g_refill >= 0 ? g_refill : (coherent ? 1 : 24);
  if (coherent && (Job::kAny ? refill_any : refill_closest) >= 0) refill = Job::kAny ? refill_any : refill_closest;
  ParkList pl;
  pl.slots = cfg.park_slots;
  pl.ctr = counters + kCtrPark + 2 * launch_slot;
  pl.cap = cfg.park_cap;
  const bool park = g_budget > 0 && cfg.park_slots && (!Job::kAny || g_park_any || n_paths <= g_any_small) &&
                    (cfg.scene_exceeds_l2 || g_park_small);
  pl.budget = park ? (Job::kAny ? g_budget_any : g_budget) : 0x3fffffff;
  pl.lanes = Job::kAny ? g_park_lanes_any : g_park_lanes;
  pl.step_nodes = coherent ? kMaxNodesPerStep : g_step_nodes;
  if (pl.step_nodes < kMaxNodesPerStep && pl.budget < 0x3fffffff) pl.budget *= kMaxNodesPerStep / pl.step_nodes;  // the budget is in 32-node steps
  if (refill <= 1) k_trace<Job, 1><<<blocks, threads, 0, st>>>(job, pl);
  else if (refill <= 8) k_trace<Job, 8><<<blocks, threads, 0, st>>>(job, pl);
  else k_trace<Job, 24><<<blocks, threads, 0, st>>>(job, pl);
  if (pl.budget >= 0x3fffffff) return 1;
  k_coop<Job><<<cfg.blocks_coop, 128, 0, st>>>(job, pl);
  return 2;
}

}  // namespace

void query_launch_cfg(int device, LaunchCfg &cfg) {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (const char *e = std::getenv("RTP_TRACE_THREADS")) {  // tuning runs: 32, 64 or 128 threads per CTA
    const int t = std::atoi(e);
    if (t == 32 || t == 64 || t == 128) cfg.threads = t;
  }
  int b0 = 1, b1 = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b0, k_trace<FrameClosest<true>, 1>, cfg.threads, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, k_trace<FrameAny<true>, 1>, cfg.threads, 0);
  cfg.blocks_closest = sms * (b0 > 0 ? b0 : 1);
  cfg.blocks_any = sms * (b1 > 0 ? b1 : 1);
  int b4 = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b4, k_trace<RaysAny, 24>, cfg.threads, 0);
  cfg.blocks_rays_any = sms * (b4 > 0 ? b4 : 1);
  int b2 = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, k_paths<24>, 128, 0);
  cfg.blocks_paths = sms * (b2 > 0 ? b2 : 1);
  int b3 = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b3, k_camera<true>, 128, 0);
  cfg.blocks_camera = sms * (b3 > 0 ? b3 : 1);
  cfg.blocks_coop = sms * 2;
  if (const char *e = std::getenv("RTP_COOP_BLOCKS")) cfg.blocks_coop = sms * std::max(1, std::atoi(e));  // per SM (tuning runs)
}

void launch_peer_barrier(unsigned *counter, unsigned target, unsigned long long timeout_ns, int *timeouts, cudaStream_t st) {
  k_peer_barrier<<<1, 1, 0, st>>>(counter, target, timeout_ns, timeouts);
}

void launch_inflate_nodes(float4 *nodes, int n_nodes, float extra, cudaStream_t st) {
  if (n_nodes > 0) k_inflate_nodes<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes, n_nodes, extra);
}

void launch_raygen(const FrameParams &p, cudaStream_t st) { k_raygen<<<(p.n_px + 255) / 256, 256, 0, st>>>(p); }

template <bool CAM>
static int launch_closest_t(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  FrameClosest<CAM> j;
  j.S = p.S;
  j.cam = p.cam;
  j.tile_ids = p.tile_ids;
  j.tiles_x = p.tiles_x; j.width = p.width; j.height = p.height;
  j.ro = p.ray_o[depth & 1];
  j.rd = p.ray_d[depth & 1];
  j.hit_slot = p.hit_slot;
  j.pbuf = p.pbuf; j.dbuf = p.dbuf; j.hitq = p.hitq;
  j.hits_ctr = p.counters + kCtrHits + depth;
  j.primface = depth == 0 ? p.primface : nullptr;
  j.ksdeep = p.ksdeep;
  j.cursor = p.counters + kCtrFetchClosest + depth;
  j.count_ptr = depth == 0 ? nullptr : p.counters + kCtrPaths + depth;
  j.count_fixed = p.n_px;
  return launch_trace(j, cfg.blocks_closest, cfg, p.counters, depth, depth == 0 && !p.user_rays, p.n_px, st);
}

int launch_closest(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  if (depth == 0 && !p.user_rays) return launch_closest_t<true>(p, depth, cfg, st);
  return launch_closest_t<false>(p, depth, cfg, st);
}

template <bool DENSE>
static int launch_anyhit_t(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  FrameAny<DENSE> j;
  j.S = p.S;
  j.pbuf = p.pbuf;
  j.occl = p.occl;
  j.cursor = p.counters + kCtrFetchAny + depth;
  j.hits_ptr = p.counters + kCtrHits + depth;
  j.hit_slot = p.hit_slot;
  j.n_px = p.n_px;
  return launch_trace(j, cfg.blocks_any, cfg, p.counters, 8 + depth, depth == 0 && !p.user_rays, p.n_px, st);
}

int launch_anyhit(const FrameParams &p, int depth, const LaunchCfg &cfg, cudaStream_t st) {
  if (depth == 0 && !p.user_rays) return launch_anyhit_t<true>(p, depth, cfg, st);  // camera level: dense records
  return launch_anyhit_t<false>(p, depth, cfg, st);
}

// Would the camera level's shadow launch park rays (launch_trace's policy)?  Then their outcomes arrive in k_coop and the
// shading cannot ride in the launch.
bool anyhit_shade_ok(const FrameParams &p, const LaunchCfg &cfg) {
  read_policy_env();
  static int on = -1;
  if (on < 0) {
    // RTP_ANY_SHADE=1 (experiment, measured level on the device and 3 % slower end to end - profiles/r02_camera.md
    // section 7): the shadow launch shades; default: k_accum / k_resolve_image after it
    const char *e = std::getenv("RTP_ANY_SHADE");
    on = e ? std::atoi(e) : 0;
  }
  const bool parks = g_budget > 0 && cfg.park_slots && (g_park_any || p.n_px <= g_any_small) && (cfg.scene_exceeds_l2 || g_park_small);
  return on && !parks && g_refill < 0 && !p.user_rays && p.tile_ids && p.S.n_lights >= 1 && p.S.n_lights <= kCamMaxLights;
}

// shadow() for the camera level's (pixel, light) pairs with the level's shading in the same launch (FrameAnyShade)
int launch_anyhit_shade(const FrameParams &p, const LaunchCfg &cfg, cudaStream_t st) {
  ParkList pl;
  pl.slots = nullptr; pl.ctr = p.counters + kCtrPark + 2 * 8; pl.cap = 0;
  pl.budget = 0x3fffffff; pl.lanes = 32; pl.step_nodes = kMaxNodesPerStep;  // never parks
  auto fill = [&](auto &j) {
    j.S = p.S;
    j.pbuf = p.pbuf;
    j.occl = p.occl;
    j.cursor = p.counters + kCtrFetchAny;
    j.hits_ptr = p.counters + kCtrHits;
    j.hit_slot = p.hit_slot;
    j.n_px = p.n_px;
    j.p = p;
  };
  if (p.max_depth == 1) {
    FrameAnyShade<true> j;
    fill(j);
    k_trace<FrameAnyShade<true>, 1><<<cfg.blocks_any, cfg.threads, 0, st>>>(j, pl);
  } else {
    FrameAnyShade<false> j;
    fill(j);
    k_trace<FrameAnyShade<false>, 1><<<cfg.blocks_any, cfg.threads, 0, st>>>(j, pl);
  }
  return 1;
}

bool camera_fusable(const FrameParams &p) {
  return !p.user_rays && p.tile_ids && p.S.n_lights <= kCamMaxLights && p.n_px < kCamMaxPixels;
}

int launch_camera(const FrameParams &p, const LaunchCfg &cfg, cudaStream_t st, int phase) {
  read_policy_env();
  CamPark pk;
  pk.slots = cfg.park_slots;
  pk.ctr = p.counters + kCtrPark;  // launch slot 0
  pk.cap = cfg.park_cap;
  // the straggler policy of launch_trace: closest-hit rays park on scenes that spill out of the L2, shadow rays only
  // when the launch is small as well
  const bool park = g_budget > 0 && cfg.park_slots && (cfg.scene_exceeds_l2 || g_park_small);
  const bool park_any = park && (g_park_any || p.n_px <= g_any_small);
  // a parked pixel is dear here (k_coop_camera also traces its shadow rays and shades it): above 600 k pixels per launch
  // sequence only the really long rays are parked (C3 1/2 share: 0.441 ms at 6 steps, 0.391 at 14; 1/8 share: 0.172 / 0.235)
  const int patient = g_budget_from_env || p.n_px * cfg.groups <= 600000 ? 1 : 2;
  pk.budget = park ? patient * g_budget + (patient - 1) * 2 : 0x3fffffff;  // 6 steps, 14 for the bigger shares
  pk.lanes = g_park_lanes;
  pk.budget_any = park_any ? patient * g_budget_any + (patient - 1) * 2 : 0x3fffffff;
  pk.lanes_any = g_park_lanes_any;
  int n = 0;
  if (phase == 0 || phase == 1) {
    if (p.max_depth == 1) k_camera<true><<<cfg.blocks_camera, 128, 0, st>>>(p, pk);
    else k_camera<false><<<cfg.blocks_camera, 128, 0, st>>>(p, pk);
    ++n;
  }
  if (park && (phase == 0 || phase == 2)) {
    if (p.max_depth == 1) k_coop_camera<true><<<cfg.blocks_coop, 128, 0, st>>>(p, pk);
    else k_coop_camera<false><<<cfg.blocks_coop, 128, 0, st>>>(p, pk);
    ++n;
  }
  return n;
}

void launch_paths(const FrameParams &p, int first_depth, const LaunchCfg &cfg, cudaStream_t st) {
  static int refill = -1;
  if (refill < 0) {
    const char *e = std::getenv("RTP_PATH_REFILL");
    refill = e ? std::atoi(e) : 24;
    if ((e = std::getenv("RTP_PATH_WAIT"))) g_path_wait = std::atoi(e);
    if ((e = std::getenv("RTP_PATH_RUN"))) g_path_run = std::atoi(e);
  }
  static int help = -1;
  if (help < 0) {
    const char *e = std::getenv("RTP_PATH_HELP");
    help = e ? std::atoi(e) : 0;
  }
  // RTP_PATH_HELP=1 (experiment, measured slower - profiles/r02_camera.md section 6): idle lanes trace shadow rays of
  // their warp's paths
  if (help && p.S.n_lights <= kCamMaxLights) {
    if (refill <= 8) k_paths_help<8><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
    else if (refill <= 24) k_paths_help<24><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
    else k_paths_help<32><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
    return;
  }
  if (refill <= 8) k_paths<8><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
  else if (refill <= 24) k_paths<24><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
  else k_paths<32><<<cfg.blocks_paths, 128, 0, st>>>(p, first_depth, g_path_wait, g_path_run);
}

void launch_accum(const FrameParams &p, int depth, cudaStream_t st) {
  k_accum<<<(p.n_px + 255) / 256, 256, 0, st>>>(p, depth);
}

void launch_resolve(const FrameParams &p, cudaStream_t st) {
  const int blocks = (p.n_px + 255) / 256;
  const bool image = p.img_rgb || p.img_rgb8 || p.img_hit;
  if (p.max_depth == 1) {  // the level's shading happens in the same kernel (no k_accum)
    if (image) k_resolve_image<true><<<blocks, 256, 0, st>>>(p);
    else k_resolve<true><<<blocks, 256, 0, st>>>(p);
  } else {
    if (image) k_resolve_image<false><<<blocks, 256, 0, st>>>(p);
    else k_resolve<false><<<blocks, 256, 0, st>>>(p);
  }
}

void launch_untile(const float4 *all, size_t rank_stride_f4, const int *local_of_tile, int width, int height,
                   int world, int only_rank, float *rgb, unsigned char *rgb8, int *hit, cudaStream_t st) {
  const long long n = (long long)width * height;
  const int tiles_x = (width + kTileDim - 1) / kTileDim;
  k_untile<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(all, rank_stride_f4, local_of_tile, width, height, tiles_x,
                                                        world, only_rank, rgb, rgb8, hit);
}

int launch_rays_closest(const DeviceScene &S, const float *rays, long long n, int *face, float *t, int *cursor,
                         const LaunchCfg &cfg, cudaStream_t st) {
  RaysClosest j;
  j.S = S; j.rays = rays; j.face = face; j.tout = t; j.cursor = cursor; j.n = (int)n;
  return launch_trace(j, cfg.blocks_closest, cfg, cursor - kCtrAux, 16, false, n, st);
}

int launch_rays_anyhit(const DeviceScene &S, const float *rays, long long n, unsigned char *occ, int *cursor,
                        const LaunchCfg &cfg, cudaStream_t st) {
  RaysAny j;
  j.S = S; j.rays = rays; j.occ = occ; j.cursor = cursor; j.n = (int)n;
  return launch_trace(j, cfg.blocks_rays_any, cfg, cursor - kCtrAux, 16, false, 1ll << 40, st);  // ray batches: never (C5)
}

// ---- ray batches without any allocation (rtp_trace_rays_device_noalloc) ----------------------------------------
// The scratch a ray-batch launch needs - the ray cursor, the parked-ray counters and the parked-ray slots that the
// scene otherwise owns (cudaMalloc at rtp_scene_create) - lives in this file-scope array instead: every GPU that loads
// the library holds kNoAllocBytes until exit.  Head = one counters block (same layout as a frame's, so launch_trace is
// shared), then kNoAllocParkCap parked-ray slots.  In host code `g_noalloc_scratch` is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kNoAllocHeadInts = 128;  // >= kCtrInts, keeps the slots 256-byte aligned
static_assert(kNoAllocHeadInts >= kCtrInts, "counters block must fit the head of the no-alloc scratch");
constexpr size_t kNoAllocInts = kNoAllocHeadInts + (size_t)kNoAllocParkCap * (kParkSlotBytes / sizeof(int));
static __device__ __align__(256) int g_noalloc_scratch[kNoAllocInts];

cudaError_t launch_rays_noalloc(int device, const DeviceScene &S, const float *rays, long long n, int any, int *face,
                                float *t, unsigned char *occ, LaunchCfg cfg, cudaStream_t st, int *launches) {
  constexpr int kMaxDev = 64;
  static int *ptr_of[kMaxDev];  // per device: the array's address there (0 = not asked yet)
  static std::mutex mu;
  if (device < 0 || device >= kMaxDev) return cudaErrorInvalidDevice;
  int *scratch;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr_of[device]) {  // the caller made `device` current (DeviceGuard)
      int *p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
      if (e != cudaSuccess) return e;
      ptr_of[device] = p;
    }
    scratch = ptr_of[device];
  }
  // what must start at zero: the ray cursor and the launch's (rays parked, k_coop cursor) pair; the slots are
  // written before they are read and keep the previous call's contents
  cudaError_t e = cudaMemsetAsync(scratch, 0, kNoAllocHeadInts * sizeof(int), st);
  if (e != cudaSuccess) return e;
  cfg.park_slots = scratch + kNoAllocHeadInts;
  cfg.park_cap = kNoAllocParkCap;
  int *cursor = scratch + kCtrAux;
  *launches = any ? launch_rays_anyhit(S, rays, n, occ, cursor, cfg, st) : launch_rays_closest(S, rays, n, face, t, cursor, cfg, st);
  return cudaGetLastError();
}

}  // namespace rtp

#ifdef RTP_CHECKED
// failed bounds checks since the library was loaded: out[0] count, out[1..3] code / value / limit of the first
extern "C" int rtp_debug_checks(unsigned *out) { return (int)cudaMemcpyFromSymbol(out, rtp::g_rtp_check, 4 * sizeof(unsigned)); }
#endif
#ifdef RTP_TRACE_TIMING
extern "C" int rtp_debug_timing(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, rtp::g_warp_timing, sizeof(unsigned long long) * 2 * 8192 * 8);
}
#endif
