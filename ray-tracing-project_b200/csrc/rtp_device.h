// rtp_device.h - device-side scene layout, BVH traversal and the reference's exact
// ray/triangle, ray/BoundingBox and shading arithmetic.
//
// Everything here is __host__ __device__: the CUDA kernels (rtp_kernels.cu) are the product;
// tests/hostsim compiles the same functions with g++ to check them against the oracle on a
// machine without a GPU (test infrastructure, never linked into the product library).
//
// Layout in HBM (all 16-byte aligned, read through the read-only path, L2-resident for
// scenes up to ~1 M triangles):
//   nodes  : 4 x float4 per BVH2 node (64 B)  n0 = c0.lo.x c0.hi.x c0.lo.y c0.hi.y
//                                              n1 = c1.lo.x c1.hi.x c1.lo.y c1.hi.y
//                                              n2 = c0.lo.z c0.hi.z c1.lo.z c1.hi.z
//                                              n3 = child0, child1 (int bits), -, -
//            child >= 0: inner node index; child < 0: leaf, ~child = first_tri << 3 | (count-1)
//   tris   : 4 x float4 per triangle in leaf order (64 B)
//                                              T0 = world v0, reference visitation rank
//                                              T1 = world v1, reference BoundingBox id
//                                              T2 = world v2, (material id + 1) | flags
//                                              T3 = unit face normal, plane constant n.v0
//   vnorm  : 3 x float4 per triangle (48 B): the three unit vertex normals (shading only)
//   boxes  : 2 x float4 per reference BoundingBox (object-space low / high)
//   face_of: int per triangle slot -> reference face id
#pragma once
#include "rtp_math.h"

// float4 / make_float4 come from the CUDA toolkit's header-only vector types, for nvcc and for
// plain g++ translation units alike (same 16-byte aligned type everywhere).
#include <vector_functions.h>
#include <vector_types.h>

namespace rtp {

constexpr int kStackSize = 64;
constexpr int kSentinel = 0x7fffffff;
constexpr int kMaxNodesPerStep = 32;
constexpr unsigned kTriFlagNormalCheck = 0x80000000u;  // vertex normals may cancel: run the Q12 check
constexpr unsigned kTriMatMask = 0x0000ffffu;
constexpr int kMaxDepth = 8;   // direct-colour slots per pixel
constexpr int kTileDim = 32;   // screen tile edge (multi-GPU interleave unit)
constexpr int kTilePixels = kTileDim * kTileDim;

// Traversal stack access.  Default: a per-thread array in local memory (L1-resident).  -DRTP_SMEM_STACK=<entries>
// (experiment build, profiles/r02_design_items.md): the stack lives in shared memory, interleaved over the 128 threads
// of the CTA (entry i of thread t at [i * 128 + t]: conflict-free when a warp's lanes sit at the same depth).
#if defined(RTP_SMEM_STACK) && defined(__CUDA_ARCH__)
#define RTP_STK(s, i) (s)[(i) * 128]
#else
#define RTP_STK(s, i) (s)[(i)]
#endif

struct DeviceScene {
  const float4 *nodes;
  const float4 *tris;
  const float4 *vnorm;
  const float4 *boxes;
  const int *face_of;
  const float4 *materials;  // 3 x float4 per entry: (ka, Ns) (kd, -) (ks, -); entry 0 = Flyscene defaults
  const float4 *lights;     // 2 x float4 per light: (pos-or-vector, is_directional) (rgb, -)
  float minv[12];           // getShapeModelMatrix().inverse(), 3x4 row-major
  int n_lights;
  int n_tris, n_nodes, n_boxes;
};

struct DeviceCamera {
  float eye[3];
  float inv_view[12];
  float vp[4];
  float sx, sy;  // aspect_ratio*scale, scale (camera.hpp:166-167)
};

#if defined(__CUDA_ARCH__)
#define RTP_LDG4(p) __ldg(p)
#define RTP_LDGI(p) __ldg(p)
#else
#define RTP_LDG4(p) (*(p))
#define RTP_LDGI(p) (*(p))
#endif

RTP_HD v3 xyz(float4 a) { return mk3(a.x, a.y, a.z); }

// Pins a value to the point where it is used: the compiler may not hoist arithmetic that depends on it out
// of the enclosing branch.  (ptxas otherwise evaluates the loop-invariant part of the reference BoundingBox
// test - two 3x4 transforms, ~40 instructions - at every leaf visit, although only a hit candidate needs it.)
#if defined(__CUDA_ARCH__)
#define RTP_PIN(x) asm volatile("" : "+f"(x))
#else
#define RTP_PIN(x) ((void)0)
#endif
RTP_HD v3 pinned(v3 a) { RTP_PIN(a.x); RTP_PIN(a.y); RTP_PIN(a.z); return a; }

// -DRTP_CHECKED (variant build libflyscene_b200_checked.so, profiles/r02_checked.md): bounds checks of the library's own
// on every index the kernels form - traversal stack, node / triangle / BoundingBox ids, parked-ray slots, k_coop's
// shared stack, hit records, ray lists, image pixels.  (compute-sanitizer is closed on the pool this is developed on.)
// A failed check is counted, its first code and value kept; rtp_debug_checks() reads the block.  Off in the product.
#if defined(RTP_CHECKED) && defined(__CUDACC__)
static __device__ unsigned g_rtp_check[4];  // [0] failed checks, [1] code of the first, [2] its value, [3] its limit
#define RTP_CHECK(cond, code, val, lim)                                                                       \
  do {                                                                                                        \
    if (!(cond)) {                                                                                            \
      if (atomicAdd(&g_rtp_check[0], 1u) == 0u) { g_rtp_check[1] = (code); g_rtp_check[2] = (unsigned)(val); g_rtp_check[3] = (unsigned)(lim); } \
    }                                                                                                         \
  } while (0)
#else
#define RTP_CHECK(cond, code, val, lim) ((void)0)
#endif
#if defined(RTP_CHECKED) && defined(__CUDA_ARCH__)
#define RTP_CHECK_DEV(cond, code, val, lim) RTP_CHECK(cond, code, val, lim)
#else
#define RTP_CHECK_DEV(cond, code, val, lim) ((void)0)
#endif

// Traversal work counters: compiled in only for tests/hostsim (CPU, -DRTP_TRAV_STATS), where they
// give nodes / leaves / triangle tests per ray for the roofline discussion in DESIGN.md.
#if defined(RTP_TRAV_STATS) && !defined(__CUDA_ARCH__)
struct TravStats { long long nodes, leaves, tris, careful; };
extern thread_local TravStats g_trav_stats;
#define RTP_STAT(x) (++g_trav_stats.x)
#else
#define RTP_STAT(x) ((void)0)
#endif

// ---- Camera::screenToWorld + direction (camera.hpp:155-173, flyscene.cpp:308) ----------------
RTP_HD v3 primary_dir(const DeviceCamera &c, int i, int j) {
  float ri = (float)i, rj = (float)j;
  v3 nc;
  // NDC in double precision, rounded to float when stored in the Vector3f
  nc.x = (float)(2.0 * (double)fsub(ri, c.vp[0]) / (double)c.vp[2] - 1.0);
  nc.y = (float)(1.0 - 2.0 * (double)fsub(rj, c.vp[1]) / (double)c.vp[3]);
  nc.z = -1.0f;
  nc.x = fmul(nc.x, c.sx);
  nc.y = fmul(nc.y, c.sy);
  v3 w = affine_mul(c.inv_view, nc);
  return vnormalized(vsub(w, mk3(c.eye[0], c.eye[1], c.eye[2])));
}

// ---- the reference's BoundingBox slab test, exactly (flyscene.cpp:484-507) ------------------
// Decides whether the reference would have looked at the faces of box `b` at all for this ray.
RTP_HD bool ref_box_pass(const DeviceScene &S, v3 bo, v3 d, int b) {
  float4 lo = RTP_LDG4(S.boxes + 2 * b), hi = RTP_LDG4(S.boxes + 2 * b + 1);
  v3 o2 = affine_mul(S.minv, bo);
  v3 d2 = vnormalized(linear_mul(S.minv, d));
  v3 tmin = vsub(xyz(lo), o2), tmax = vsub(xyz(hi), o2);
  tmin.x = fdiv(tmin.x, d2.x); tmax.x = fdiv(tmax.x, d2.x);
  tmin.y = fdiv(tmin.y, d2.y); tmax.y = fdiv(tmax.y, d2.y);
  tmin.z = fdiv(tmin.z, d2.z); tmax.z = fdiv(tmax.z, d2.z);
  float tin = std_max(std_min(tmin.x, tmax.x), std_max(std_min(tmin.y, tmax.y), std_min(tmin.z, tmax.z)));
  float tout = std_min(std_max(tmin.x, tmax.x), std_min(std_max(tmin.y, tmax.y), std_max(tmin.z, tmax.z)));
  return !(tin > tout || tout < 0.f);
}

// ---- interpolateNormal (flyscene.cpp:572-600) given world vertices and unit vertex normals ----
RTP_HD v3 interpolate_normal(v3 v0, v3 v1, v3 v2, v3 fn, v3 n0, v3 n1, v3 n2, v3 P) {
  v3 e0 = vsub(v1, v0), e1 = vsub(v2, v1), e2 = vsub(v0, v2);
  v3 a0v = vcross(e0, vsub(P, v0)), a1v = vcross(e1, vsub(P, v1)), a2v = vcross(e2, vsub(P, v2));
  if (vdot(fn, a0v) < 0.f || vdot(fn, a1v) < 0.f || vdot(fn, a2v) < 0.f) return mk3(0.f, 0.f, 0.f);
  float area0 = fdiv(vnorm(a0v), 2.f), area1 = fdiv(vnorm(a1v), 2.f), area2 = fdiv(vnorm(a2v), 2.f);
  float area = fdiv(vnorm(vcross(e0, vneg(e2))), 2.f);
  v3 s = vadd(vadd(vdivs(vmuls(n0, area1), area), vdivs(vmuls(n1, area2), area)), vdivs(vmuls(n2, area0), area));
  return vnormalized(s);
}

RTP_HD v3 interpolate_normal_slot(const DeviceScene &S, int slot, v3 P) {
  const float4 *T = S.tris + 4 * (size_t)slot;
  const float4 *N = S.vnorm + 3 * (size_t)slot;
  return interpolate_normal(xyz(RTP_LDG4(T)), xyz(RTP_LDG4(T + 1)), xyz(RTP_LDG4(T + 2)), xyz(RTP_LDG4(T + 3)),
                            xyz(RTP_LDG4(N)), xyz(RTP_LDG4(N + 1)), xyz(RTP_LDG4(N + 2)), P);
}

// ---- calculateDistance (flyscene.cpp:444-478) on one flattened triangle ----------------------
// Returns true when the reference would report t >= 0 for this face, with 0 <= t <= tmax.
RTP_HD bool tri_exact(const DeviceScene &S, int slot, v3 o, v3 d, float tmax, float &t_out, float4 &T0o, float4 &T1o) {
  const float4 *T = S.tris + 4 * (size_t)slot;
  float4 T3 = RTP_LDG4(T + 3);
  v3 n = xyz(T3);
  float dn = vdot(n, d);
  if (dn == 0.f) return false;                       // :453
  float t = fdiv(fsub(T3.w, vdot(o, n)), dn);         // :459-463
  if (!(t >= 0.f) || t > tmax) return false;           // caller's 0 <= t (< best) test, :387
  v3 P = vadd(o, vmuls(d, t));                        // :464
  float4 T0 = RTP_LDG4(T), T1 = RTP_LDG4(T + 1), T2 = RTP_LDG4(T + 2);
  v3 v0 = xyz(T0), v1 = xyz(T1), v2 = xyz(T2);
  v3 e0 = vsub(v1, v0), e1 = vsub(v2, v1), e2 = vsub(v0, v2);
  if (vdot(n, vcross(e0, vsub(P, v0))) < 0.f) return false;   // :591
  if (vdot(n, vcross(e1, vsub(P, v1))) < 0.f) return false;
  if (vdot(n, vcross(e2, vsub(P, v2))) < 0.f) return false;
  if ((unsigned)f2i(T2.w) & kTriFlagNormalCheck) {            // :467 (only when it can trigger)
    const float4 *N = S.vnorm + 3 * (size_t)slot;
    v3 nn = interpolate_normal(v0, v1, v2, n, xyz(RTP_LDG4(N)), xyz(RTP_LDG4(N + 1)), xyz(RTP_LDG4(N + 2)), P);
    if (vnorm(nn) == 0.f) return false;
  }
  t_out = t;
  T0o = T0; T1o = T1;
  return true;
}

// ---- BVH2 traversal ---------------------------------------------------------------------------
// ANY = false: calculateMinimumFace (flyscene.cpp:373-396): smallest t >= 0 over the faces of
//   every reference box whose slab test passes, ties to the lowest visitation rank.
// ANY = true : shadow (flyscene.cpp:510-526): any such face.
// o/d: origin and direction used for the face tests; bo: origin used for the reference box test
// (== o for closest-hit, the un-offset point P for shadow rays).
//
// The traversal is resumable (Trav + trav_step) so that the persistent-warp kernels can refill
// idle lanes with new rays between steps; traverse<> is the plain run-to-completion loop.
struct Trav {
  float idx, idy, idz, oodx, oody, oodz;
  v3 o, d;
  int node, sp;
  float tbest;
  unsigned rbest;
  int sbest;
  bool careful;  // closest-hit re-run: filter every candidate by the reference's box test
  bool any;      // trav_step<2> only (k_paths: lanes of one warp mix closest-hit and shadow rays): this ray is a shadow ray
};

RTP_HD void trav_init(Trav &t, v3 o, v3 d, int *stack) {
  const float ooeps = 1.0e-24f;
  float dx = fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x);
  float dy = fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y);
  float dz = fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z);
  t.idx = 1.0f / dx; t.idy = 1.0f / dy; t.idz = 1.0f / dz;
  t.oodx = o.x * t.idx; t.oody = o.y * t.idy; t.oodz = o.z * t.idz;
  t.o = o; t.d = d;
  t.node = 0;  // the root is always an inner node
  t.sp = 0;
  RTP_STK(stack, 0) = kSentinel;
  t.tbest = INFINITY;
  t.rbest = 0xffffffffu;
  t.sbest = -1;
  t.careful = false;
  t.any = false;
}

// One macro step: descend through inner nodes to the next leaf, then test that leaf (and any
// leaves popped right after it).  Returns true when the ray is finished.
// t.careful = true : every candidate face is filtered by the reference's own box test before it
//                    may win (exact, slow).
// t.careful = false: closest-hit only - candidates are accepted unfiltered; the caller checks the
//                    final winner's box once (closest_needs_rerun) and re-runs in careful mode if
//                    that test fails.  The unfiltered minimum, when its box passes, is also the
//                    filtered minimum.  A runtime flag (not a second instantiation) keeps the whole
//                    traversal state in registers: no out-of-line re-trace function.
// A step visits at most kMaxNodesPerStep inner nodes (it is resumable: t.node = the next inner node),
// so that a ray with a long leaf-free stretch - an unoccluded shadow ray - still comes back to the
// caller, who may hand a straggler to the whole warp (rtp_kernels.cu).
// box_origin: callable returning the origin for the reference box test.  It is evaluated only when a candidate
// face has to be filtered (for a frame's shadow rays it is a division and a global load that a step otherwise
// paid on entry, every time).
// MODE 0: closest hit, 1: any hit (compile-time), 2: per ray at run time (t.any) - the deep-path kernel, whose warps hold
// closest-hit and shadow rays side by side; same arithmetic and the same decisions per ray in every mode.
// max_trips: inner nodes a step may visit (<= kMaxNodesPerStep).  Lanes that reach their leaf early idle until the
// slowest lane of the warp leaves the node loop, so scattered rays want a short leash (k_trace's step_nodes).
template <int MODE, class BoxOrigin>
RTP_HD bool trav_step(const DeviceScene &S, Trav &t, int *stack, BoxOrigin box_origin, const int max_trips = kMaxNodesPerStep) {
  const bool ANY = MODE == 2 ? t.any : MODE == 1;
  int node = t.node, sp = t.sp;
  for (int trip = 0; node >= 0 && node != kSentinel && trip < max_trips; ++trip) {
    const float4 *Np = S.nodes + 4 * (size_t)node;
    RTP_CHECK_DEV(node < S.n_nodes, 1, node, S.n_nodes);
    RTP_STAT(nodes);
    float4 n0 = RTP_LDG4(Np), n1 = RTP_LDG4(Np + 1), n2 = RTP_LDG4(Np + 2), n3 = RTP_LDG4(Np + 3);
    float c0lox = fmaf(n0.x, t.idx, -t.oodx), c0hix = fmaf(n0.y, t.idx, -t.oodx);
    float c0loy = fmaf(n0.z, t.idy, -t.oody), c0hiy = fmaf(n0.w, t.idy, -t.oody);
    float c0loz = fmaf(n2.x, t.idz, -t.oodz), c0hiz = fmaf(n2.y, t.idz, -t.oodz);
    float c1lox = fmaf(n1.x, t.idx, -t.oodx), c1hix = fmaf(n1.y, t.idx, -t.oodx);
    float c1loy = fmaf(n1.z, t.idy, -t.oody), c1hiy = fmaf(n1.w, t.idy, -t.oody);
    float c1loz = fmaf(n2.z, t.idz, -t.oodz), c1hiz = fmaf(n2.w, t.idz, -t.oodz);
    float c0min = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), 0.f));
    float c0max = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), t.tbest));
    float c1min = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), 0.f));
    float c1max = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), t.tbest));
    bool h0 = c0min <= c0max, h1 = c1min <= c1max;
    int ch0 = f2i(n3.x), ch1 = f2i(n3.y);
    if (!h0 && !h1) {
      node = RTP_STK(stack, sp); --sp;  // stack[0] is the sentinel: popping it ends the ray, sp is not used after
    } else {
      node = h0 ? ch0 : ch1;
      if (h0 && h1) {
        if (!ANY && c1min < c0min) { int tmp = node; node = ch1; ch1 = tmp; }
        ++sp; RTP_STK(stack, sp) = ch1;
        RTP_CHECK_DEV(sp < kStackSize, 2, sp, kStackSize);
#if defined(RTP_PREFETCH) && defined(__CUDA_ARCH__)  // experiment build: pull the postponed child towards the L2 now
        {
          const void *a = ch1 >= 0 ? (const void *)(S.nodes + 4 * (size_t)ch1) : (const void *)(S.tris + 4 * (size_t)((~ch1) >> 3));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
        }
#endif
      }
    }
  }
  while (node < 0) {
    int enc = ~node;
    int first = enc >> 3, cnt = (enc & 7) + 1;
    RTP_CHECK_DEV(first + cnt <= S.n_tris, 3, first + cnt, S.n_tris);
    RTP_STAT(leaves);
    for (int k = 0; k < cnt; ++k) {
      RTP_STAT(tris);
      float tt;
      float4 T0, T1;
      if (!tri_exact(S, first + k, t.o, t.d, t.tbest, tt, T0, T1)) continue;
      unsigned rank = (unsigned)f2i(T0.w);
      // tt <= tbest here.  Equal t: the first-visited face wins (flyscene.cpp:387 is a strict <),
      // and t == +inf never beats the initial INFINITY.
      if (!ANY && !(tt < t.tbest || (t.sbest >= 0 && rank < t.rbest))) continue;
      RTP_CHECK_DEV((unsigned)f2i(T1.w) < (unsigned)S.n_boxes, 4, f2i(T1.w), S.n_boxes);
      if ((ANY || t.careful) && !ref_box_pass(S, pinned(box_origin()), pinned(t.d), f2i(T1.w))) continue;  // invisible to the reference
      t.tbest = ANY ? t.tbest : tt;
      t.rbest = rank;
      t.sbest = first + k;
      if (ANY) { t.node = kSentinel; t.sp = 0; return true; }
    }
    RTP_CHECK_DEV(sp >= 0, 5, sp, 0);
    node = RTP_STK(stack, sp); --sp;
  }
  t.node = node;
  t.sp = sp;
  return node == kSentinel;
}

// Deferred visibility check of a finished closest-hit ray: when the reference's own slab test
// rejects the winner's BoundingBox the ray is re-initialised in careful mode (rare) and the caller
// keeps stepping it.
RTP_HD bool closest_needs_rerun(const DeviceScene &S, Trav &t, int *stack) {
  if (t.careful || t.sbest < 0) return false;
  const int box = f2i(RTP_LDG4(S.tris + 4 * (size_t)t.sbest + 1).w);
  if (ref_box_pass(S, t.o, t.d, box)) return false;
  const v3 o = t.o, d = t.d;
  trav_init(t, o, d, stack);
  t.careful = true;
  RTP_STAT(careful);
  return true;
}

template <bool ANY>
RTP_HD bool traverse(const DeviceScene &S, v3 o, v3 d, v3 bo, float &best_t, int &best_slot) {
  int stack[kStackSize];
  Trav t;
  trav_init(t, o, d, stack);
  do {
    while (!trav_step<ANY ? 1 : 0>(S, t, stack, [&] { return bo; })) {}
  } while (!ANY && closest_needs_rerun(S, t, stack));
  best_t = t.tbest;
  best_slot = t.sbest;
  return t.sbest >= 0;
}

// ---- calcSingleColor after the shadow gate (flyscene.cpp:546-565) -----------------------------
RTP_HD v3 phong(v3 ka, v3 kd, v3 ks, float shininess, v3 normal, v3 L, v3 I, v3 origin, v3 P) {
  v3 lightReflection = vsub(L, vmuls(normal, fmul(2.f, vdot(normal, L))));
  v3 eye = vnormalized(vsub(origin, P));
  v3 ambient = mk3(fmul(I.x, ka.x), fmul(I.y, ka.y), fmul(I.z, ka.z));
  v3 diffuse = vmuls(mk3(fmul(I.x, kd.x), fmul(I.y, kd.y), fmul(I.z, kd.z)), std_max(vdot(L, normal), 0.f));
  float sp = std_max(powf(vdot(lightReflection, eye), shininess), 0.f);
  v3 specular = vmuls(mk3(fmul(I.x, ks.x), fmul(I.y, ks.y), fmul(I.z, ks.z)), sp);
  return vadd(vadd(ambient, diffuse), specular);
}

// Flyscene::reflect on dir.normalized() (flyscene.cpp:361,480-482)
RTP_HD v3 reflect_dir(v3 dir, v3 n) {
  v3 dn = vnormalized(dir);
  return vnormalized(vsub(dn, vmuls(vmuls(n, vdot(dn, n)), 2.f)));
}

// local pixel index -> position inside a 32x32 screen tile: 8x4-pixel warp footprints
RTP_HD void tile_pixel(int p, int &x, int &y) {
  int sub = p >> 5, lane = p & 31;
  x = ((sub & 3) << 3) + (lane & 7);
  y = ((sub >> 2) << 2) + (lane >> 3);
}
// Owner rank of screen tile (tx, ty): diagonal interleave.  A rank's tiles touch at their corners, so neighbouring tiles of
// a rank share hierarchy nodes in its L2.  (Measured against a hashed rotation of the ranks per run of `world` tiles,
// profiles/r02_camera.md section 5: the C3 shares at world 8 got slower, 0.171-0.190 -> 0.178-0.204 ms, and the C2 shares
// no better balanced - their spread is the longest path chain a rank happens to own, not a screen-space pattern.)
RTP_HD int tile_owner(int tx, int ty, int world) { return (tx + ty) % world; }

}  // namespace rtp
