// scene_host.cpp - see scene_host.h.  Compiled with -ffp-contract=off: world vertices, unit
// normals and plane constants must carry exactly the bits the reference recomputes per test.
#include "scene_host.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <chrono>
#include <cstdio>
#include <cstring>

#include "parallel_for.h"
#include "ref_boxes.h"

namespace rtp {

static inline float4 f4(v3 a, float w) { return make_float4(a.x, a.y, a.z, w); }
static inline v3 ld3(const float *p) { return mk3(p[0], p[1], p[2]); }

std::string flatten_scene(const rtp_mesh &m, HostScene &out, int n_threads, bool with_bvh) {
  if (m.n_verts <= 0 || m.n_faces <= 0 || !m.verts || !m.normals || !m.faces || !m.face_mat)
    return "rtp_mesh: verts/normals/faces/face_mat must be non-empty";
  if (m.n_mats > 0 && !m.materials) return "rtp_mesh: n_mats > 0 but materials is NULL";
  if (m.n_mats >= (int)kTriMatMask) return "rtp_mesh: too many materials";
  if (m.n_faces >= (1 << 28)) return "rtp_mesh: too many faces";
  for (int i = 0; i < 3 * m.n_faces; ++i)
    if (m.faces[i] < 0 || m.faces[i] >= m.n_verts) return "rtp_mesh: face vertex id out of range";
  for (int i = 0; i < m.n_faces; ++i)
    if (m.face_mat[i] < -1 || m.face_mat[i] >= m.n_mats) return "rtp_mesh: material id out of range";

  const bool timing = std::getenv("RTP_BUILD_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[flatten_scene] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };
  // ---- the reference's BoundingBox list (given, or built by the reference's splitter) ----
  if (m.n_boxes > 0) {
    if (!m.boxes || !m.box_off || !m.box_faces) return "rtp_mesh: n_boxes > 0 but box arrays are NULL";
    out.box_bounds.assign(m.boxes, m.boxes + 6 * (size_t)m.n_boxes);
    out.box_off.assign(m.box_off, m.box_off + m.n_boxes + 1);
    if (out.box_off[0] != 0) return "rtp_mesh: box_off[0] must be 0";
    for (int b = 0; b < m.n_boxes; ++b)
      if (out.box_off[b + 1] < out.box_off[b]) return "rtp_mesh: box_off must be non-decreasing";
    out.box_faces.assign(m.box_faces, m.box_faces + out.box_off[m.n_boxes]);
    for (int32_t f : out.box_faces)
      if (f < 0 || f >= m.n_faces) return "rtp_mesh: box face id out of range";
  } else {
    std::vector<RefBox> rb;
    build_reference_boxes(m.verts, m.faces, m.n_faces, m.min_faces > 0 ? m.min_faces : 300,
                          m.max_boxes > 0 ? m.max_boxes : INT32_MAX, rb, n_threads);
    out.box_bounds.clear(); out.box_off.assign(1, 0); out.box_faces.clear();
    for (const RefBox &b : rb) {
      for (int k = 0; k < 3; ++k) out.box_bounds.push_back(b.low[k]);
      for (int k = 0; k < 3; ++k) out.box_bounds.push_back(b.high[k]);
      out.box_faces.insert(out.box_faces.end(), b.faces.begin(), b.faces.end());
      out.box_off.push_back((int32_t)out.box_faces.size());
    }
  }
  lap("reference BoundingBox list");
  out.n_boxes = (int)out.box_off.size() - 1;
  out.boxes.resize(2 * (size_t)out.n_boxes);
  for (int b = 0; b < out.n_boxes; ++b) {
    const float *p = &out.box_bounds[6 * (size_t)b];
    out.boxes[2 * (size_t)b] = make_float4(p[0], p[1], p[2], 0.f);
    out.boxes[2 * (size_t)b + 1] = make_float4(p[3], p[4], p[5], 0.f);
  }

  // ---- materials: entry 0 = the Flyscene member defaults (flyscene.hpp:179-182), used for -1 ----
  out.materials.resize(3 * (size_t)(m.n_mats + 1));
  out.materials[0] = make_float4((float)0.2, (float)0.2, (float)0.2, 0.f);
  out.materials[1] = make_float4((float)0.9, (float)0.9, 0.f, 0.f);
  out.materials[2] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = 0; i < m.n_mats; ++i) {
    const float *p = m.materials + 12 * (size_t)i;
    out.materials[3 * (size_t)(i + 1)] = make_float4(p[0], p[1], p[2], p[9]);
    out.materials[3 * (size_t)(i + 1) + 1] = make_float4(p[3], p[4], p[5], 0.f);
    out.materials[3 * (size_t)(i + 1) + 2] = make_float4(p[6], p[7], p[8], 0.f);
  }
  affine_inverse(m.shape, out.minv);  // flyscene.cpp:486

  // ---- per visible face (rank order = the reference's visitation order) ----
  const int n = (int)out.box_faces.size();
  if (n == 0) return "rtp_mesh: the BoundingBox list holds no faces";
  out.n_tris = n;
  BigVec<float4> T((size_t)n * 4), N((size_t)n * 3);
  std::vector<Aabb> bounds((size_t)n);
  std::vector<int32_t> box_of_rank((size_t)n);
  parallel_for((size_t)out.n_boxes, n_threads, 64, [&](size_t b0, size_t b1, int) {
    for (size_t b = b0; b < b1; ++b)
      for (int r = out.box_off[b]; r < out.box_off[b + 1]; ++r) box_of_rank[(size_t)r] = (int32_t)b;
  });
  // faces are independent; each thread keeps its own extent / counters, merged below
  struct Partial {
    float maxabs = 1.f, lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    int flagged = 0, degenerate = 0;
  };
  std::vector<Partial> part((size_t)host_threads(n_threads));
  parallel_for((size_t)n, n_threads, 1 << 14, [&](size_t r0, size_t r1, int tid) {
    Partial &acc = part[(size_t)tid];
    for (size_t r = r0; r < r1; ++r) {
      const int f = out.box_faces[r];
      const int32_t *id = m.faces + 3 * (size_t)f;
      // world vertices: M * v.head<3>() (flyscene.cpp:449,574-576)
      v3 w[3];
      for (int k = 0; k < 3; ++k) w[k] = affine_mul(m.shape, ld3(m.verts + 4 * (size_t)id[k]));
      v3 fn;
      if (m.face_normal) fn = ld3(m.face_normal + 3 * (size_t)f);
      else {  // Mesh::createFaces, mesh.hpp:475-477 (object-space, normalised edges)
        v3 a = ld3(m.verts + 4 * (size_t)id[0]), b = ld3(m.verts + 4 * (size_t)id[1]), c = ld3(m.verts + 4 * (size_t)id[2]);
        v3 e1 = vnormalized_dyn(vsub(c, a));
        v3 e0 = vnormalized_dyn(vsub(b, a));
        fn = vnormalized(vcross(e0, e1));
      }
      v3 nhat = vnormalized(fn);          // face.normal.normalized(), flyscene.cpp:450,577
      float dP = vdot(nhat, w[0]);        // :459
      v3 vn[3];
      for (int k = 0; k < 3; ++k) vn[k] = vnormalized(ld3(m.normals + 3 * (size_t)id[k]));  // :599
      unsigned flags = 0;
      bool unit = true;
      for (int k = 0; k < 3; ++k) {
        float q = vdot(vn[k], vn[k]);
        if (!(q > 0.5f && q < 1.5f)) unit = false;
      }
      if (!unit || !(vdot(vn[0], vn[1]) >= -0.49f) || !(vdot(vn[1], vn[2]) >= -0.49f) || !(vdot(vn[0], vn[2]) >= -0.49f)) {
        flags |= kTriFlagNormalCheck;
        acc.flagged++;
      }
      float nq = vdot(nhat, nhat);
      if (!(nq > 0.5f && nq < 1.5f)) acc.degenerate++;
      unsigned mat = (unsigned)(m.face_mat[f] + 1);
      T[4 * r + 0] = f4(w[0], i2f((int)r));
      T[4 * r + 1] = f4(w[1], i2f(box_of_rank[r]));
      T[4 * r + 2] = f4(w[2], i2f((int)(mat | flags)));
      T[4 * r + 3] = f4(nhat, dP);
      for (int k = 0; k < 3; ++k) N[3 * r + k] = f4(vn[k], 0.f);
      Aabb &bb = bounds[r];
      bool finite = true;
      for (int a = 0; a < 3; ++a) {
        float c0 = (&w[0].x)[a], c1 = (&w[1].x)[a], c2 = (&w[2].x)[a];
        if (!std::isfinite(c0) || !std::isfinite(c1) || !std::isfinite(c2)) finite = false;
        bb.lo[a] = std::min(c0, std::min(c1, c2));
        bb.hi[a] = std::max(c0, std::max(c1, c2));
        if (finite) acc.maxabs = std::max(acc.maxabs, std::max(std::fabs(bb.lo[a]), std::fabs(bb.hi[a])));
      }
      if (finite)
        for (int a = 0; a < 3; ++a) {
          acc.lo[a] = std::min(acc.lo[a], bb.lo[a]);
          acc.hi[a] = std::max(acc.hi[a], bb.hi[a]);
        }
      if (!finite)
        for (int a = 0; a < 3; ++a) { bb.lo[a] = -FLT_MAX; bb.hi[a] = FLT_MAX; }
    }
  });
  float maxabs = 1.f;
  for (int a = 0; a < 3; ++a) { out.scene_lo[a] = FLT_MAX; out.scene_hi[a] = -FLT_MAX; }
  out.n_flagged = out.n_degenerate = 0;
  for (const Partial &acc : part) {
    maxabs = std::max(maxabs, acc.maxabs);
    for (int a = 0; a < 3; ++a) {
      out.scene_lo[a] = std::min(out.scene_lo[a], acc.lo[a]);
      out.scene_hi[a] = std::max(out.scene_hi[a], acc.hi[a]);
    }
    out.n_flagged += acc.flagged;
    out.n_degenerate += acc.degenerate;
  }
  lap("per-face flatten");
  // Conservative padding: the exact test accepts points P = o + t*d that carry a few ulps of
  // rounding at |o| + t, so a face may be "hit" by a ray that misses its tight box by ~1e-6.
  float pad_rel = 3e-5f;
  if (const char *e = std::getenv("RTP_BVH_PAD")) pad_rel = (float)std::atof(e);
  out.pad = pad_rel * maxabs;
  out.maxabs = maxabs;
  if (!with_bvh) {  // rank order; the device builder sorts, fits and permutes (bvh_build_gpu.cu)
    out.n_nodes = 0;
    out.bvh_depth = 0;
    out.nodes.clear();
    out.tris.swap(T);
    out.vnorm.swap(N);
    out.face_of.assign(out.box_faces.begin(), out.box_faces.end());
    return "";
  }
  for (Aabb &bb : bounds)
    for (int a = 0; a < 3; ++a)
      if (bb.hi[a] < FLT_MAX) { bb.lo[a] -= out.pad; bb.hi[a] += out.pad; }

  Bvh bvh;
  int max_leaf = 4;
  if (const char *e = std::getenv("RTP_BVH_MAX_LEAF")) max_leaf = std::atoi(e);
  build_bvh(bounds, max_leaf, n_threads, bvh);
  lap("binned-SAH BVH");
  if (bvh.max_depth >= kStackSize - 2) return "BVH deeper than the traversal stack";
  out.bvh_depth = bvh.max_depth;
  out.n_nodes = (int)bvh.nodes.size();
  out.nodes.resize(4 * bvh.nodes.size());
  static_assert(sizeof(BvhNode) == 64, "BvhNode must be 64 bytes");
  std::memcpy(out.nodes.data(), bvh.nodes.data(), bvh.nodes.size() * sizeof(BvhNode));
  out.tris.resize((size_t)n * 4);
  out.vnorm.resize((size_t)n * 3);
  out.face_of.resize((size_t)n);
  parallel_for((size_t)n, n_threads, 1 << 14, [&](size_t s0, size_t s1, int) {
    for (size_t s = s0; s < s1; ++s) {
      const size_t r = (size_t)bvh.order[s];
      for (int k = 0; k < 4; ++k) out.tris[4 * s + k] = T[4 * r + k];
      for (int k = 0; k < 3; ++k) out.vnorm[3 * s + k] = N[3 * r + k];
      out.face_of[s] = out.box_faces[r];
    }
  });
  lap("leaf-order permutation");
  return "";
}

void derive_camera(const rtp_camera &c, DeviceCamera &out) {
  float li[9];  // camera.hpp:115-118: view.linear().inverse() * (-view.translation())
  inverse3(c.view, 4, li, 3);
  const float nt[3] = {-c.view[3], -c.view[7], -c.view[11]};
  for (int r = 0; r < 3; ++r) out.eye[r] = li[3 * r] * nt[0] + (li[3 * r + 1] * nt[1] + li[3 * r + 2] * nt[2]);
  affine_inverse(c.view, out.inv_view);
  for (int k = 0; k < 4; ++k) out.vp[k] = c.viewport[k];
  // camera.hpp:263-266 and :165-167
  float pscale = (float)((double)1.0f / std::tan((double)(c.fovy / 2.0f) * (M_PI / (double)180.0f)));
  float scale = (float)(1.0 / (double)pscale);
  out.sx = c.aspect * scale;
  out.sy = scale;
}

void pack_lights(const rtp_lights &l, std::vector<float4> &out) {
  out.clear();
  for (int i = 0; i < l.n_point; ++i) {
    const float *p = l.point + 6 * (size_t)i;
    out.push_back(make_float4(p[0], p[1], p[2], 0.f));
    out.push_back(make_float4(p[3], p[4], p[5], 0.f));
  }
  for (int i = 0; i < l.n_dir; ++i) {
    const float *p = l.dir + 6 * (size_t)i;
    out.push_back(make_float4(p[0], p[1], p[2], 1.f));
    out.push_back(make_float4(p[3], p[4], p[5], 0.f));
  }
}

void tile_tables(int width, int height, int rank, int world, std::vector<int32_t> &tiles_of_rank,
                 std::vector<int32_t> &local_of_tile) {
  const int tx = (width + kTileDim - 1) / kTileDim, ty = (height + kTileDim - 1) / kTileDim;
  std::vector<int32_t> count((size_t)world, 0);
  local_of_tile.assign((size_t)tx * ty, 0);
  tiles_of_rank.clear();
  for (int y = 0; y < ty; ++y)
    for (int x = 0; x < tx; ++x) {
      int owner = tile_owner(x, y, world);
      local_of_tile[(size_t)y * tx + x] = count[(size_t)owner]++;
      if (owner == rank) tiles_of_rank.push_back(y * tx + x);
    }
}

}  // namespace rtp
