// TEST INFRASTRUCTURE ONLY - never linked into the product library.
//
// Runs the product's __host__ __device__ code (flatten_scene, the BVH2 traversal, the exact
// triangle / BoundingBox tests, interpolate_normal, phong, reflect_dir from
// ray-tracing-project_b200/csrc/rtp_device.h) serially on the CPU, pixel by pixel, following the
// same per-depth sequence the CUDA kernels in rtp_kernels.cu use.  Its only purpose is to let
// `pytest -m "not gpu"` compare that shared code with the oracle on a machine without a GPU, so
// that GPU minutes are spent on measuring rather than on debugging host logic.
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/rtp.h"
#include "../../ray-tracing-project_b200/csrc/scene_host.h"

using namespace rtp;
namespace rtp { thread_local TravStats g_trav_stats; }

extern "C" {

struct hostsim_scene {
  HostScene h;
  DeviceScene d;
  std::vector<float4> lights;
};

const char *hostsim_create(const rtp_mesh *m, hostsim_scene **out) {
  static std::string err;
  hostsim_scene *s = new hostsim_scene();
  const char *nt = std::getenv("HOSTSIM_THREADS");  // the build must not depend on the thread count
  err = flatten_scene(*m, s->h, nt ? std::atoi(nt) : 1);
  if (!err.empty()) { delete s; return err.c_str(); }
  DeviceScene &d = s->d;
  d.nodes = s->h.nodes.data(); d.tris = s->h.tris.data(); d.vnorm = s->h.vnorm.data();
  d.boxes = s->h.boxes.data(); d.face_of = s->h.face_of.data(); d.materials = s->h.materials.data();
  d.lights = nullptr; d.n_lights = 0;
  std::memcpy(d.minv, s->h.minv, sizeof d.minv);
  d.n_tris = s->h.n_tris; d.n_nodes = s->h.n_nodes; d.n_boxes = s->h.n_boxes;
  *out = s;
  return nullptr;
}
void hostsim_destroy(hostsim_scene *s) { delete s; }
void hostsim_info(hostsim_scene *s, int *out /* n_tris n_nodes n_boxes depth flagged degenerate */) {
  out[0] = s->h.n_tris; out[1] = s->h.n_nodes; out[2] = s->h.n_boxes; out[3] = s->h.bvh_depth;
  out[4] = s->h.n_flagged; out[5] = s->h.n_degenerate;
}
void hostsim_get_boxes(hostsim_scene *s, float *b, int *off, int *faces) {
  std::memcpy(b, s->h.box_bounds.data(), s->h.box_bounds.size() * 4);
  std::memcpy(off, s->h.box_off.data(), s->h.box_off.size() * 4);
  std::memcpy(faces, s->h.box_faces.data(), s->h.box_faces.size() * 4);
}
void hostsim_get_layout(hostsim_scene *s, float *nodes /* 16 per node */, float *tris /* 16 per tri */, int *face_of) {
  std::memcpy(nodes, s->h.nodes.data(), s->h.nodes.size() * sizeof(float4));
  std::memcpy(tris, s->h.tris.data(), s->h.tris.size() * sizeof(float4));
  std::memcpy(face_of, s->h.face_of.data(), s->h.face_of.size() * 4);
}
void hostsim_set_lights(hostsim_scene *s, const rtp_lights *l) {
  pack_lights(*l, s->lights);
  s->d.lights = s->lights.data();
  s->d.n_lights = l->n_point + l->n_dir;
}

// pixels: n x 2 (i, j); outputs as oracle_render. counts: primary shadow reflection hits
void hostsim_render(hostsim_scene *s, const rtp_camera *c, const int *pixels, long long n, int max_depth,
                    const float *bg, float *rgb, int *hit, float *tprim, float *dirs, long long *counts) {
  DeviceCamera cam;
  derive_camera(*c, cam);
  const DeviceScene &S = s->d;
  long long cnt[4] = {0, 0, 0, 0};
  for (long long k = 0; k < n; ++k) {
    v3 o = mk3(cam.eye[0], cam.eye[1], cam.eye[2]);
    v3 d = primary_dir(cam, pixels[2 * k], pixels[2 * k + 1]);
    if (dirs) { dirs[3 * k] = d.x; dirs[3 * k + 1] = d.y; dirs[3 * k + 2] = d.z; }
    cnt[0]++;
    v3 direct[kMaxDepth];
    v3 ksdeep = mk3(0, 0, 0);
    int chain = 0, face0 = -1;
    float t0 = INFINITY;
    for (int depth = 0; depth < max_depth; ++depth) {
      float t; int slot;
      traverse<false>(S, o, d, o, t, slot);            // k_closest
      if (depth == 0) { face0 = slot >= 0 ? S.face_of[slot] : -1; t0 = t; }
      if (slot < 0) break;
      cnt[3]++;
      v3 P = vadd(o, vmuls(d, t));                       // k_shade
      v3 normal = interpolate_normal_slot(S, slot, P);
      unsigned mat = (unsigned)f2i(S.tris[4 * (size_t)slot + 2].w) & kTriMatMask;
      float4 m0 = S.materials[3 * mat], m1 = S.materials[3 * mat + 1], m2 = S.materials[3 * mat + 2];
      v3 sum = mk3(0, 0, 0);
      for (int li = 0; li < S.n_lights; ++li) {
        float4 l0 = S.lights[2 * li], l1 = S.lights[2 * li + 1];
        v3 L = xyz(l0);
        if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));
        v3 inter = vadd(P, vmuls(L, (float)0.003));
        float ts; int ss;
        cnt[1]++;
        if (traverse<true>(S, inter, L, P, ts, ss)) continue;
        sum = vadd(sum, phong(xyz(m0), xyz(m1), xyz(m2), m0.w, normal, L, xyz(l1), o, P));
      }
      direct[depth] = mk3(clamp01(sum.x), clamp01(sum.y), clamp01(sum.z));
      ksdeep = xyz(m2);
      chain = depth + 1;
      if (depth + 1 < max_depth) {
        cnt[2]++;
        v3 r = reflect_dir(d, normal);
        o = vadd(P, vmuls(r, (float)0.001));
        d = r;
      }
    }
    v3 c;                                                // k_resolve
    if (chain == 0) c = mk3(bg[0], bg[1], bg[2]);
    else {
      v3 dc = direct[chain - 1];
      c = mk3(clamp01(fadd(dc.x, fmul(0.f, ksdeep.x))), clamp01(fadd(dc.y, fmul(0.f, ksdeep.y))),
              clamp01(fadd(dc.z, fmul(0.f, ksdeep.z))));
      for (int dd = chain - 2; dd >= 0; --dd) {
        dc = direct[dd];
        c = mk3(clamp01(fadd(dc.x, fmul(c.x, ksdeep.x))), clamp01(fadd(dc.y, fmul(c.y, ksdeep.y))),
                clamp01(fadd(dc.z, fmul(c.z, ksdeep.z))));
      }
    }
    if (rgb) { rgb[3 * k] = c.x; rgb[3 * k + 1] = c.y; rgb[3 * k + 2] = c.z; }
    if (hit) hit[k] = face0;
    if (tprim) tprim[k] = t0;
  }
  if (counts) std::memcpy(counts, cnt, sizeof cnt);
}

void hostsim_rays(hostsim_scene *s, const float *rays, long long n, int anyhit, int *out, float *t) {
  for (long long k = 0; k < n; ++k) {
    v3 o = mk3(rays[6 * k], rays[6 * k + 1], rays[6 * k + 2]), d = mk3(rays[6 * k + 3], rays[6 * k + 4], rays[6 * k + 5]);
    float tt; int slot;
    if (anyhit) {
      v3 inter = vadd(o, vmuls(d, (float)0.003));
      out[k] = traverse<true>(s->d, inter, d, o, tt, slot) ? 1 : 0;
    } else {
      traverse<false>(s->d, o, d, o, tt, slot);
      out[k] = slot >= 0 ? s->d.face_of[slot] : -1;
      if (t) t[k] = tt;
    }
  }
}

// per primary ray: nodes visited, leaves visited, triangles tested, careful re-runs (out: n x 4)
void hostsim_trav_stats(hostsim_scene *s, const rtp_camera *c, const int *pixels, long long n, int *out) {
  DeviceCamera cam;
  derive_camera(*c, cam);
  for (long long k = 0; k < n; ++k) {
    v3 o = mk3(cam.eye[0], cam.eye[1], cam.eye[2]);
    v3 d = primary_dir(cam, pixels[2 * k], pixels[2 * k + 1]);
    g_trav_stats = TravStats{0, 0, 0, 0};
    float t; int slot;
    traverse<false>(s->d, o, d, o, t, slot);
    out[4 * k] = (int)g_trav_stats.nodes; out[4 * k + 1] = (int)g_trav_stats.leaves;
    out[4 * k + 2] = (int)g_trav_stats.tris; out[4 * k + 3] = (int)g_trav_stats.careful;
  }
}
// same for the shadow ray of light 0 from each primary hit (out: n x 4, last column: 1 = occluded, -1 = no hit)
void hostsim_shadow_stats(hostsim_scene *s, const rtp_camera *c, const int *pixels, long long n, int *out) {
  DeviceCamera cam;
  derive_camera(*c, cam);
  const DeviceScene &S = s->d;
  for (long long k = 0; k < n; ++k) {
    v3 o = mk3(cam.eye[0], cam.eye[1], cam.eye[2]);
    v3 d = primary_dir(cam, pixels[2 * k], pixels[2 * k + 1]);
    float t; int slot;
    traverse<false>(S, o, d, o, t, slot);
    out[4 * k] = out[4 * k + 1] = out[4 * k + 2] = 0; out[4 * k + 3] = -1;
    if (slot < 0 || S.n_lights < 1) continue;
    v3 P = vadd(o, vmuls(d, t));
    float4 l0 = S.lights[0];
    v3 L = xyz(l0);
    if (l0.w == 0.f) L = vneg(vnormalized(vsub(P, L)));
    g_trav_stats = TravStats{0, 0, 0, 0};
    float ts; int ss;
    bool occ = traverse<true>(S, vadd(P, vmuls(L, (float)0.003)), L, P, ts, ss);
    out[4 * k] = (int)g_trav_stats.nodes; out[4 * k + 1] = (int)g_trav_stats.leaves;
    out[4 * k + 2] = (int)g_trav_stats.tris; out[4 * k + 3] = occ ? 1 : 0;
  }
}

void hostsim_tile_tables(int w, int h, int rank, int world, int *tiles, int *n_tiles, int *local_of_tile) {
  std::vector<int32_t> a, b;
  tile_tables(w, h, rank, world, a, b);
  *n_tiles = (int)a.size();
  std::memcpy(tiles, a.data(), a.size() * 4);
  std::memcpy(local_of_tile, b.data(), b.size() * 4);
}
}
