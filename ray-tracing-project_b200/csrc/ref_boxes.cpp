// ref_boxes.cpp - see ref_boxes.h.  Compiled with -ffp-contract=off: the float comparisons and
// the FP32 running mean must round exactly like the reference's.
#include "ref_boxes.h"

#include "parallel_for.h"

#include <cmath>
#include <cstddef>
#include <deque>
#include <memory>

namespace rtp {

static void reshape(RefBox &b) {  // BoundingBox.cpp:18-24
  for (int k = 0; k < 3; ++k) b.shape[k] = b.high[k] - b.low[k];
}

bool refbox_has_face(const RefBox &b, const float *verts, const int32_t *f) {  // :26-39
  for (int k = 0; k < 3; ++k) {
    const float *v = verts + 4 * (size_t)f[k];
    if (!(v[0] >= b.low[0] && v[0] <= b.high[0] && v[1] >= b.low[1] && v[1] <= b.high[1] &&
          v[2] >= b.low[2] && v[2] <= b.high[2]))
      return false;
  }
  return true;
}

// Boxes with at least this many faces are also cut into chunks inside one split (the first
// passes of generateBoundingBoxes hold one, two, four ... boxes, far fewer than there are cores).
constexpr size_t kChunkedSplitMin = 1 << 16;

void refbox_fit_faces(RefBox &b, const float *verts, const int32_t *faces, int n_threads) {  // :48-90
  if (b.faces.empty()) return;
  const float *t = verts + 4 * (size_t)faces[3 * (size_t)b.faces[0]];
  float mn[3] = {t[0], t[1], t[2]}, mx[3] = {t[0], t[1], t[2]};
  auto sweep = [&](size_t lo, size_t hi, float *mn, float *mx) {
    for (size_t i = lo; i < hi; ++i)
      for (int j = 0; j < 3; ++j) {
        const float *v = verts + 4 * (size_t)faces[3 * (size_t)b.faces[i] + j];
        for (int k = 0; k < 3; ++k) {
          if (v[k] < mn[k]) mn[k] = v[k];
          else if (v[k] > mx[k]) mx[k] = v[k];
        }
      }
  };
  const size_t n = b.faces.size();
  if (n_threads > 1 && n >= kChunkedSplitMin) {
    // The reference's loop keeps, per coordinate, the first-seen smallest and largest value
    // (strict compares; NaN never wins; a NaN start value never loses).  Chunk results merged in
    // chunk order with the same strict compares give the same bits, signed zeros included.
    const size_t chunks = (size_t)n_threads * 4, grain = (n + chunks - 1) / chunks;
    struct MinMax { float mn[3], mx[3]; };
    std::vector<MinMax> part((n + grain - 1) / grain);
    parallel_for(n, n_threads, grain, [&](size_t lo, size_t hi, int) {
      MinMax &p = part[lo / grain];
      for (int k = 0; k < 3; ++k) { p.mn[k] = INFINITY; p.mx[k] = -INFINITY; }
      for (size_t i = lo; i < hi; ++i)
        for (int j = 0; j < 3; ++j) {
          const float *v = verts + 4 * (size_t)faces[3 * (size_t)b.faces[i] + j];
          for (int k = 0; k < 3; ++k) {
            if (v[k] < p.mn[k]) p.mn[k] = v[k];
            if (v[k] > p.mx[k]) p.mx[k] = v[k];
          }
        }
    });
    for (const MinMax &p : part)
      for (int k = 0; k < 3; ++k) {
        if (p.mn[k] < mn[k]) mn[k] = p.mn[k];
        if (p.mx[k] > mx[k]) mx[k] = p.mx[k];
      }
  } else {
    sweep(0, n, mn, mx);
  }
  for (int k = 0; k < 3; ++k) { b.low[k] = mn[k]; b.high[k] = mx[k]; }
  reshape(b);
}

float refbox_average_vertex_coord(const RefBox &b, const float *verts, const int32_t *faces, int axis) {
  float average = 0;  // :151-161: FP32 running sum, one add per vertex use
  for (int32_t f : b.faces) {
    const int32_t *id = faces + 3 * (size_t)f;
    average += verts[4 * (size_t)id[0] + axis];
    average += verts[4 * (size_t)id[1] + axis];
    average += verts[4 * (size_t)id[2] + axis];
  }
  average /= (float)(b.faces.size() * 3);
  return average;
}

int refbox_split(RefBox &b, RefBox &created, const float *verts, const int32_t *faces, int n_threads) {  // :109-149
  float old_low[3] = {b.low[0], b.low[1], b.low[2]}, old_high[3] = {b.high[0], b.high[1], b.high[2]};
  const float width = b.shape[0], height = b.shape[1], depth = b.shape[2];
  int choice;
  if ((width >= height || b.failed[1]) && (width >= depth || b.failed[2]) && !b.failed[0]) choice = 0;
  else if ((height >= width || b.failed[0]) && (height >= depth || b.failed[2]) && !b.failed[1]) choice = 1;
  else if (!(b.failed[0] && b.failed[1] && b.failed[2])) choice = 2;
  else return 0;

  b.high[choice] = refbox_average_vertex_coord(b, verts, faces, choice);
  reshape(b);
  std::vector<int32_t> inside, outside;  // outsideFaces(), :92-107
  const size_t n = b.faces.size();
  if (n_threads > 1 && n >= kChunkedSplitMin) {  // same two lists, built per chunk and joined in order
    const size_t chunks = (size_t)n_threads * 4, grain = (n + chunks - 1) / chunks;
    std::vector<std::vector<int32_t>> in((n + grain - 1) / grain), outv(in.size());
    parallel_for(n, n_threads, grain, [&](size_t lo, size_t hi, int) {
      std::vector<int32_t> &ci = in[lo / grain], &co = outv[lo / grain];
      for (size_t i = lo; i < hi; ++i) {
        const int32_t f = b.faces[i];
        (refbox_has_face(b, verts, faces + 3 * (size_t)f) ? ci : co).push_back(f);
      }
    });
    for (auto &c : in) inside.insert(inside.end(), c.begin(), c.end());
    for (auto &c : outv) outside.insert(outside.end(), c.begin(), c.end());
  } else {
    for (int32_t f : b.faces) (refbox_has_face(b, verts, faces + 3 * (size_t)f) ? inside : outside).push_back(f);
  }
  b.faces.swap(inside);
  if (b.faces.empty() || outside.empty()) {
    if (b.faces.empty()) b.faces.swap(outside);
    b.failed[choice] = true;
    for (int k = 0; k < 3; ++k) { b.low[k] = old_low[k]; b.high[k] = old_high[k]; }
    reshape(b);
    return 1;
  }
  b.failed[0] = b.failed[1] = b.failed[2] = false;
  created = RefBox();
  created.faces.swap(outside);
  refbox_fit_faces(created, verts, faces, n_threads);
  refbox_fit_faces(b, verts, faces, n_threads);
  return 2;
}

void build_reference_boxes(const float *verts, const int32_t *faces, int n_faces, int min_faces,
                           int max_boxes, std::vector<RefBox> &out, int n_threads) {
  std::deque<RefBox> boxes;  // stable references while appending
  boxes.emplace_back();
  boxes[0].faces.resize((size_t)n_faces);  // fitMesh, :41-46
  for (int i = 0; i < n_faces; ++i) boxes[0].faces[(size_t)i] = i;
  refbox_fit_faces(boxes[0], verts, faces);
  bool not_done = true;
  while (not_done && (long long)boxes.size() < (long long)max_boxes) {  // flyscene.cpp:405-418
    // One pass of the reference's loop splits every box that existed when the pass began and
    // appends the new halves behind them.  A split reads and writes its own box only, so the
    // boxes of a pass are split concurrently and the halves appended afterwards in box order -
    // the list (order included) is the one the sequential loop produces.
    const size_t current = boxes.size();
    std::vector<RefBox> created(current);
    std::vector<char> made(current, 0), tried(current, 0);
    const int total = host_threads(n_threads);
    const int inner = current < (size_t)total ? (int)((size_t)total / current) : 1;
    parallel_for(current, n_threads, 1, [&](size_t lo, size_t hi, int) {
      for (size_t i = lo; i < hi; ++i) {
        RefBox &b = boxes[i];
        if ((int)b.faces.size() > min_faces && (!b.failed[0] || !b.failed[1] || !b.failed[2])) {
          int r;
          do r = refbox_split(b, created[i], verts, faces, inner);
          while (r == 1);
          made[i] = r == 2;
          tried[i] = 1;
        }
      }
    });
    not_done = false;
    for (size_t i = 0; i < current; ++i) {
      if (made[i]) boxes.push_back(std::move(created[i]));
      if (tried[i]) not_done = true;
    }
  }
  out.assign(std::make_move_iterator(boxes.begin()), std::make_move_iterator(boxes.end()));
}

}  // namespace rtp
