// ref_boxes.cpp - see ref_boxes.h.  Compiled with -ffp-contract=off: the float comparisons and
// the FP32 running mean must round exactly like the reference's.
#include "ref_boxes.h"

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

void refbox_fit_faces(RefBox &b, const float *verts, const int32_t *faces) {  // :48-90
  if (b.faces.empty()) return;
  const float *t = verts + 4 * (size_t)faces[3 * (size_t)b.faces[0]];
  float mn[3] = {t[0], t[1], t[2]}, mx[3] = {t[0], t[1], t[2]};
  for (int32_t f : b.faces)
    for (int j = 0; j < 3; ++j) {
      const float *v = verts + 4 * (size_t)faces[3 * (size_t)f + j];
      for (int k = 0; k < 3; ++k) {
        if (v[k] < mn[k]) mn[k] = v[k];
        else if (v[k] > mx[k]) mx[k] = v[k];
      }
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

int refbox_split(RefBox &b, RefBox &created, const float *verts, const int32_t *faces) {  // :109-149
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
  for (int32_t f : b.faces) (refbox_has_face(b, verts, faces + 3 * (size_t)f) ? inside : outside).push_back(f);
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
  refbox_fit_faces(created, verts, faces);
  refbox_fit_faces(b, verts, faces);
  return 2;
}

void build_reference_boxes(const float *verts, const int32_t *faces, int n_faces, int min_faces,
                           int max_boxes, std::vector<RefBox> &out) {
  std::deque<RefBox> boxes;  // stable references while appending
  boxes.emplace_back();
  boxes[0].faces.resize((size_t)n_faces);  // fitMesh, :41-46
  for (int i = 0; i < n_faces; ++i) boxes[0].faces[(size_t)i] = i;
  refbox_fit_faces(boxes[0], verts, faces);
  bool not_done = true;
  while (not_done && (long long)boxes.size() < (long long)max_boxes) {  // flyscene.cpp:405-418
    not_done = false;
    const size_t current = boxes.size();
    for (size_t i = 0; i < current; ++i) {
      RefBox &b = boxes[i];
      if ((int)b.faces.size() > min_faces && (!b.failed[0] || !b.failed[1] || !b.failed[2])) {
        int r;
        do {
          RefBox nb;
          r = refbox_split(b, nb, verts, faces);
          if (r == 2) boxes.push_back(std::move(nb));
        } while (r == 1);
        not_done = true;
      }
    }
  }
  out.assign(std::make_move_iterator(boxes.begin()), std::make_move_iterator(boxes.end()));
}

}  // namespace rtp
