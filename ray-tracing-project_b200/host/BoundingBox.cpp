// BoundingBox.cpp - see BoundingBox.h.  Each member maps the object onto an rtp::RefBox, runs the
// shared step in csrc/ref_boxes.cpp (the restatement of src/BoundingBox.cpp:18-161) and maps back.
#include "BoundingBox.h"

#include <cassert>
#include <cstdlib>

#include "../csrc/ref_boxes.h"

std::vector<BoundingBox *> BoundingBox::boxes;
Tucano::Mesh *BoundingBox::mesh = nullptr;

namespace {

// Flat views of BoundingBox::mesh (V x 4 floats, F x 3 ints), rebuilt when the mesh changes.
struct FlatMesh {
  const Tucano::Mesh *of = nullptr;
  int n_verts = 0, n_faces = 0;
  std::vector<float> verts;
  std::vector<int32_t> faces;
} g_flat;

const FlatMesh &flat() {
  Tucano::Mesh *m = BoundingBox::mesh;
  assert(m && "BoundingBox::mesh must be set before any box is used");
  if (g_flat.of != m || g_flat.n_verts != m->getNumberOfVertices() || g_flat.n_faces != m->getNumberOfFaces()) {
    g_flat.of = m;
    g_flat.n_verts = m->getNumberOfVertices();
    g_flat.n_faces = m->getNumberOfFaces();
    g_flat.verts.resize(4 * (size_t)g_flat.n_verts);
    g_flat.faces.resize(3 * (size_t)g_flat.n_faces);
    for (int i = 0; i < g_flat.n_verts; ++i) {
      const Eigen::Vector4f v = m->getVertex(i);
      for (int c = 0; c < 4; ++c) g_flat.verts[4 * (size_t)i + c] = v[c];
    }
    for (int i = 0; i < g_flat.n_faces; ++i)
      for (int c = 0; c < 3; ++c) g_flat.faces[3 * (size_t)i + c] = (int32_t)m->getFace(i).vertex_ids[(size_t)c];
  }
  return g_flat;
}

int32_t face_id(const Tucano::Face *f) { return (int32_t)(f - &BoundingBox::mesh->getFace(0)); }

void to_ref(const BoundingBox &b, rtp::RefBox &r) {
  for (int k = 0; k < 3; ++k) {
    r.low[k] = b.low[k]; r.high[k] = b.high[k]; r.shape[k] = b.shape[k]; r.failed[k] = b.failed[k];
  }
  r.faces.resize(b.faces.size());
  for (size_t i = 0; i < b.faces.size(); ++i) r.faces[i] = face_id(b.faces[i]);
}

void from_ref(const rtp::RefBox &r, BoundingBox &b) {
  for (int k = 0; k < 3; ++k) {
    b.low[k] = r.low[k]; b.high[k] = r.high[k]; b.shape[k] = r.shape[k]; b.failed[k] = r.failed[k];
  }
  b.width = r.shape[0]; b.height = r.shape[1]; b.depth = r.shape[2];
  b.faces.resize(r.faces.size());
  for (size_t i = 0; i < r.faces.size(); ++i) b.faces[i] = &BoundingBox::mesh->getFace(r.faces[i]);
}

}  // namespace

BoundingBox::BoundingBox(bool remember) {  // BoundingBox.cpp:7-9
  if (remember) boxes.push_back(this);
}

BoundingBox::~BoundingBox() {}

void BoundingBox::deconstruct() {  // BoundingBox.cpp:11-16
  for (BoundingBox *b : boxes) delete b;
  boxes.clear();
}

void BoundingBox::reshape() {  // BoundingBox.cpp:18-24
  shape = high - low;
  width = shape[0]; height = shape[1]; depth = shape[2];
}

bool BoundingBox::hasVertex(Eigen::Vector4f &v) {  // BoundingBox.cpp:34-39, bounds inclusive
  assert(v[3] == 1);
  return v[0] >= low[0] && v[0] <= high[0] && v[1] >= low[1] && v[1] <= high[1] && v[2] >= low[2] && v[2] <= high[2];
}

bool BoundingBox::hasFace(Tucano::Face &face) {  // BoundingBox.cpp:26-32
  for (int k = 0; k < 3; ++k) {
    Eigen::Vector4f v = mesh->getVertex((int)face.vertex_ids[(size_t)k]);
    if (!hasVertex(v)) return false;
  }
  return true;
}

void BoundingBox::fitMesh() {  // BoundingBox.cpp:41-46
  faces.clear();
  for (int i = 0; i < mesh->getNumberOfFaces(); ++i) faces.push_back(&mesh->getFace(i));
  fitFaces();
}

void BoundingBox::fitFaces() {  // BoundingBox.cpp:48-90
  const FlatMesh &fm = flat();
  rtp::RefBox r;
  to_ref(*this, r);
  rtp::refbox_fit_faces(r, fm.verts.data(), fm.faces.data());
  from_ref(r, *this);
}

std::vector<Tucano::Face *> BoundingBox::outsideFaces() {  // BoundingBox.cpp:92-107
  std::vector<Tucano::Face *> inside, outside;
  for (Tucano::Face *f : faces) (hasFace(*f) ? inside : outside).push_back(f);
  faces.swap(inside);
  return outside;
}

float BoundingBox::averageVertexCoord(int axis) {  // BoundingBox.cpp:151-161
  const FlatMesh &fm = flat();
  rtp::RefBox r;
  to_ref(*this, r);
  return rtp::refbox_average_vertex_coord(r, fm.verts.data(), fm.faces.data(), axis);
}

BoundingBox *BoundingBox::splitBox() {  // BoundingBox.cpp:109-149
  const FlatMesh &fm = flat();
  rtp::RefBox r, created;
  to_ref(*this, r);
  const int rc = rtp::refbox_split(r, created, fm.verts.data(), fm.faces.data());
  if (rc == 0) return nullptr;  // every axis failed
  from_ref(r, *this);
  if (rc == 1) return this;     // this axis failed: the caller retries (flyscene.cpp:413)
  BoundingBox *nb = new BoundingBox(true);
  from_ref(created, *nb);
  return nb;
}

void BoundingBox::setRandomColor() {  // BoundingBox.cpp:163-169 (debug colouring only)
  color = Eigen::Vector3f((float)rand() / (float)RAND_MAX, (float)rand() / (float)RAND_MAX, (float)rand() / (float)RAND_MAX);
}

int BoundingBox::getNumberOfFaces() { return (int)faces.size(); }

void BoundingBox::exportBoxes(std::vector<float> &bounds, std::vector<int> &box_off, std::vector<int> &box_faces) {
  bounds.clear(); box_faces.clear();
  box_off.assign(1, 0);
  for (BoundingBox *b : boxes) {
    for (int k = 0; k < 3; ++k) bounds.push_back(b->low[k]);
    for (int k = 0; k < 3; ++k) bounds.push_back(b->high[k]);
    for (Tucano::Face *f : b->faces) box_faces.push_back(face_id(f));
    box_off.push_back((int)box_faces.size());
  }
}
