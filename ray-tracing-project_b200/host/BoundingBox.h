// BoundingBox.h - mirror of the reference's acceleration-structure class (src/BoundingBox.h:10-67):
// same statics (`boxes`, `mesh`), same fields (low / high / shape / faces / failed[3]) and the same
// member functions, so Flyscene::generateBoundingBoxes reads like flyscene.cpp:399-428.  The list
// is part of the reference's observable behaviour (which faces a ray may see, first-visited tie
// order - SURVEY.md Q11), so it is handed to the CUDA library verbatim (rtp_mesh::boxes); the GPU
// traverses its own linear BVH built over the same triangles.  The split arithmetic lives once,
// in csrc/ref_boxes.cpp, shared with the library's own builder.  GL debug drawing
// (generateShape / render) is declared and empty: no graphics stack on this box.
#pragma once
#include <vector>

#include "tucano_lite.hpp"

class BoundingBox {
 public:
  static std::vector<BoundingBox *> boxes;
  static Tucano::Mesh *mesh;
  bool failed[3] = {false, false, false};

  // raw coordinates (object space)
  Eigen::Vector3f low;
  Eigen::Vector3f high;
  Eigen::Vector3f shape;
  float width = 0.f;
  float height = 0.f;
  float depth = 0.f;
  Eigen::Vector3f color = Eigen::Vector3f(1, 1, 1);
  // all faces that should be inside this box
  std::vector<Tucano::Face *> faces;

  // remember: the box registers itself in `boxes` and is freed by deconstruct()
  BoundingBox(bool remember);
  ~BoundingBox();
  static void deconstruct();

  void reshape();
  bool hasFace(Tucano::Face &face);
  bool hasVertex(Eigen::Vector4f &vertex);
  void fitMesh();
  void fitFaces();
  std::vector<Tucano::Face *> outsideFaces();
  BoundingBox *splitBox();
  float averageVertexCoord(int axis);
  void setRandomColor();
  int getNumberOfFaces();
  void generateShape() {}                                  // GL debug box: unavailable on this box
  void render(Tucano::Flycamera &, Tucano::Camera &) {}    // GL debug box: unavailable on this box

  // Flat copies of the list for the C ABI (rtp_mesh::boxes / box_off / box_faces).
  static void exportBoxes(std::vector<float> &bounds, std::vector<int> &box_off, std::vector<int> &box_faces);
};
