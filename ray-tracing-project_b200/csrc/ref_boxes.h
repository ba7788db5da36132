// ref_boxes.h - the reference's BoundingBox list, built with the reference's own splitter
// (src/flyscene.cpp:399-418 generateBoundingBoxes; src/BoundingBox.cpp:41-161 fitMesh, fitFaces,
// outsideFaces, hasFace, hasVertex, splitBox, averageVertexCoord).  The list is part of the
// reference's observable behaviour: it decides which faces a ray may see and the tie order.
#pragma once
#include <cstdint>
#include <vector>

namespace rtp {

struct RefBox {
  float low[3], high[3];
  float shape[3];  // width, height, depth
  bool failed[3] = {false, false, false};
  std::vector<int32_t> faces;  // face ids, in the reference's order
};

// One step of the reference algorithm, exposed so that the BoundingBox mirror class
// (host/BoundingBox.cpp) and the array builder below share a single implementation.
void refbox_fit_faces(RefBox &b, const float *verts, const int32_t *faces, int n_threads = 1);
bool refbox_has_face(const RefBox &b, const float *verts, const int32_t *face3);
float refbox_average_vertex_coord(const RefBox &b, const float *verts, const int32_t *faces, int axis);
// splitBox: returns 0 = all axes failed (nullptr), 1 = this axis failed (returns this),
// 2 = split done, `created` holds the new box.
// n_threads > 1 lets a large box cut its face sweeps into chunks; the result does not change.
int refbox_split(RefBox &b, RefBox &created, const float *verts, const int32_t *faces, int n_threads = 1);

// generateBoundingBoxes over a whole mesh. verts: V x 4, faces: F x 3.  n_threads <= 0: all host
// cores; the list does not depend on the thread count.
void build_reference_boxes(const float *verts, const int32_t *faces, int n_faces, int min_faces,
                           int max_boxes, std::vector<RefBox> &out, int n_threads = 0);

}  // namespace rtp
