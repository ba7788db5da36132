// tucano_lite.hpp - the CPU-side slice of Tucano that Flyscene::raytraceScene touches, with no
// OpenGL in it: Mesh / Face / Model data accessors (a16), Material::Mtl + loadMTL (a17), Camera /
// Flycamera ray-generation math (a3), the OBJ importer (a19) and the ASCII PPM writer (a18).
// Same class and member names as the reference's vendored headers (cited per function, paths
// relative to /root/reference/dependencies/tucano/tucano) so that code written against
// src/flyscene.hpp compiles against this header; every GL buffer upload the originals also do
// (createAttribute, loadIndices, texture loads) is simply absent - no graphics stack on this box.
// The arithmetic follows the originals expression by expression: the flattened scene handed to
// the CUDA library must carry the bits the reference computes.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "../csrc/parallel_for.h"
#include "eigen_lite.hpp"

typedef unsigned int GLuint;

namespace Tucano {

// mesh.hpp:238-248
class Face {
 public:
  std::vector<GLuint> vertex_ids;
  std::vector<Eigen::Vector2f> texcoord;
  int material_id = -1;
  Eigen::Vector3f normal = Eigen::Vector3f::Zero();
  Face() {}
};

// model.hpp:16-200 (the matrices and the normalisation only)
class Model {
 protected:
  Eigen::Affine3f model_matrix = Eigen::Affine3f::Identity();
  Eigen::Affine3f shape_matrix = Eigen::Affine3f::Identity();
  Eigen::Vector3f objectCenter = Eigen::Vector3f::Zero();
  Eigen::Vector3f centroid = Eigen::Vector3f::Zero();
  float radius = 1.f;
  float normalization_scale = 1.f;

 public:
  virtual ~Model() {}
  Eigen::Vector3f getObjectCenter() const { return objectCenter; }
  Eigen::Vector3f getCentroid() const { return centroid; }
  float getBoundingSphereRadius() const { return radius; }
  float getNormalizationScale() const { return normalization_scale; }
  Eigen::Affine3f getModelMatrix() const { return model_matrix; }
  Eigen::Affine3f *modelMatrix() { return &model_matrix; }
  Eigen::Affine3f getShapeMatrix() const { return shape_matrix; }
  // model.hpp:102-105
  virtual Eigen::Affine3f getShapeModelMatrix() const { return model_matrix * shape_matrix; }
  void resetModelMatrix() { model_matrix = Eigen::Affine3f::Identity(); }
  void resetShapeMatrix() { shape_matrix = Eigen::Affine3f::Identity(); }
  // model.hpp:169-173
  void normalizeModelMatrix() {
    shape_matrix.scale(normalization_scale);
    shape_matrix.translate(-objectCenter);
  }
};

// mesh.hpp:257-... : vertex / normal / face storage and the derived model information
class Mesh : public Model {
 protected:
  unsigned int numberOfVertices = 0;
  std::vector<Eigen::Vector4f> vertices;
  std::vector<Eigen::Vector3f> normals;
  std::vector<std::vector<GLuint>> indices_vertices;
  std::vector<int> material_ids;
  std::vector<Face> faces;

 public:
  Mesh() {}
  void reset() {
    numberOfVertices = 0;
    vertices.clear(); normals.clear(); indices_vertices.clear(); material_ids.clear(); faces.clear();
    resetModelMatrix(); resetShapeMatrix();
  }
  int getNumberOfVertices() const { return (int)numberOfVertices; }
  int getNumberOfFaces() const { return (int)faces.size(); }          // mesh.hpp:431
  int getNumberOfMaterials() const { return (int)material_ids.size(); }
  Eigen::Vector4f getVertex(int i) const { return vertices[(size_t)i]; }  // mesh.hpp:387
  Eigen::Vector3f getNormal(int i) const { return normals[(size_t)i]; }   // mesh.hpp:395
  Face &getFace(int i) { return faces[(size_t)i]; }                      // mesh.hpp:426
  const Face &getFace(int i) const { return faces[(size_t)i]; }
  std::vector<Face> &getFaces() { return faces; }

  void storeVertexData(std::vector<Eigen::Vector4f> &verts) { vertices = verts; }
  void storeNormalData(std::vector<Eigen::Vector3f> &norms) { normals = norms; }
  void storeVertexIdsData(std::vector<GLuint> &ids) { indices_vertices.push_back(ids); }
  void storeMaterialIdsData(std::vector<int> &ids) { material_ids = ids; }
  // mesh.hpp:720-748 minus the GL element buffer: remembers the material of this index group
  void loadIndices(std::vector<GLuint> &, int material_id) { material_ids.push_back(material_id); }
  void loadNormals(std::vector<Eigen::Vector3f> &) {}

  // mesh.hpp:592-644: bounding box centre, normalisation scale, centroid, bounding-sphere radius
  void loadVertices(std::vector<Eigen::Vector4f> &vert) {
    if (vert.empty()) return;
    numberOfVertices = (unsigned int)vert.size();
    float xMax = vert[0][0], xMin = vert[0][0];
    float yMax = vert[0][1], yMin = vert[0][1];
    float zMax = vert[0][2], zMin = vert[0][2];
    for (unsigned int i = 0; i < numberOfVertices; ++i) {
      xMax = std::max(vert[i][0], xMax); yMax = std::max(vert[i][1], yMax); zMax = std::max(vert[i][2], zMax);
      xMin = std::min(vert[i][0], xMin); yMin = std::min(vert[i][1], yMin); zMin = std::min(vert[i][2], zMin);
    }
    // `fabs` on floats is the double overload in the original; the ratio is a double rounded to float
    normalization_scale =
        (float)(1.0 / (std::max(std::max(std::fabs((double)(xMax - xMin)), std::fabs((double)(yMax - yMin))),
                                std::fabs((double)(zMax - zMin)))));
    float centerX = (float)((xMax + xMin) / 2.0);
    float centerY = (float)((yMax + yMin) / 2.0);
    float centerZ = (float)((zMax + zMin) / 2.0);
    objectCenter = Eigen::Vector3f(centerX, centerY, centerZ);
    centroid = Eigen::Vector3f::Zero();
    for (unsigned int i = 0; i < numberOfVertices; i++) centroid = centroid + vert[i].head3();
    centroid = centroid / (float)numberOfVertices;
    radius = 0.f;
    for (unsigned int i = 0; i < numberOfVertices; i++) radius = std::max(radius, (vert[i].head3() - centroid).norm());
  }

  // mesh.hpp:448-482: one Face per index triple; normal from the NORMALISED object-space edges.
  // Faces are independent of each other, so the index groups are filled by all host threads (same bits).
  void createFaces() {
    if (vertices.empty() || indices_vertices.empty()) return;
    std::vector<size_t> first(indices_vertices.size() + 1, faces.size());
    for (size_t id = 0; id < indices_vertices.size(); ++id)
      first[id + 1] = first[id] + indices_vertices[id].size() / 3;  // a trailing partial triple is dropped (UB in the original)
    faces.resize(first.back());
    for (size_t id = 0; id < indices_vertices.size(); ++id) {
      const std::vector<GLuint> &ix = indices_vertices[id];
      const int mat = material_ids.empty() ? -1 : material_ids[id];
      const bool has_mat = !material_ids.empty();
      Face *out = faces.data() + first[id];
      rtp::parallel_for(ix.size() / 3, 0, 8192, [&](size_t b, size_t e, int) {
        for (size_t k = b; k < e; ++k) {
          const size_t i = 3 * k;
          Face &f = out[k];
          f.vertex_ids.reserve(3);
          f.vertex_ids.push_back(ix[i]);
          f.vertex_ids.push_back(ix[i + 1]);
          f.vertex_ids.push_back(ix[i + 2]);
          if (has_mat) f.material_id = mat;
          const rtp::v3 a = vertices[ix[i]].head3().r(), b3 = vertices[ix[i + 1]].head3().r(), c = vertices[ix[i + 2]].head3().r();
          // .head(3) differences are dynamic-size expressions: sequential squaredNorm reduction
          const rtp::v3 v1 = rtp::vnormalized_dyn(rtp::vsub(c, a));
          const rtp::v3 v0 = rtp::vnormalized_dyn(rtp::vsub(b3, a));
          f.normal = Eigen::Vector3f(rtp::vnormalized(rtp::vcross(v0, v1)));
        }
      });
    }
  }
};

namespace Material {
// materials/mtl.hpp:17-128 (texture object left out: GL)
class Mtl {
 protected:
  Eigen::Vector3f ka = Eigen::Vector3f(0.3, 0.3, 0.3);
  Eigen::Vector3f kd = Eigen::Vector3f(0.5, 0.5, 0.5);
  Eigen::Vector3f ks = Eigen::Vector3f(1.0, 1.0, 1.0);
  float shininess = 10;
  float optical_density = 0.0;
  float dissolve_factor = 1.0;
  int illumination_model = 0;
  std::string diffuse_tex_filename;
  std::string name;

 public:
  Mtl() {}
  void setAmbient(const Eigen::Vector3f &v) { ka = v; }
  void setDiffuse(const Eigen::Vector3f &v) { kd = v; }
  void setSpecular(const Eigen::Vector3f &v) { ks = v; }
  void setShininess(float v) { shininess = v; }
  void setOpticalDensity(float v) { optical_density = v; }
  void setDissolveFactor(float v) { dissolve_factor = v; }
  void setIlluminationModel(int v) { illumination_model = v; }
  void setDiffuseTextureFilename(const std::string &s) { diffuse_tex_filename = s; }
  void setName(const std::string &s) { name = s; }
  Eigen::Vector3f getAmbient() const { return ka; }
  Eigen::Vector3f getDiffuse() const { return kd; }
  Eigen::Vector3f getSpecular() const { return ks; }
  float getShininess() const { return shininess; }
  float getOpticalDensity() const { return optical_density; }
  float getDissolveFactor() const { return dissolve_factor; }
  int getIlluminationModel() const { return illumination_model; }
  std::string getDiffuseTextureFilename() const { return diffuse_tex_filename; }
  std::string getName() const { return name; }
};
}  // namespace Material

}  // namespace Tucano

#include "tucano_camera.hpp"
#include "tucano_io.hpp"
