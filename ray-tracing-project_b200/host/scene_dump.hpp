// scene_dump.hpp - writes the Mesh / materials / BoundingBox list a model loads to, as raw arrays (PREFIX.*.f32 / .i32).
// Used by `raytrace --dump-scene` (CPU-only tests compare it with the reference's goldens) and by the tests' own
// cross-check tool around the line-by-line loader (tests/hostsim/obj_dump_serial.cpp).  No CUDA.
#pragma once
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "BoundingBox.h"
#include "tucano_lite.hpp"

namespace scene_dump {

template <class T>
inline void dump(const std::string &path, const std::vector<T> &v) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(3); }
  if (!v.empty()) fwrite(v.data(), sizeof(T), v.size(), f);
  fclose(f);
}

// Loads (through `loader(mesh, materials, path)`) and flattens exactly as Flyscene::initialize does, without touching CUDA.
template <class Loader>
int dump_scene(const std::string &model, const std::string &prefix, int min_faces, bool load_only, Loader loader, const char *loader_name) {
  Tucano::Mesh mesh;
  std::vector<Tucano::Material::Mtl> materials;
  const auto t0 = std::chrono::steady_clock::now();
  loader(mesh, materials, model);
  fprintf(stderr, "obj load (%s): %.3f s\n", loader_name,
          std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
  if (load_only) min_faces = INT_MAX;  // --dump-scene-noboxes: one box, no splitter (timing / large synthetic meshes)
  mesh.normalizeModelMatrix();
  const int V = mesh.getNumberOfVertices(), F = mesh.getNumberOfFaces();
  if (F == 0) return 1;
  std::vector<float> verts, normals, fnormal, mats, shape;
  std::vector<int> faces, fmat;
  for (int i = 0; i < V; ++i) {
    for (int c = 0; c < 4; ++c) verts.push_back(mesh.getVertex(i)[c]);
    for (int c = 0; c < 3; ++c) normals.push_back(mesh.getNormal(i)[c]);
  }
  for (int i = 0; i < F; ++i) {
    const Tucano::Face &f = mesh.getFace(i);
    for (int c = 0; c < 3; ++c) { faces.push_back((int)f.vertex_ids[(size_t)c]); fnormal.push_back(f.normal[c]); }
    fmat.push_back(f.material_id);
  }
  for (const Tucano::Material::Mtl &m : materials) {
    for (int c = 0; c < 3; ++c) mats.push_back(m.getAmbient()[c]);
    for (int c = 0; c < 3; ++c) mats.push_back(m.getDiffuse()[c]);
    for (int c = 0; c < 3; ++c) mats.push_back(m.getSpecular()[c]);
    mats.push_back(m.getShininess()); mats.push_back(m.getOpticalDensity()); mats.push_back(m.getDissolveFactor());
  }
  const Eigen::Affine3f sm = mesh.getShapeModelMatrix();
  shape.assign(sm.data(), sm.data() + 12);
  // the BoundingBox list through the mirror class, as Flyscene::generateBoundingBoxes drives it
  BoundingBox::mesh = &mesh;
  BoundingBox *root = new BoundingBox(true);
  root->fitMesh();
  bool notDone = true;
  while (notDone) {
    notDone = false;
    std::vector<BoundingBox *> current = BoundingBox::boxes;
    for (BoundingBox *b : current)
      if (b->getNumberOfFaces() > min_faces && (!b->failed[0] || !b->failed[1] || !b->failed[2])) {
        BoundingBox *nb = b->splitBox();
        while (b == nb) nb = b->splitBox();
        notDone = true;
      }
  }
  std::vector<float> bounds;
  std::vector<int> box_off, box_faces;
  BoundingBox::exportBoxes(bounds, box_off, box_faces);
  BoundingBox::deconstruct();
  dump(prefix + ".verts.f32", verts); dump(prefix + ".normals.f32", normals); dump(prefix + ".fnormal.f32", fnormal);
  dump(prefix + ".materials.f32", mats); dump(prefix + ".shape.f32", shape); dump(prefix + ".faces.i32", faces);
  dump(prefix + ".fmat.i32", fmat); dump(prefix + ".boxes.f32", bounds); dump(prefix + ".box_off.i32", box_off);
  dump(prefix + ".box_faces.i32", box_faces);
  return 0;
}

}  // namespace scene_dump
