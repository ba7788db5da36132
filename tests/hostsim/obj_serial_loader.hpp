// obj_serial_loader.hpp - TEST INFRASTRUCTURE (not part of the product library or CLI).
// A line-by-line iostream OBJ loader shaped like the reference's (tucano/utils/objimporter.hpp:117-351), kept only
// as the cross-check of the product's mapped, multi-threaded MeshImporter::loadObjFile (host/tucano_io.hpp):
// tests/test_host_mirror.py builds tests/hostsim/obj_dump_serial.cpp around it and compares the Mesh bits.
#pragma once
#include "../../ray-tracing-project_b200/host/tucano_lite.hpp"

namespace Tucano {
namespace MeshImporter {

inline void loadObjFileSerial(Tucano::Mesh &mesh, std::vector<Tucano::Material::Mtl> &mtls, std::string filename) {
  std::string path = MaterialImporter::getPathName(filename);
  std::vector<Eigen::Vector4f> vert;
  std::vector<Eigen::Vector3f> norm;
  std::vector<int> elements_material_id;
  std::vector<std::vector<GLuint>> elementsVertices;
  std::ifstream in(filename.c_str(), std::ios::in);
  if (!in) {
    std::cerr << "Cannot open " << filename.c_str() << std::endl;
    return;
  }
  elementsVertices.push_back(std::vector<GLuint>());
  elements_material_id.push_back(-1);
  std::string line;
  int current_mat = -1;
  while (getline(in, line)) {
    if (line.substr(0, 6) == "mtllib" && line.size() > 7) {
      std::string mtlfn = path + line.substr(7);
      MaterialImporter::stripEol(mtlfn);
      MaterialImporter::loadMTL(mtls, mtlfn);
    } else if (line.substr(0, 6) == "usemtl" && line.size() > 7) {
      if (!elementsVertices.back().empty()) {
        elementsVertices.push_back(std::vector<GLuint>());
        elements_material_id.push_back(-1);
      }
      std::string name = line.substr(7);
      MaterialImporter::stripEol(name);
      for (size_t i = 0; i < mtls.size(); ++i)
        if (mtls[i].getName().compare(name) == 0) current_mat = (int)i;
      elements_material_id.back() = current_mat;
    } else if (line.substr(0, 2) == "v ") {
      std::istringstream s(line.substr(2));
      Eigen::Vector4f v;
      s >> v[0]; s >> v[1]; s >> v[2]; v[3] = 1.0f;
      vert.push_back(v);  // per-vertex colours after xyz are only used by the GL preview
    } else if (line.substr(0, 2) == "vn" && line.size() > 3) {
      std::istringstream s(line.substr(3));
      Eigen::Vector3f vn;
      s >> vn[0]; s >> vn[1]; s >> vn[2];
      norm.push_back(vn);
    } else if (line.substr(0, 2) == "f ") {
      std::stringstream liness(line.substr(2));
      std::string buf, element;
      while (liness >> buf) {  // "v", "v/t", "v//n" or "v/t/n": only the vertex id feeds the ray tracer
        std::stringstream elementss(buf);
        std::getline(elementss, element, '/');
        int vertexId = 0;
        try { vertexId = std::stoi(element); } catch (...) { continue; }
        elementsVertices.back().push_back((GLuint)(vertexId - 1));
      }
    }
  }
  in.close();
  if (vert.size() > 0) {
    mesh.loadVertices(vert);
    mesh.storeVertexData(vert);
  }
  if (norm.size() != vert.size()) {
    norm.clear();
    computeNormals(vert, elementsVertices, norm);
  }
  if (norm.size() == vert.size()) {
    mesh.loadNormals(norm);
    mesh.storeNormalData(norm);
  }
  for (size_t i = 0; i < elementsVertices.size(); ++i) {
    if (elementsVertices[i].size() > 0) {
      mesh.loadIndices(elementsVertices[i], elements_material_id[i]);
      mesh.storeVertexIdsData(elementsVertices[i]);
    }
  }
  mesh.createFaces();
  std::cout << "OBJ info:" << std::endl;
  std::cout << "number vertices : " << mesh.getNumberOfVertices() << std::endl;
  std::cout << "number faces : " << mesh.getNumberOfFaces() << std::endl;
  std::cout << "number materials : " << mesh.getNumberOfMaterials() << std::endl;
}

}  // namespace MeshImporter
}  // namespace Tucano
