// tucano_io.hpp - input and output formats of the ray-tracing path (included by tucano_lite.hpp):
//   MaterialImporter::loadMTL      utils/mtlIO.hpp:49-140
//   MeshImporter::computeNormals   utils/objimporter.hpp:81-106
//   MeshImporter::loadObjFile      utils/objimporter.hpp:117-351
//   ImageImporter::writePPMImage   utils/ppmIO.hpp:135-156
// The parsing rules are the reference's, quirks included: `vn` lines are used per
// VERTEX index and only when #vn == #v, otherwise smooth normals are rebuilt from unit face
// normals; every `f` line appends all of its vertex ids to one flat list that is later cut into
// triples; `usemtl` with an unknown name keeps the previous material; MTL lines are split on
// single spaces and converted with atof.  GL uploads of the originals are absent.
#pragma once
#include "tucano_lite.hpp"

namespace Tucano {

namespace MaterialImporter {

inline std::string getPathName(const std::string &s) {
  size_t found = s.find_last_of("/\\");
  return s.substr(0, found + 1);
}

inline void stripEol(std::string &s) {
  s.erase(std::remove(s.begin(), s.end(), '\n'), s.end());
  s.erase(std::remove(s.begin(), s.end(), '\r'), s.end());
}

inline bool loadMTL(std::vector<Tucano::Material::Mtl> &materials, std::string filename) {
  std::string path = getPathName(filename);
  std::ifstream in(filename.c_str(), std::ios::in);
  if (!in) {
    std::cerr << "Cannot open " << filename.c_str() << std::endl;
    return false;
  }
  std::cout << "Reading:" << filename.c_str() << std::endl;
  for (std::string line; getline(in, line);) {
    if (line.empty()) continue;
    std::stringstream ss(line);
    std::string s;
    std::vector<std::string> tokens;
    while (std::getline(ss, s, ' ')) tokens.push_back(s);
    if (tokens.empty()) continue;
    const std::string &key = tokens[0];
    auto num = [&](size_t i) { return i < tokens.size() ? atof(tokens[i].c_str()) : 0.0; };
    if (key == "#") continue;
    if (key == "newmtl") {
      materials.push_back(Tucano::Material::Mtl());
      std::string name = tokens.size() > 1 ? tokens[1] : std::string();
      stripEol(name);
      materials.back().setName(name);
      continue;
    }
    if (materials.empty()) continue;  // attribute before any newmtl: back() on an empty vector in the original
    if (key == "Ns") materials.back().setShininess((float)num(1));
    else if (key == "Ka") materials.back().setAmbient(Eigen::Vector3f(num(1), num(2), num(3)));
    else if (key == "Kd") materials.back().setDiffuse(Eigen::Vector3f(num(1), num(2), num(3)));
    else if (key == "Ks") materials.back().setSpecular(Eigen::Vector3f(num(1), num(2), num(3)));
    else if (key == "Ni") materials.back().setOpticalDensity((float)num(1));
    else if (key == "d") materials.back().setDissolveFactor((float)num(1));
    else if (key == "illum") materials.back().setIlluminationModel(tokens.size() > 1 ? atoi(tokens[1].c_str()) : 0);
    else if (key == "map_kd" && tokens.size() > 1) materials.back().setDiffuseTextureFilename(path + tokens[1]);
  }
  if (materials.empty()) materials.push_back(Tucano::Material::Mtl());  // mtlIO.hpp:127-131
  return true;
}

}  // namespace MaterialImporter

namespace MeshImporter {

// objimporter.hpp:81-106: normalised sum of unit face normals (edges normalised first)
inline void computeNormals(const std::vector<Eigen::Vector4f> &vertices, const std::vector<std::vector<GLuint>> &indices,
                           std::vector<Eigen::Vector3f> &normals) {
  for (size_t i = 0; i < vertices.size(); ++i) normals.push_back(Eigen::Vector3f::Zero());
  for (size_t id = 0; id < indices.size(); ++id) {
    const std::vector<GLuint> &ix = indices[id];
    for (size_t i = 0; i + 2 < ix.size(); i = i + 3) {
      if (ix[i] >= vertices.size() || ix[i + 1] >= vertices.size() || ix[i + 2] >= vertices.size()) continue;
      const rtp::v3 a = vertices[ix[i]].head3().r(), b = vertices[ix[i + 1]].head3().r(), c = vertices[ix[i + 2]].head3().r();
      const rtp::v3 v0 = rtp::vnormalized_dyn(rtp::vsub(b, a));
      const rtp::v3 v1 = rtp::vnormalized_dyn(rtp::vsub(c, a));
      const Eigen::Vector3f n(rtp::vnormalized(rtp::vcross(v0, v1)));
      normals[ix[i]] += n;
      normals[ix[i + 1]] += n;
      normals[ix[i + 2]] += n;
    }
  }
  for (size_t i = 0; i < normals.size(); ++i) normals[i].normalize();
}

inline void loadObjFile(Tucano::Mesh &mesh, std::vector<Tucano::Material::Mtl> &mtls, std::string filename) {
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

namespace ImageImporter {

// The reference's `(int)(255*c)`: x86 cvttss2si semantics (NaN / out of range -> INT_MIN),
// which is what the shipped binary prints for the NaN pixels of SURVEY.md Q8.
inline int ppmQuantise(float c) {
  const float s = 255 * c;
  if (!(s > -2147483648.0f && s < 2147483648.0f)) return INT_MIN;
  return std::min(255, (int)s);
}

// ppmIO.hpp:135-156: ASCII P3, one image row per line, a space after every value
inline void writePPMImage(std::string filename, const std::vector<std::vector<Eigen::Vector3f>> &data) {
  if (data.empty()) return;
  int width = (int)data[0].size();
  int height = (int)data.size();
  std::string text;
  text.reserve((size_t)width * height * 12 + 32);
  text += "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
  for (int j = 0; j < height; ++j) {
    for (int i = 0; i < width; ++i)
      for (int c = 0; c < 3; ++c) {
        text += std::to_string(ppmQuantise(data[j][i][c]));
        text += ' ';
      }
    text += '\n';
  }
  std::ofstream out_stream(filename.c_str(), std::ios::binary);
  out_stream.write(text.data(), (std::streamsize)text.size());
  out_stream.close();
}

}  // namespace ImageImporter

}  // namespace Tucano
