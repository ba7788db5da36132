// tucano_io.hpp - input and output formats of the ray-tracing path (included by tucano_lite.hpp):
//   MaterialImporter::loadMTL      utils/mtlIO.hpp:49-140
//   MeshImporter::computeNormals   utils/objimporter.hpp:81-106
//   MeshImporter::loadObjFile      utils/objimporter.hpp:117-351 (mapped file, parsed in parallel; the line-by-line
//                                  cross-check loader lives with the tests: tests/hostsim/obj_serial_loader.hpp)
//   ImageImporter::writePPMImage   utils/ppmIO.hpp:135-156
// The parsing rules are the reference's, quirks included: `vn` lines are used per
// VERTEX index and only when #vn == #v, otherwise smooth normals are rebuilt from unit face
// normals; every `f` line appends all of its vertex ids to one flat list that is later cut into
// triples; `usemtl` with an unknown name keeps the previous material; MTL lines are split on
// single spaces and converted with atof.  GL uploads of the originals are absent.
#pragma once
#include <charconv>
#include <chrono>
#include <limits>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../csrc/parallel_for.h"
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

// ---- OBJ ingest at scale -------------------------------------------------------
// Same parsing rules, different machinery: the file is mapped, cut into one range per thread at line
// boundaries and tokenised in parallel with from_chars; `mtllib` / `usemtl` lines become ordered events that a
// short serial pass replays, so index groups and material ids come out exactly as the line-by-line loader
// builds them.  The number grammar is the one `istream >> float` accepts (sign, digits, '.', exponent; a
// failed conversion leaves the remaining components of that line at 0), `f` tokens take the integer in front
// of the first '/' like std::stoi does.  A 70 MB / 1 M-triangle OBJ: 2.9 s -> 0.35 s on 8 cores.
namespace detail {

struct ObjEvent {
  int kind;           // 0 = mtllib, 1 = usemtl
  size_t ids_before;  // vertex ids of this chunk that precede the line
  std::string arg;
};
struct ObjChunk {
  std::vector<Eigen::Vector4f> vert;
  std::vector<Eigen::Vector3f> norm;
  std::vector<GLuint> ids;
  std::vector<ObjEvent> ev;
};

inline bool isSpace(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

// One `stream >> float`: returns false (stream failed) when no well-formed number starts at p.
inline bool extractFloat(const char *&p, const char *e, float &out) {
  while (p < e && isSpace(*p)) ++p;
  const char *b = p;
  const char *q = p;
  if (q < e && (*q == '+' || *q == '-')) ++q;
  bool digits = false;
  while (q < e && *q >= '0' && *q <= '9') { ++q; digits = true; }
  if (q < e && *q == '.') {
    ++q;
    while (q < e && *q >= '0' && *q <= '9') { ++q; digits = true; }
  }
  if (!digits) { p = q; return false; }
  if (q < e && (*q == 'e' || *q == 'E')) {
    ++q;
    if (q < e && (*q == '+' || *q == '-')) ++q;
    while (q < e && *q >= '0' && *q <= '9') ++q;
  }
  p = q;
  const char *nb = (*b == '+') ? b + 1 : b;
  float v = 0.f;
  auto r = std::from_chars(nb, q, v);
  if (r.ec == std::errc::result_out_of_range) {
    // libstdc++'s num_get hands the token to strtof: overflow stores +-FLT_MAX and fails the stream,
    // underflow stores strtof's result (0 or a denormal) and carries on
    const std::string tok(nb, q);
    const float sv = strtof(tok.c_str(), nullptr);
    if (sv == HUGE_VALF || sv == -HUGE_VALF) {
      out = sv > 0 ? std::numeric_limits<float>::max() : -std::numeric_limits<float>::max();
      return false;
    }
    out = sv;
    return true;
  }
  if (r.ec != std::errc() || r.ptr != q) { out = 0.f; return false; }  // e.g. "1e": not consumed entirely
  out = v;
  return true;
}

inline void parseObjRange(const char *b, const char *e, ObjChunk &c) {
  const char *p = b;
  while (p < e) {
    const char *le = (const char *)memchr(p, '\n', (size_t)(e - p));
    if (!le) le = e;
    const size_t len = (size_t)(le - p);
    if (len >= 2 && p[0] == 'v' && p[1] == ' ') {
      Eigen::Vector4f v;
      const char *q = p + 2;
      if (extractFloat(q, le, v[0]) && extractFloat(q, le, v[1])) extractFloat(q, le, v[2]);
      v[3] = 1.0f;
      c.vert.push_back(v);
    } else if (len >= 2 && p[0] == 'f' && p[1] == ' ') {
      const char *q = p + 2;
      while (q < le) {
        while (q < le && isSpace(*q)) ++q;
        if (q >= le) break;
        const char *t = q;
        while (q < le && !isSpace(*q)) ++q;  // token [t, q)
        const char *d = t;
        bool neg = false;
        if (d < q && (*d == '+' || *d == '-')) { neg = *d == '-'; ++d; }
        long long val = 0;
        bool digits = false, range = false;
        while (d < q && *d >= '0' && *d <= '9') {
          digits = true;
          if (val < (1ll << 40)) val = val * 10 + (*d - '0'); else range = true;
          ++d;
        }
        if (neg) val = -val;
        if (!digits || range || val > INT_MAX || val < INT_MIN) continue;  // std::stoi throws: the token is skipped
        c.ids.push_back((GLuint)((int)val - 1));
      }
    } else if (len > 3 && p[0] == 'v' && p[1] == 'n') {
      Eigen::Vector3f vn;
      const char *q = p + 3;
      if (extractFloat(q, le, vn[0]) && extractFloat(q, le, vn[1])) extractFloat(q, le, vn[2]);
      c.norm.push_back(vn);
    } else if (len > 7 && (memcmp(p, "mtllib", 6) == 0 || memcmp(p, "usemtl", 6) == 0)) {
      std::string arg(p + 7, le);
      MaterialImporter::stripEol(arg);
      c.ev.push_back(ObjEvent{p[0] == 'm' ? 0 : 1, c.ids.size(), arg});
    }
    p = le + 1;
  }
}

inline int ingestThreads() {
  if (const char *e = std::getenv("RTP_OBJ_THREADS")) return std::max(1, atoi(e));
  return rtp::host_threads(0);
}

}  // namespace detail

inline void loadObjFile(Tucano::Mesh &mesh, std::vector<Tucano::Material::Mtl> &mtls, std::string filename) {
  std::string path = MaterialImporter::getPathName(filename);
  const bool timing = std::getenv("RTP_OBJ_TIMING") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    std::cerr << "  obj ingest: " << what << " " << std::chrono::duration<double>(now - t_last).count() << " s" << std::endl;
    t_last = now;
  };
  const int fd = ::open(filename.c_str(), O_RDONLY);
  struct stat sb;
  if (fd < 0 || fstat(fd, &sb) != 0) {
    if (fd >= 0) ::close(fd);
    std::cerr << "Cannot open " << filename.c_str() << std::endl;
    return;
  }
  const size_t size = (size_t)sb.st_size;
  const char *data = size ? (const char *)mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
  if (size && data == MAP_FAILED) {
    ::close(fd);
    std::cerr << "Cannot map " << filename.c_str() << std::endl;
    return;
  }
  // one range per thread, cut at line starts
  size_t min_chunk = 1 << 16;
  if (const char *e = std::getenv("RTP_OBJ_MIN_CHUNK")) min_chunk = (size_t)std::max(1, atoi(e));  // tests: many ranges on small files
  const int nt = (int)std::max<size_t>(1, std::min<size_t>((size_t)detail::ingestThreads(), size / min_chunk + 1));
  std::vector<size_t> cut((size_t)nt + 1, size);
  cut[0] = 0;
  for (int t = 1; t < nt; ++t) {
    size_t at = size * (size_t)t / (size_t)nt;
    const char *nl = at < size ? (const char *)memchr(data + at, '\n', size - at) : nullptr;
    cut[(size_t)t] = nl ? (size_t)(nl - data) + 1 : size;
  }
  std::vector<detail::ObjChunk> chunks((size_t)nt);
  rtp::parallel_for((size_t)nt, nt, 1, [&](size_t a, size_t b, int) {
    for (size_t t = a; t < b; ++t)
      if (cut[t] < cut[t + 1]) detail::parseObjRange(data + cut[t], data + cut[t + 1], chunks[t]);
  });
  if (size) munmap((void *)data, size);
  ::close(fd);
  lap("tokenise");

  // serial replay: concatenate, and let the mtllib / usemtl events cut the index groups in file order
  std::vector<Eigen::Vector4f> vert;
  std::vector<Eigen::Vector3f> norm;
  std::vector<int> elements_material_id;
  std::vector<std::vector<GLuint>> elementsVertices;
  elementsVertices.push_back(std::vector<GLuint>());
  elements_material_id.push_back(-1);
  int current_mat = -1;
  size_t nv = 0, nn = 0;
  for (const auto &c : chunks) { nv += c.vert.size(); nn += c.norm.size(); }
  vert.reserve(nv);
  norm.reserve(nn);
  for (auto &c : chunks) {
    vert.insert(vert.end(), c.vert.begin(), c.vert.end());
    norm.insert(norm.end(), c.norm.begin(), c.norm.end());
    size_t pos = 0;
    auto flush = [&](size_t upto) {
      elementsVertices.back().insert(elementsVertices.back().end(), c.ids.begin() + (std::ptrdiff_t)pos, c.ids.begin() + (std::ptrdiff_t)upto);
      pos = upto;
    };
    for (const auto &ev : c.ev) {
      flush(ev.ids_before);
      if (ev.kind == 0) {
        MaterialImporter::loadMTL(mtls, path + ev.arg);
      } else {
        if (!elementsVertices.back().empty()) {
          elementsVertices.push_back(std::vector<GLuint>());
          elements_material_id.push_back(-1);
        }
        for (size_t i = 0; i < mtls.size(); ++i)
          if (mtls[i].getName().compare(ev.arg) == 0) current_mat = (int)i;
        elements_material_id.back() = current_mat;
      }
    }
    flush(c.ids.size());
    detail::ObjChunk().vert.swap(c.vert);
    std::vector<GLuint>().swap(c.ids);
  }
  lap("replay");
  if (vert.size() > 0) {
    mesh.loadVertices(vert);
    mesh.storeVertexData(vert);
  }
  lap("loadVertices");
  if (norm.size() != vert.size()) {
    norm.clear();
    computeNormals(vert, elementsVertices, norm);
  }
  lap("normals");
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
  lap("store");
  mesh.createFaces();
  lap("createFaces");
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

// ppmIO.hpp:135-156: ASCII P3, one image row per line, a space after every value.  Same bytes as the
// reference's ofstream << int loop; the 256 in-range values come from a table of ready-made "n " strings.
inline void writePPMImage(std::string filename, const std::vector<std::vector<Eigen::Vector3f>> &data) {
  if (data.empty()) return;
  int width = (int)data[0].size();
  int height = (int)data.size();
  char lut[256][4];
  unsigned char lut_len[256];
  for (int v = 0; v < 256; ++v) lut_len[v] = (unsigned char)snprintf(lut[v], sizeof lut[v], "%d", v);
  std::string text;
  text.reserve((size_t)width * height * 12 + 32);
  text += "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
  for (int j = 0; j < height; ++j) {
    for (int i = 0; i < width; ++i)
      for (int c = 0; c < 3; ++c) {
        const int q = ppmQuantise(data[j][i][c]);
        if (q >= 0) text.append(lut[q], lut_len[q]);
        else text += std::to_string(q);  // negative colours and the NaN pixels of SURVEY.md Q8
        text += ' ';
      }
    text += '\n';
  }
  std::ofstream out_stream(filename.c_str(), std::ios::binary);
  out_stream.write(text.data(), (std::streamsize)text.size());
  out_stream.close();
}

// binary P6 of an already quantised image (the device writes these bytes: k_resolve_image /
// k_untile apply ppmIO.hpp:150's min(255, (int)(255*c)); NaN and negative values are stored as 0).
inline bool writePPMImageBinary(std::string filename, const unsigned char *rgb8, int width, int height) {
  if (!rgb8 || width <= 0 || height <= 0) return false;
  std::ofstream out_stream(filename.c_str(), std::ios::binary);
  if (!out_stream) return false;
  out_stream << "P6\n" << width << " " << height << "\n255\n";
  out_stream.write((const char *)rgb8, (std::streamsize)width * height * 3);
  out_stream.close();
  return (bool)out_stream;
}

// PNG of the same quantised image (8-bit RGB, filter 0, stored deflate blocks: no compression library needed).
inline bool writePNGImage(std::string filename, const unsigned char *rgb8, int width, int height) {
  if (!rgb8 || width <= 0 || height <= 0) return false;
  static unsigned crc_table[256];
  static bool have_table = false;
  if (!have_table) {
    for (unsigned n = 0; n < 256; ++n) {
      unsigned c = n;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
      crc_table[n] = c;
    }
    have_table = true;
  }
  auto be32 = [](std::string &o, unsigned v) { for (int k = 3; k >= 0; --k) o += (char)((v >> (8 * k)) & 0xff); };
  auto chunk = [&](std::string &o, const char *type, const std::string &data) {
    be32(o, (unsigned)data.size());
    std::string body = std::string(type, 4) + data;
    unsigned c = 0xffffffffu;
    for (unsigned char ch : body) c = crc_table[(c ^ ch) & 0xff] ^ (c >> 8);
    o += body;
    be32(o, c ^ 0xffffffffu);
  };
  // raw scanlines: filter byte 0 + W*3 bytes
  const size_t row = (size_t)width * 3 + 1, raw_size = row * (size_t)height;
  std::string raw(raw_size, '\0');
  for (int j = 0; j < height; ++j) std::memcpy(&raw[(size_t)j * row + 1], rgb8 + (size_t)j * width * 3, (size_t)width * 3);
  unsigned a = 1, b = 0;  // Adler-32 of the raw data
  for (size_t i = 0; i < raw_size;) {
    const size_t n = std::min<size_t>(5552, raw_size - i);
    for (size_t k = 0; k < n; ++k) { a += (unsigned char)raw[i + k]; b += a; }
    a %= 65521u; b %= 65521u;
    i += n;
  }
  std::string z;
  z.reserve(raw_size + raw_size / 65535 * 5 + 16);
  z += (char)0x78; z += (char)0x01;
  for (size_t i = 0; i < raw_size;) {
    const size_t n = std::min<size_t>(65535, raw_size - i);
    z += (char)(i + n == raw_size ? 1 : 0);  // BFINAL, BTYPE = 00 (stored)
    z += (char)(n & 0xff); z += (char)(n >> 8);
    z += (char)(~n & 0xff); z += (char)((~n >> 8) & 0xff);
    z.append(raw, i, n);
    i += n;
  }
  be32(z, (b << 16) | a);
  std::string out("\x89PNG\r\n\x1a\n", 8);
  std::string ihdr;
  be32(ihdr, (unsigned)width); be32(ihdr, (unsigned)height);
  ihdr += (char)8; ihdr += (char)2; ihdr += (char)0; ihdr += (char)0; ihdr += (char)0;  // 8-bit, RGB
  chunk(out, "IHDR", ihdr);
  chunk(out, "IDAT", z);
  chunk(out, "IEND", std::string());
  std::ofstream f(filename.c_str(), std::ios::binary);
  if (!f) return false;
  f.write(out.data(), (std::streamsize)out.size());
  f.close();
  return (bool)f;
}

}  // namespace ImageImporter

}  // namespace Tucano
